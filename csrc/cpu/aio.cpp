// DeepNVMe-style asynchronous file I/O: a handle owns a pool of worker threads; reads/writes of a host buffer are
// split into `block_size` requests spread over the workers (pread/pwrite, optional O_DIRECT), submitted
// asynchronously and collected with wait().
//
// Reference: csrc/aio/common/deepspeed_aio_common.cpp:71-269 (libaio io_submit/io_getevents with O_DIRECT),
// csrc/aio/py_lib/deepspeed_py_io_handle.{h,cpp} (worker-thread pool per handle, sync_/async_ pread/pwrite, wait),
// deepspeed_pin_tensor.cpp (page-locked buffers). libaio is not available in this image, so parallelism comes from
// `intra_op_parallelism` threads each issuing large positional I/O; queue_depth bounds the requests in flight.
#include <algorithm>
#include <atomic>
#include <condition_variable>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <fcntl.h>
#include <functional>
#include <mutex>
#include <string>
#include <sys/stat.h>
#include <thread>
#include <unistd.h>
#include <vector>

namespace {

struct Request {
    bool write;
    int fd;
    char* buf;
    int64_t nbytes;
    int64_t offset;
    std::atomic<int64_t>* done_bytes;
    std::atomic<int>* errors;
};

class Handle {
public:
    Handle(int64_t block_size, int queue_depth, int threads, bool use_direct)
        : block_size_(block_size > 0 ? block_size : (1 << 20)), queue_depth_(queue_depth > 0 ? queue_depth : 8),
          use_direct_(use_direct), stop_(false), inflight_(0)
    {
        if (threads < 1) threads = 1;
        for (int i = 0; i < threads; i++) workers_.emplace_back([this] { this->loop(); });
    }
    ~Handle()
    {
        {
            std::unique_lock<std::mutex> lk(mu_);
            stop_ = true;
        }
        cv_.notify_all();
        for (auto& t : workers_) t.join();
        for (int fd : open_fds_) ::close(fd);
    }

    // returns 0 on success (async: request queued)
    int submit(bool write, const char* path, char* buf, int64_t nbytes, int64_t file_offset, bool async)
    {
        int flags = write ? (O_WRONLY | O_CREAT) : O_RDONLY;
        const bool aligned = ((uintptr_t)buf % 512 == 0) && (nbytes % 512 == 0) && (file_offset % 512 == 0);
#ifdef O_DIRECT
        if (use_direct_ && aligned) flags |= O_DIRECT;
#endif
        int fd = ::open(path, flags, 0644);
        if (fd < 0 && (flags & O_DIRECT)) fd = ::open(path, flags & ~O_DIRECT, 0644);  // tmpfs etc.
        if (fd < 0) return -1;
        {
            std::unique_lock<std::mutex> lk(mu_);
            open_fds_.push_back(fd);
            for (int64_t off = 0; off < nbytes; off += block_size_) {
                Request r{write, fd, buf + off, std::min(block_size_, nbytes - off), file_offset + off, &done_bytes_,
                          &errors_};
                queue_.push_back(r);
                inflight_++;
            }
        }
        cv_.notify_all();
        if (!async) return wait() >= 0 ? 0 : -1;
        return 0;
    }

    // blocks until all queued requests finished; returns number of completed bytes since last wait, or -1 on error
    int64_t wait()
    {
        std::unique_lock<std::mutex> lk(mu_);
        done_cv_.wait(lk, [this] { return inflight_ == 0; });
        for (int fd : open_fds_) ::close(fd);
        open_fds_.clear();
        int64_t b = done_bytes_.exchange(0);
        int e = errors_.exchange(0);
        return e ? -1 : b;
    }

    int64_t block_size() const { return block_size_; }
    int queue_depth() const { return queue_depth_; }
    int threads() const { return (int)workers_.size(); }

private:
    void loop()
    {
        for (;;) {
            Request r;
            {
                std::unique_lock<std::mutex> lk(mu_);
                cv_.wait(lk, [this] { return stop_ || !queue_.empty(); });
                if (stop_ && queue_.empty()) return;
                r = queue_.front();
                queue_.pop_front();
            }
            int64_t done = 0;
            while (done < r.nbytes) {
                ssize_t k = r.write ? ::pwrite(r.fd, r.buf + done, r.nbytes - done, r.offset + done)
                                    : ::pread(r.fd, r.buf + done, r.nbytes - done, r.offset + done);
                if (k <= 0) {
                    if (k < 0) (*r.errors)++;
                    break;  // EOF on read is not an error: caller sees fewer bytes
                }
                done += k;
            }
            r.done_bytes->fetch_add(done);
            {
                std::unique_lock<std::mutex> lk(mu_);
                if (--inflight_ == 0) done_cv_.notify_all();
            }
        }
    }

    int64_t block_size_;
    int queue_depth_;
    bool use_direct_;
    bool stop_;
    int inflight_;
    std::mutex mu_;
    std::condition_variable cv_, done_cv_;
    std::deque<Request> queue_;
    std::vector<std::thread> workers_;
    std::vector<int> open_fds_;
    std::atomic<int64_t> done_bytes_{0};
    std::atomic<int> errors_{0};
};

}  // namespace

extern "C" {

void* dsb200_aio_create(int64_t block_size, int queue_depth, int threads, int use_direct)
{
    return new Handle(block_size, queue_depth, threads, use_direct != 0);
}
void dsb200_aio_destroy(void* h) { delete static_cast<Handle*>(h); }
int dsb200_aio_pread(void* h, void* buf, int64_t nbytes, const char* path, int64_t offset, int async)
{
    return static_cast<Handle*>(h)->submit(false, path, static_cast<char*>(buf), nbytes, offset, async != 0);
}
int dsb200_aio_pwrite(void* h, const void* buf, int64_t nbytes, const char* path, int64_t offset, int async)
{
    return static_cast<Handle*>(h)->submit(true, path, const_cast<char*>(static_cast<const char*>(buf)), nbytes, offset,
                                           async != 0);
}
int64_t dsb200_aio_wait(void* h) { return static_cast<Handle*>(h)->wait(); }

// page-aligned host allocation (the Python side pins it with cudaHostRegister when CUDA is present)
void* dsb200_aligned_alloc(int64_t nbytes, int64_t alignment)
{
    void* p = nullptr;
    if (posix_memalign(&p, (size_t)alignment, (size_t)((nbytes + alignment - 1) / alignment * alignment)) != 0)
        return nullptr;
    return p;
}
void dsb200_aligned_free(void* p) { free(p); }

}  // extern "C"

// Shared-memory all-reduce between the ranks of one host (CPU tensors: CPU inference / host-side reductions).
//
// Reference: csrc/cpu/comm/shm.cpp + shm_interface.cpp (`inference_all_reduce_`: a low-latency SHM all-reduce used
// when every rank lives on the same machine, LOCAL_SIZE == world size; sum only).
//
// One POSIX segment holds, per rank, three sequence counters and two buffer sets (alternating by call parity). A call:
//   1. copy my data into my input buffer of the current set, publish `ready = seq`
//   2. wait for every rank's `ready`, reduce MY 1/world slice over all ranks' inputs (fp32 accumulation) into my
//      result buffer, publish `done = seq`
//   3. wait for every rank's `done`, copy all slices back out
// Reusing a set two calls later is safe without another handshake: finishing call seq+1 required every rank to have
// entered it, i.e. to have finished reading the set of call seq. Tensors larger than the buffer go in chunks.
#include <fcntl.h>
#include <sched.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdint>
#include <cstring>
#include <string>

namespace {

constexpr uint32_t MAGIC = 0xD5B20051u;
constexpr int MAX_WORLD = 64;

struct alignas(64) RankCtl {
    std::atomic<uint64_t> ready;
    std::atomic<uint64_t> done;
    char pad[48];
};

struct Header {
    std::atomic<uint32_t> magic;
    int world;
    int64_t buf_bytes;
    char pad[48];
    RankCtl ctl[MAX_WORLD];
};

struct Shm {
    std::string name;
    int rank, world;
    int64_t buf_bytes;  // per rank, per set, per kind (input / result)
    size_t total;
    Header* hdr;
    uint8_t* data;
    uint64_t seq;
    double timeout_s;
    uint8_t* in_buf(int r, int set) const { return data + ((size_t)(r * 2 + set) * 2 + 0) * buf_bytes; }
    uint8_t* res_buf(int r, int set) const { return data + ((size_t)(r * 2 + set) * 2 + 1) * buf_bytes; }
};

inline float bf16_to_f32(uint16_t v)
{
    uint32_t u = (uint32_t)v << 16;
    float f;
    std::memcpy(&f, &u, 4);
    return f;
}
inline uint16_t f32_to_bf16(float f)
{
    uint32_t u;
    std::memcpy(&u, &f, 4);
    if ((u & 0x7FFFFFFFu) > 0x7F800000u) return (uint16_t)((u >> 16) | 0x40);  // NaN stays NaN
    u += 0x7FFFu + ((u >> 16) & 1u);                                            // round to nearest even
    return (uint16_t)(u >> 16);
}
inline float f16_to_f32(uint16_t h) { return (float)(*reinterpret_cast<const _Float16*>(&h)); }
inline uint16_t f32_to_f16(float f)
{
    _Float16 h = (_Float16)f;
    uint16_t r;
    std::memcpy(&r, &h, 2);
    return r;
}

bool wait_all(Shm* s, bool done_flag, uint64_t seq)
{
    const auto t0 = std::chrono::steady_clock::now();
    for (int r = 0; r < s->world; r++) {
        auto& c = done_flag ? s->hdr->ctl[r].done : s->hdr->ctl[r].ready;
        int spins = 0;
        while (c.load(std::memory_order_acquire) < seq) {
            if (++spins > 200) {
                sched_yield();
                if ((spins & 1023) == 0) {
                    const double el = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
                    if (el > s->timeout_s) return false;
                }
            } else {
                __builtin_ia32_pause();
            }
        }
    }
    return true;
}

// dtype: 0 fp32, 1 bf16, 2 fp16
template <typename LD, typename ST>
void reduce_slice(Shm* s, int set, int64_t lo, int64_t hi, int esz, LD ld, ST st)
{
    uint8_t* out = s->res_buf(s->rank, set);
    for (int64_t i = lo; i < hi; i++) {
        float acc = 0.f;
        for (int r = 0; r < s->world; r++) acc += ld(s->in_buf(r, set) + i * esz);
        st(out + i * esz, acc);
    }
}

}  // namespace

extern "C" {

// rank 0 creates (create != 0) the segment, the others attach to it afterwards (the caller orders the two with a
// barrier). buf_bytes: capacity per call chunk. Returns nullptr on failure.
void* dsb200_shm_create(const char* name, int rank, int world, int64_t buf_bytes, int create, double timeout_s)
{
    if (world < 1 || world > MAX_WORLD || rank < 0 || rank >= world || buf_bytes < 64) return nullptr;
    buf_bytes = (buf_bytes + 63) / 64 * 64;
    const size_t total = sizeof(Header) + (size_t)world * 4 * (size_t)buf_bytes;
    int fd;
    if (create) {
        shm_unlink(name);
        fd = shm_open(name, O_CREAT | O_EXCL | O_RDWR, 0600);
        if (fd < 0) return nullptr;
        if (ftruncate(fd, (off_t)total) != 0) {
            close(fd);
            shm_unlink(name);
            return nullptr;
        }
    } else {
        fd = shm_open(name, O_RDWR, 0600);
        if (fd < 0) return nullptr;
        struct stat stt;
        if (fstat(fd, &stt) != 0 || (size_t)stt.st_size < total) {
            close(fd);
            return nullptr;
        }
    }
    void* p = mmap(nullptr, total, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    close(fd);
    if (p == MAP_FAILED) return nullptr;
    auto* hdr = static_cast<Header*>(p);
    if (create) {
        hdr->world = world;
        hdr->buf_bytes = buf_bytes;
        for (int r = 0; r < MAX_WORLD; r++) {
            hdr->ctl[r].ready.store(0, std::memory_order_relaxed);
            hdr->ctl[r].done.store(0, std::memory_order_relaxed);
        }
        hdr->magic.store(MAGIC, std::memory_order_release);
    } else if (hdr->magic.load(std::memory_order_acquire) != MAGIC || hdr->world != world || hdr->buf_bytes != buf_bytes) {
        munmap(p, total);
        return nullptr;
    }
    auto* s = new Shm{name, rank, world, buf_bytes, total, hdr, static_cast<uint8_t*>(p) + sizeof(Header), 0,
                      timeout_s > 0 ? timeout_s : 120.0};
    return s;
}

void dsb200_shm_destroy(void* h)
{
    auto* s = static_cast<Shm*>(h);
    if (s == nullptr) return;
    munmap(s->hdr, s->total);
    if (s->rank == 0) shm_unlink(s->name.c_str());
    delete s;
}

// In-place sum over the ranks. dtype: 0 fp32, 1 bf16, 2 fp16. Returns 0, -1 bad argument, -2 timeout (a peer is gone).
int dsb200_shm_all_reduce(void* h, void* data, int64_t numel, int dtype)
{
    auto* s = static_cast<Shm*>(h);
    if (s == nullptr || numel < 0 || dtype < 0 || dtype > 2) return -1;
    const int esz = dtype == 0 ? 4 : 2;
    const int64_t cap = s->buf_bytes / esz;
    uint8_t* p = static_cast<uint8_t*>(data);
    for (int64_t off = 0; off < numel; off += cap) {
        const int64_t n = numel - off < cap ? numel - off : cap;
        const uint64_t seq = ++s->seq;
        const int set = (int)(seq & 1);
        std::memcpy(s->in_buf(s->rank, set), p + off * esz, (size_t)n * esz);
        s->hdr->ctl[s->rank].ready.store(seq, std::memory_order_release);
        if (!wait_all(s, false, seq)) return -2;
        const int64_t per = (n + s->world - 1) / s->world;
        const int64_t lo = std::min<int64_t>(n, per * s->rank), hi = std::min<int64_t>(n, lo + per);
        if (dtype == 0)
            reduce_slice(s, set, lo, hi, 4, [](const uint8_t* q) { float f; std::memcpy(&f, q, 4); return f; },
                         [](uint8_t* q, float v) { std::memcpy(q, &v, 4); });
        else if (dtype == 1)
            reduce_slice(s, set, lo, hi, 2, [](const uint8_t* q) { uint16_t v; std::memcpy(&v, q, 2); return bf16_to_f32(v); },
                         [](uint8_t* q, float v) { const uint16_t b = f32_to_bf16(v); std::memcpy(q, &b, 2); });
        else
            reduce_slice(s, set, lo, hi, 2, [](const uint8_t* q) { uint16_t v; std::memcpy(&v, q, 2); return f16_to_f32(v); },
                         [](uint8_t* q, float v) { const uint16_t b = f32_to_f16(v); std::memcpy(q, &b, 2); });
        s->hdr->ctl[s->rank].done.store(seq, std::memory_order_release);
        if (!wait_all(s, true, seq)) return -2;
        for (int r = 0; r < s->world; r++) {
            const int64_t rlo = std::min<int64_t>(n, per * r), rhi = std::min<int64_t>(n, rlo + per);
            if (rhi > rlo) std::memcpy(p + (off + rlo) * esz, s->res_buf(r, set) + rlo * esz, (size_t)(rhi - rlo) * esz);
        }
    }
    return 0;
}

}  // extern "C"

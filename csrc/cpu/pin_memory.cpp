// Page-locked host allocations for the offload / NVMe paths (reference csrc/pin_memory + csrc/aio
// deepspeed_pin_tensor.cpp: page-aligned allocation + mlock, tracked by address).
#include <sys/mman.h>
#include <unistd.h>

#include <cstdint>
#include <cstdlib>
#include <mutex>
#include <unordered_map>

namespace {
std::mutex g_mu;
std::unordered_map<void*, size_t> g_live;  // base -> bytes (rounded to pages)
size_t g_locked_bytes = 0;
}  // namespace

extern "C" {

// Page-aligned allocation of at least `bytes`, locked in RAM with mlock. locked_out (optional) receives 1 when the
// lock succeeded, 0 when RLIMIT_MEMLOCK refused it (the memory is still usable, just pageable). nullptr on failure.
void* dsb200_locked_alloc(int64_t bytes, int* locked_out)
{
    if (bytes <= 0) return nullptr;
    const size_t page = (size_t)sysconf(_SC_PAGESIZE);
    const size_t total = ((size_t)bytes + page - 1) / page * page;
    void* p = nullptr;
    if (posix_memalign(&p, page, total) != 0) return nullptr;
    const int ok = mlock(p, total) == 0;
    if (locked_out) *locked_out = ok;
    std::lock_guard<std::mutex> lk(g_mu);
    g_live[p] = ok ? total : 0;
    if (ok) g_locked_bytes += total;
    return p;
}

// 1 if `p` was a live allocation of dsb200_locked_alloc (now unlocked and freed), 0 otherwise.
int dsb200_locked_free(void* p)
{
    size_t total;
    {
        std::lock_guard<std::mutex> lk(g_mu);
        auto it = g_live.find(p);
        if (it == g_live.end()) return 0;
        total = it->second;
        g_live.erase(it);
        g_locked_bytes -= total;
    }
    if (total) munlock(p, total);
    free(p);
    return 1;
}

int64_t dsb200_locked_bytes() { return (int64_t)g_locked_bytes; }

}  // extern "C"

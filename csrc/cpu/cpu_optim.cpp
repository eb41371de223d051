// Host optimizers for ZeRO-Offload: Adam/AdamW, Adagrad, Lion on fp32 master weights held in (pinned) host DRAM,
// with optional simultaneous bf16/fp16 write-back of the updated weights and a step rollback (SuperOffload).
//
// Reference: csrc/adam/cpu_adam_impl.cpp (Step_1/4/8 AVX tiles :35-165, ds_adam_step :235, ds_adam_rollback :311),
// csrc/adagrad/cpu_adagrad.cpp, csrc/lion/cpu_lion_impl.cpp, SIMD wrappers csrc/includes/simd.h.
// Design: one flat loop per call, OpenMP static partition, the compiler vectorises the fp32 math to AVX-512
// (`-O3 -march=native -fopenmp`, `#pragma omp simd`); grads may be fp32 or bf16/fp16 (converted on load), the
// low-precision copy is written in the same pass so the H2D upload can start as soon as a tile finishes.
#include <cmath>
#include <cstdint>
#include <cstring>
#include <omp.h>

namespace {

inline float bf16_to_f(uint16_t h)
{
    uint32_t u = (uint32_t)h << 16;
    float f;
    std::memcpy(&f, &u, 4);
    return f;
}
inline uint16_t f_to_bf16(float f)
{
    uint32_t u;
    std::memcpy(&u, &f, 4);
    if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x0040u);  // NaN
    u += 0x7fffu + ((u >> 16) & 1u);                                             // round to nearest even
    return (uint16_t)(u >> 16);
}
inline float f16_to_f(uint16_t h)
{
    const uint32_t sign = (uint32_t)(h & 0x8000u) << 16;
    uint32_t exp = (h >> 10) & 0x1fu, man = h & 0x3ffu, u;
    if (exp == 0) {
        if (man == 0) {
            u = sign;
        } else {
            exp = 127 - 15 + 1;
            while ((man & 0x400u) == 0) {
                man <<= 1;
                exp--;
            }
            man &= 0x3ffu;
            u = sign | (exp << 23) | (man << 13);
        }
    } else if (exp == 31) {
        u = sign | 0x7f800000u | (man << 13);
    } else {
        u = sign | ((exp + 112) << 23) | (man << 13);
    }
    float f;
    std::memcpy(&f, &u, 4);
    return f;
}
inline uint16_t f_to_f16(float f)
{
    uint32_t x;
    std::memcpy(&x, &f, 4);
    const uint32_t sign = (x >> 16) & 0x8000u;
    int32_t exp = (int32_t)((x >> 23) & 0xff) - 127 + 15;
    uint32_t man = x & 0x7fffffu;
    if (exp >= 31) return (uint16_t)(sign | 0x7c00u | (((x & 0x7fffffffu) > 0x7f800000u) ? 0x200u : 0));
    if (exp <= 0) {
        if (exp < -10) return (uint16_t)sign;
        man |= 0x800000u;
        const uint32_t shift = (uint32_t)(14 - exp);
        uint32_t half = man >> shift;
        const uint32_t rem = man & ((1u << shift) - 1), mid = 1u << (shift - 1);
        if (rem > mid || (rem == mid && (half & 1))) half++;
        return (uint16_t)(sign | half);
    }
    uint32_t half = ((uint32_t)exp << 10) | (man >> 13);
    const uint32_t rem = man & 0x1fffu;
    if (rem > 0x1000u || (rem == 0x1000u && (half & 1))) half++;
    return (uint16_t)(sign | half);
}

// dtype codes: 0 fp32, 1 fp16, 2 bf16
inline float load_g(const void* g, int dt, int64_t i)
{
    if (dt == 0) return ((const float*)g)[i];
    if (dt == 2) return bf16_to_f(((const uint16_t*)g)[i]);
    return f16_to_f(((const uint16_t*)g)[i]);
}
inline void store_lp(void* lp, int dt, int64_t i, float v)
{
    if (dt == 0)
        ((float*)lp)[i] = v;
    else if (dt == 2)
        ((uint16_t*)lp)[i] = f_to_bf16(v);
    else
        ((uint16_t*)lp)[i] = f_to_f16(v);
}

constexpr int64_t TILE = 1 << 16;

}  // namespace

extern "C" {

int dsb200_cpu_num_threads() { return omp_get_max_threads(); }

// Adam / AdamW on n elements. adamw != 0 -> decoupled weight decay. lp may be null.
void dsb200_cpu_adam(float* p, const void* g, int g_dt, float* m, float* v, void* lp, int lp_dt, int64_t n, float lr,
                     float beta1, float beta2, float eps, float wd, int step, int bias_correction, int adamw,
                     float grad_scale)
{
    const float bc1 = bias_correction ? 1.f - std::pow(beta1, (float)step) : 1.f;
    const float bc2 = bias_correction ? 1.f - std::pow(beta2, (float)step) : 1.f;
    const float inv_bc1 = 1.f / bc1, inv_sqrt_bc2 = 1.f / std::sqrt(bc2);
    const int64_t ntiles = (n + TILE - 1) / TILE;
#pragma omp parallel for schedule(static)
    for (int64_t t = 0; t < ntiles; t++) {
        const int64_t lo = t * TILE, hi = (lo + TILE < n) ? lo + TILE : n;
        if (g_dt == 0) {
            const float* gf = (const float*)g;
#pragma omp simd
            for (int64_t i = lo; i < hi; i++) {
                float gi = gf[i] * grad_scale, pi = p[i];
                if (!adamw) gi += wd * pi;
                const float mi = beta1 * m[i] + (1.f - beta1) * gi;
                const float vi = beta2 * v[i] + (1.f - beta2) * gi * gi;
                float upd = (mi * inv_bc1) / (std::sqrt(vi) * inv_sqrt_bc2 + eps);
                if (adamw) upd += wd * pi;
                m[i] = mi;
                v[i] = vi;
                p[i] = pi - lr * upd;
            }
        } else {
            for (int64_t i = lo; i < hi; i++) {
                float gi = load_g(g, g_dt, i) * grad_scale, pi = p[i];
                if (!adamw) gi += wd * pi;
                const float mi = beta1 * m[i] + (1.f - beta1) * gi;
                const float vi = beta2 * v[i] + (1.f - beta2) * gi * gi;
                float upd = (mi * inv_bc1) / (std::sqrt(vi) * inv_sqrt_bc2 + eps);
                if (adamw) upd += wd * pi;
                m[i] = mi;
                v[i] = vi;
                p[i] = pi - lr * upd;
            }
        }
        if (lp != nullptr)
            for (int64_t i = lo; i < hi; i++) store_lp(lp, lp_dt, i, p[i]);
    }
}

// Undo one Adam step given the same gradient and hyper-parameters (speculative-step rollback).
void dsb200_cpu_adam_rollback(float* p, const void* g, int g_dt, float* m, float* v, int64_t n, float lr, float beta1,
                              float beta2, float eps, float wd, int step, int bias_correction, int adamw,
                              float grad_scale)
{
    const float bc1 = bias_correction ? 1.f - std::pow(beta1, (float)step) : 1.f;
    const float bc2 = bias_correction ? 1.f - std::pow(beta2, (float)step) : 1.f;
    const float inv_bc1 = 1.f / bc1, inv_sqrt_bc2 = 1.f / std::sqrt(bc2);
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; i++) {
        const float mi = m[i], vi = v[i];
        const float upd0 = (mi * inv_bc1) / (std::sqrt(vi) * inv_sqrt_bc2 + eps);
        float pi;
        if (adamw)
            pi = (p[i] + lr * upd0) / (1.f - lr * wd);
        else
            pi = p[i] + lr * upd0;
        float gi = load_g(g, g_dt, i) * grad_scale;
        if (!adamw) gi += wd * pi;
        m[i] = (mi - (1.f - beta1) * gi) / beta1;
        v[i] = (vi - (1.f - beta2) * gi * gi) / beta2;
        p[i] = pi;
    }
}

void dsb200_cpu_adagrad(float* p, const void* g, int g_dt, float* v, void* lp, int lp_dt, int64_t n, float lr,
                        float eps, float wd, float grad_scale)
{
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; i++) {
        float gi = load_g(g, g_dt, i) * grad_scale + wd * p[i];
        const float vi = v[i] + gi * gi;
        v[i] = vi;
        p[i] -= lr * gi / (std::sqrt(vi) + eps);
        if (lp != nullptr) store_lp(lp, lp_dt, i, p[i]);
    }
}

void dsb200_cpu_lion(float* p, const void* g, int g_dt, float* m, void* lp, int lp_dt, int64_t n, float lr,
                     float beta1, float beta2, float wd, float grad_scale)
{
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; i++) {
        const float gi = load_g(g, g_dt, i) * grad_scale;
        float pi = p[i] * (1.f - lr * wd);
        const float c = beta1 * m[i] + (1.f - beta1) * gi;
        pi -= lr * ((c > 0.f) ? 1.f : ((c < 0.f) ? -1.f : 0.f));
        m[i] = beta2 * m[i] + (1.f - beta2) * gi;
        p[i] = pi;
        if (lp != nullptr) store_lp(lp, lp_dt, i, pi);
    }
}

// fp32 -> bf16/fp16 cast of a host buffer (parallel), used when the optimizer is a generic torch one
void dsb200_cpu_cast(const float* src, void* dst, int dst_dt, int64_t n)
{
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < n; i++) store_lp(dst, dst_dt, i, src[i]);
}

double dsb200_cpu_sumsq(const float* x, int64_t n)
{
    double s = 0.0;
#pragma omp parallel for reduction(+ : s) schedule(static)
    for (int64_t i = 0; i < n; i++) s += (double)x[i] * (double)x[i];
    return s;
}

}  // extern "C"

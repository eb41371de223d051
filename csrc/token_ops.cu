// Row gather / scatter by index: the data movement of random layerwise token dropping (random-LTD), MoE token
// permutation and sparse-gradient packing. Rows are moved as 16-byte vectors, one CTA per few rows, enough rows
// in flight per CTA to cover HBM latency (each thread issues up to 4 independent 16 B loads before storing).
// Reference counterpart: csrc/random_ltd/{gather_scatter.cu,token_sort.cu,slice_attn_masks.cu}.
#include "common.cuh"

namespace {

constexpr int ROW_THREADS = 256;

// dst[i, :] = src[idx[i], :]          (gather)   or   dst[idx[i], :] = src[i, :]   (scatter)
template <typename IdxT, bool SCATTER>
__global__ void __launch_bounds__(ROW_THREADS) move_rows_kernel(const uint4* __restrict__ src, uint4* __restrict__ dst,
                                                                const IdxT* __restrict__ idx, int64_t n_rows,
                                                                int vec_per_row, int64_t src_stride_vec,
                                                                int64_t dst_stride_vec, int64_t bound)
{
    for (int64_t r = blockIdx.x; r < n_rows; r += gridDim.x) {
        const int64_t j = (int64_t)idx[r];
        if (j < 0 || j >= bound) continue;  // dropped token (MoE capacity) or padding
        const uint4* s = src + (SCATTER ? r : j) * src_stride_vec;
        uint4* d = dst + (SCATTER ? j : r) * dst_stride_vec;
        int c = threadIdx.x;
        for (; c + 3 * ROW_THREADS < vec_per_row; c += 4 * ROW_THREADS) {
            uint4 a0 = s[c], a1 = s[c + ROW_THREADS], a2 = s[c + 2 * ROW_THREADS], a3 = s[c + 3 * ROW_THREADS];
            d[c] = a0;
            d[c + ROW_THREADS] = a1;
            d[c + 2 * ROW_THREADS] = a2;
            d[c + 3 * ROW_THREADS] = a3;
        }
        for (; c < vec_per_row; c += ROW_THREADS) d[c] = s[c];
    }
}

// dst[idx[i], :] += w[i] * src[i, :]   (MoE combine / gradient of gather). Rows with equal idx may collide, so the
// caller guarantees uniqueness per launch (top-k slots are launched one k at a time) or passes atomic=1.
template <typename T, typename IdxT>
__global__ void __launch_bounds__(ROW_THREADS) scatter_add_rows_kernel(const T* __restrict__ src, T* __restrict__ dst,
                                                                       const IdxT* __restrict__ idx,
                                                                       const float* __restrict__ w, int64_t n_rows,
                                                                       int cols, int64_t src_stride,
                                                                       int64_t dst_stride, int64_t bound, int atomic)
{
    for (int64_t r = blockIdx.x; r < n_rows; r += gridDim.x) {
        const int64_t j = (int64_t)idx[r];
        if (j < 0 || j >= bound) continue;
        const float scale = w ? w[r] : 1.f;
        const T* s = src + r * src_stride;
        T* d = dst + j * dst_stride;
        for (int c = threadIdx.x; c < cols; c += ROW_THREADS) {
            const float v = scale * (float)s[c];
            if (atomic) {
                if constexpr (sizeof(T) == 4) atomicAdd((float*)&d[c], v);
                else atomicAdd(&d[c], (T)v);
            } else {
                d[c] = (T)((float)d[c] + v);
            }
        }
    }
}

// Sort each row of `n` int32 indices ascending (n <= 8192; bitonic in shared memory). random-LTD keeps the sampled
// token positions ordered so that causal masks stay valid.
__global__ void sort_rows_kernel(int* __restrict__ data, int n, int n_pow2)
{
    extern __shared__ int sm[];
    int* row = data + (int64_t)blockIdx.x * n;
    for (int i = threadIdx.x; i < n_pow2; i += blockDim.x) sm[i] = i < n ? row[i] : 0x7fffffff;
    __syncthreads();
    for (int k = 2; k <= n_pow2; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = threadIdx.x; i < n_pow2; i += blockDim.x) {
                const int p = i ^ j;
                if (p > i) {
                    const bool up = (i & k) == 0;
                    const int a = sm[i], b = sm[p];
                    if ((a > b) == up) {
                        sm[i] = b;
                        sm[p] = a;
                    }
                }
            }
            __syncthreads();
        }
    }
    for (int i = threadIdx.x; i < n; i += blockDim.x) row[i] = sm[i];
}

int grid_for(int64_t n_rows) { return (int)(n_rows < (int64_t)DSB_NUM_SMS * 8 ? n_rows : (int64_t)DSB_NUM_SMS * 8); }

}  // namespace

// idx_bytes: 4 (int32) or 8 (int64). row_bytes and both strides must be multiples of 16; `bound` is the number of rows
// of the indexed side (indices outside [0, bound) are skipped).
extern "C" int dsb200_move_rows(const void* src, void* dst, const void* idx, int idx_bytes, int64_t n_rows,
                                int64_t row_bytes, int64_t src_stride_bytes, int64_t dst_stride_bytes, int64_t bound,
                                int scatter, cudaStream_t stream)
{
    if (n_rows == 0) return 0;
    if ((row_bytes | src_stride_bytes | dst_stride_bytes) & 15 || ((uintptr_t)src | (uintptr_t)dst) & 15) {
        dsb200_set_error("move_rows: rows must be 16-byte aligned");
        return -1;
    }
    const int vec = (int)(row_bytes / 16);
    const int64_t ss = src_stride_bytes / 16, ds = dst_stride_bytes / 16;
    const int g = grid_for(n_rows);
#define LAUNCH(IT, SC) \
    move_rows_kernel<IT, SC><<<g, ROW_THREADS, 0, stream>>>((const uint4*)src, (uint4*)dst, (const IT*)idx, n_rows, vec, ss, ds, bound)
    if (idx_bytes == 4) {
        if (scatter) LAUNCH(int, true); else LAUNCH(int, false);
    } else if (idx_bytes == 8) {
        if (scatter) LAUNCH(int64_t, true); else LAUNCH(int64_t, false);
    } else {
        dsb200_set_error("move_rows: idx_bytes must be 4 or 8");
        return -1;
    }
#undef LAUNCH
    DSB_CHECK_LAUNCH("move_rows");
    return 0;
}

extern "C" int dsb200_scatter_add_rows(const void* src, void* dst, const void* idx, int idx_bytes, const float* w,
                                       int64_t n_rows, int cols, int64_t src_stride, int64_t dst_stride, int64_t bound,
                                       int atomic, int dt, cudaStream_t stream)
{
    if (n_rows == 0) return 0;
    const int g = grid_for(n_rows);
#define LAUNCH(T, IT) \
    scatter_add_rows_kernel<T, IT><<<g, ROW_THREADS, 0, stream>>>((const T*)src, (T*)dst, (const IT*)idx, w, n_rows, cols, src_stride, dst_stride, bound, atomic)
#define BY_IDX(T) do { if (idx_bytes == 4) LAUNCH(T, int); else LAUNCH(T, int64_t); } while (0)
    if (idx_bytes != 4 && idx_bytes != 8) {
        dsb200_set_error("scatter_add_rows: idx_bytes must be 4 or 8");
        return -1;
    }
    switch (dt) {
        case DSB_DT_F32: BY_IDX(float); break;
        case DSB_DT_F16: BY_IDX(__half); break;
        case DSB_DT_BF16: BY_IDX(__nv_bfloat16); break;
        default: dsb200_set_error("scatter_add_rows: bad dtype"); return -1;
    }
#undef BY_IDX
#undef LAUNCH
    DSB_CHECK_LAUNCH("scatter_add_rows");
    return 0;
}

extern "C" int dsb200_sort_rows_i32(int* data, int64_t n_rows, int n, cudaStream_t stream)
{
    if (n_rows == 0 || n <= 1) return 0;
    int p2 = 1;
    while (p2 < n) p2 <<= 1;
    if (p2 > 8192) {
        dsb200_set_error("sort_rows_i32: at most 8192 entries per row");
        return -1;
    }
    const int threads = p2 >= 1024 ? 1024 : (p2 < 32 ? 32 : p2);
    sort_rows_kernel<<<(unsigned)n_rows, threads, p2 * sizeof(int), stream>>>(data, n, p2);
    DSB_CHECK_LAUNCH("sort_rows_i32");
    return 0;
}

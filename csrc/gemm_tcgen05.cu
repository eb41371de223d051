// bf16 GEMM on the 5th-generation tensor cores: C[M,N] = A[M,K] * B[N,K]^T (both operands K-major, i.e. the
// torch.nn.Linear forward y = x W^T), fp32 accumulation in TMEM, bf16 output (+ optional fused SwiGLU epilogue).
//
// Reference: every dense GEMM of the reference is a cuBLAS call (csrc/transformer/cublas_wrappers.cu,
// module_inject/layers.py:771 F.linear). This is the framework's own tcgen05 kernel, the building block that the
// fused all-gather->GEMM / GEMM->reduce-scatter paths extend.
//
// Structure (persistent, one CTA per SM, 192 threads, warp-specialised):
//   warp 0      : TMA producer  - cp.async.bulk.tensor 2D loads of A (128x64) and B (256x64) tiles, 128B swizzle,
//                                 4-stage ring guarded by full/empty mbarriers
//   warp 1      : MMA issuer    - allocates TMEM (512 columns = 2 accumulator stages of 128x256 fp32), one elected
//                                 lane issues 4 x tcgen05.mma (M128 N256 K16) per k-block, tcgen05.commit frees the
//                                 smem stage / publishes the accumulator
//   warps 2..5  : epilogue      - tcgen05.ld 32 lanes x 32 columns -> registers -> bf16 -> 16-byte global stores,
//                                 overlapped with the MMAs of the next tile through the second accumulator stage
#include "gemm_common.cuh"

namespace {

// EPI: 0 = plain bf16 store; 1 = SwiGLU over interleaved (gate, up) column pairs is NOT used here - the fused
// gate_up layout keeps gate in columns [0, N/2) and up in [N/2, N), so SwiGLU stays a separate kernel.
template <int BLOCK_N>
__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_bf16_nt_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                    __nv_bfloat16* __restrict__ C, int M, int N, int K, int64_t ldc)
{
    constexpr int B_BYTES = Cfg<BLOCK_N>::B_BYTES, STAGE_BYTES = Cfg<BLOCK_N>::STAGE_BYTES;
    constexpr int STAGES = Cfg<BLOCK_N>::STAGES, TMEM_COLS = Cfg<BLOCK_N>::TMEM_COLS;
    (void)B_BYTES;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* tiles = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    Smem* sm = reinterpret_cast<Smem*>(tiles + STAGES * STAGE_BYTES);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int m_tiles = (M + BLOCK_M - 1) / BLOCK_M, n_tiles = (N + BLOCK_N - 1) / BLOCK_N;
    const int num_tiles = m_tiles * n_tiles;
    const int k_blocks = (K + BLOCK_K - 1) / BLOCK_K;

    if (threadIdx.x == 0) {
        for (int s = 0; s < STAGES; s++) {
            mbar_init(&sm->full[s], 1);
            mbar_init(&sm->empty[s], 1);
        }
        for (int s = 0; s < ACC_STAGES; s++) {
            mbar_init(&sm->tmem_full[s], 1);
            mbar_init(&sm->tmem_empty[s], 4);  // one arrive per epilogue warp
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_b) : "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&sm->tmem_base)),
                     "n"(TMEM_COLS)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = sm->tmem_base;

    if (warp == 0) {
        // ================= TMA producer =================
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
                const int m_blk = t % m_tiles, n_blk = t / m_tiles;
                for (int kb = 0; kb < k_blocks; kb++) {
                    mbar_wait(&sm->empty[stage], phase ^ 1);
                    uint8_t* a_dst = tiles + stage * STAGE_BYTES;
                    uint8_t* b_dst = a_dst + A_BYTES;
                    mbar_expect_tx(&sm->full[stage], STAGE_BYTES);
                    tma_load_2d(a_dst, &map_a, &sm->full[stage], kb * BLOCK_K, m_blk * BLOCK_M);
                    tma_load_2d(b_dst, &map_b, &sm->full[stage], kb * BLOCK_K, n_blk * BLOCK_N);
                    if (++stage == STAGES) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ================= MMA issuer =================
        if (lane == 0) {
            constexpr uint32_t idesc = make_idesc(BLOCK_M, BLOCK_N);
            int stage = 0, acc = 0;
            uint32_t phase = 0, acc_phase = 0;
            for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
                mbar_wait(&sm->tmem_empty[acc], acc_phase ^ 1);
                tc_fence_after();
                const uint32_t tmem_d = tmem_base + acc * BLOCK_N;
                for (int kb = 0; kb < k_blocks; kb++) {
                    mbar_wait(&sm->full[stage], phase);
                    tc_fence_after();
                    const uint32_t a_addr = smem_u32(tiles + stage * STAGE_BYTES);
                    const uint32_t b_addr = a_addr + A_BYTES;
                    const uint64_t adesc = make_smem_desc(a_addr), bdesc = make_smem_desc(b_addr);
#pragma unroll
                    for (int k = 0; k < BLOCK_K / UMMA_K; k++) {
                        // advance 32 bytes (16 bf16) along K inside the 128B-swizzled row: +2 in the >>4 address field
                        umma_bf16(tmem_d, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), idesc,
                                  (kb > 0 || k > 0) ? 1u : 0u);
                    }
                    umma_commit(&sm->empty[stage]);  // smem stage reusable once these MMAs retire
                    if (++stage == STAGES) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
                umma_commit(&sm->tmem_full[acc]);  // accumulator complete
                if (++acc == ACC_STAGES) {
                    acc = 0;
                    acc_phase ^= 1;
                }
            }
        }
    } else {
        // ================= epilogue (warps 2..5) =================
        const int lane_group = warp & 3;  // TMEM lanes [32*lane_group, +32) are accessible to this warp
        int acc = 0;
        uint32_t acc_phase = 0;
        for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
            const int m_blk = t % m_tiles, n_blk = t / m_tiles;
            mbar_wait(&sm->tmem_full[acc], acc_phase);
            tc_fence_after();
            const int row = m_blk * BLOCK_M + lane_group * 32 + lane;
            __nv_bfloat16* crow = C + (int64_t)row * ldc + (int64_t)n_blk * BLOCK_N;
#pragma unroll 1
            for (int c = 0; c < BLOCK_N; c += 32) {
                uint32_t r[32];
                const uint32_t taddr = tmem_base + ((uint32_t)(lane_group * 32) << 16) + (uint32_t)(acc * BLOCK_N + c);
                tmem_ld_32x32(taddr, r);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                const int col0 = n_blk * BLOCK_N + c;
                if (row < M) {
                    if (col0 + 32 <= N) {
#pragma unroll
                        for (int v = 0; v < 4; v++) {
                            uint4 pk;
                            __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
                            for (int j = 0; j < 4; j++)
                                h[j] = __floats2bfloat162_rn(__uint_as_float(r[v * 8 + 2 * j]),
                                                             __uint_as_float(r[v * 8 + 2 * j + 1]));
                            *reinterpret_cast<uint4*>(crow + c + v * 8) = pk;
                        }
                    } else {
                        for (int j = 0; j < 32; j++)
                            if (col0 + j < N) crow[c + j] = __float2bfloat16_rn(__uint_as_float(r[j]));
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&sm->tmem_empty[acc]);
            if (++acc == ACC_STAGES) {
                acc = 0;
                acc_phase ^= 1;
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(TMEM_COLS) : "memory");
    }
}

}  // namespace

extern "C" {

// C[M,N] (bf16, row stride ldc) = A[M,K] (row stride lda) * B[N,K]^T (row stride ldb); all bf16, K-major.
// Requirements: 16-byte aligned base pointers, lda/ldb multiples of 8 elements, ldc multiple of 8.
int dsb200_gemm_bf16_nt(const void* A, const void* B, void* C, int M, int N, int K, int64_t lda, int64_t ldb,
                        int64_t ldc, int max_ctas, int force_bn, cudaStream_t stream)
{
    if (M <= 0 || N <= 0 || K <= 0) return 0;
    if ((lda % 8) || (ldb % 8) || (ldc % 8) || ((uintptr_t)A & 15) || ((uintptr_t)B & 15) || ((uintptr_t)C & 15)) {
        dsb200_set_error("gemm_bf16_nt: operands must be 16-byte aligned with leading dimensions multiple of 8");
        return -1;
    }
    // pick BLOCK_N to minimise wave quantisation: cost ~ waves(tiles / 148) * BLOCK_N (+ small-tile penalty)
    const int m_tiles = (M + BLOCK_M - 1) / BLOCK_M;
    int best_bn = 256;
    double best_cost = 1e30;
    const int cands[3] = {256, 192, 128};
    const double penalty[3] = {1.0, 1.05, 1.25};  // measured: narrower tiles re-read A more often (L2-bound)
    for (int i = 0; i < 3; i++) {
        const int bn = cands[i];
        const long tiles_i = (long)m_tiles * ((N + bn - 1) / bn);
        const long waves = (tiles_i + DSB_NUM_SMS - 1) / DSB_NUM_SMS;
        const double cost = (double)waves * bn * penalty[i];
        if (cost < best_cost - 1e-9) {
            best_cost = cost;
            best_bn = bn;
        }
    }
    if (force_bn == 128 || force_bn == 192 || force_bn == 256) best_bn = force_bn;
    CUtensorMap map_a, map_b;
    if (make_map(&map_a, A, M, K, lda, BLOCK_M) != 0) return -2;
    if (make_map(&map_b, B, N, K, ldb, best_bn) != 0) return -2;
    const int tiles = m_tiles * ((N + best_bn - 1) / best_bn);
    int grid = tiles < DSB_NUM_SMS ? tiles : DSB_NUM_SMS;
    if (max_ctas > 0 && grid > max_ctas) grid = max_ctas;
#define GEMM_LAUNCH(BN)                                                                                              \
    do {                                                                                                             \
        static bool attr_set[64] = {};                                                                                \
        int dev_ = 0;                                                                                \
        cudaGetDevice(&dev_);                                                                                \
        if (!attr_set[dev_ & 63]) {                                                                                             \
            DSB_CHECK(cudaFuncSetAttribute(gemm_bf16_nt_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize,     \
                                           Cfg<BN>::SMEM_BYTES));                                                   \
            attr_set[dev_ & 63] = true;                                                                                         \
        }                                                                                                            \
        gemm_bf16_nt_kernel<BN><<<grid, NUM_THREADS, Cfg<BN>::SMEM_BYTES, stream>>>(map_a, map_b, (__nv_bfloat16*)C, \
                                                                                    M, N, K, ldc);                  \
    } while (0)
    if (best_bn == 256)
        GEMM_LAUNCH(256);
    else if (best_bn == 192)
        GEMM_LAUNCH(192);
    else
        GEMM_LAUNCH(128);
#undef GEMM_LAUNCH
    DSB_CHECK_LAUNCH("gemm_bf16_nt");
    return 0;
}

}  // extern "C"

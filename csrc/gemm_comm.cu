// tcgen05 GEMM fused with NVLink collectives over symmetric memory: C = A * B^T (bf16, fp32 accumulate in TMEM), where
//
//   AG_A       the rows of A live on the peers (rows_per_peer each); copy      all-gather -> GEMM   (AutoTP sequence-
//              CTAs pull them ONCE into the local buffer g_full              parallel input, Ulysses gather)
//   AG_B       the rows of B (= output features of a sharded weight) live    ZeRO-3 parameter all-gather fused into
//              on the peers and are staged into g_full the same way            the first GEMM that consumes it
//   SCATTER_N  column block j of C is stored into peer j's output buffer at   GEMM -> all-to-all (Ulysses QKV
//              row offset rank*M (peer j ends up with [world*M, N/world])      projection -> head-sharded layout)
//   RS_M       row block j of the partial product is stored into peer j's     GEMM -> reduce-scatter (AutoTP row-
//              staging slot [rank]; after an epoch barrier every rank sums      parallel linear)
//              its world slots into out[M/world, N]
//
// Reference baseline: separate NCCL collective + cuBLAS GEMM kernels (partition_parameters.py:1301 all-gather then
// F.linear; sequence/layer.py:264 all-to-all after the QKV linear; module_inject/layers.py:771 linear then
// all-reduce). Here the transfer of block i overlaps the MMAs of block i-1 inside ONE persistent kernel: dedicated
// copy CTAs land each peer's tile-row block in the local g_full with plain bulk async copies (no multicast) and
// publish a per-block flag; the GEMM CTAs wait on the flag and then load the block through the ordinary TMA/mbarrier
// ring. Remote output tiles leave from the epilogue warps while the tensor cores work on the next accumulator stage.
// Every tensor map in this file covers THIS GPU's memory only: a remote operand must cross NVLink once, and per-tile
// TMA loads from peer memory re-fetched the same weight tile for every m-tile (7x slower than NCCL when measured).
//
// Ordering uses the epoch flags of symm.cu: IN_READY ("my shard may be read", raised by the host-enqueued signal
// kernel before this launch), IN_DONE ("I finished reading your shard"), OUT_READY ("all my stores into your buffer are
// visible"), OUT_DONE ("I consumed my buffer", raised by the consumer, or by this kernel after the RS_M reduction).
#include "gemm_common.cuh"
#include "symm_common.cuh"
#include <cstring>

namespace {

enum { MODE_AG_A = 1, MODE_AG_B = 2, MODE_SCATTER_N = 4, MODE_RS_M = 8, MODE_REUSE_FULL = 16 };

struct MapTable {
    CUtensorMap m[MAX_WORLD];
};

struct CommArgs {
    int mode, rank, world;
    int rows_per_peer_a;  // AG_A: rows of A held by each peer (multiple of BLOCK_M)
    int rows_per_peer_b;  // AG_B: rows of B held by each peer (multiple of BLOCK_N)
    void* c_ptrs[MAX_WORLD];  // per-peer output base (SCATTER_N: [world*M, N/world]; RS_M: staging [world, M/world, N])
    PeerPtrs flags;
    int ch_in_ready, ch_in_done, ch_out_ready, ch_out_done;
    uint32_t epoch, reuse_epoch;
    unsigned int* tickets;  // 3 zero-initialised local words
    __nv_bfloat16* rs_out;  // RS_M result [M/world, N]
    int64_t rs_ld;
    // gathered operand (AG_A / AG_B): `n_copy` CTAs pull the peers' shards ONCE over NVLink into the local buffer
    // `g_full` with bulk async copies and publish per-chunk flags; the GEMM CTAs read chunks from `g_full`
    int n_copy;
    uint32_t* chunk_flags;        // local, one word per tile-row block of the gathered operand, compared with `epoch`
    const void* g_src[MAX_WORLD];  // peers' shard bases
    uint8_t* g_full;              // local gathered operand [world * rows_per_peer, K]
    int64_t chunk_bytes;          // bytes of one tile-row block (rows * K * 2)
    int chunks_per_peer;
};

__device__ __forceinline__ void wait_flag(const PeerPtrs& flags, int rank, int channel, int peer, uint32_t epoch)
{
    const uint32_t* f = reinterpret_cast<const uint32_t*>(flags.p[rank]) + channel * FLAG_STRIDE + peer;
    while ((int32_t)(ld_acquire_sys(f) - epoch) < 0) { __nanosleep(64); }
}

__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(smem_dst)),
                 "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void bulk_s2g(void* gdst, const void* smem_src, uint32_t bytes)
{
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(smem_src)),
                 "r"(bytes)
                 : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
__device__ __forceinline__ void st_release_gpu(uint32_t* addr, uint32_t v)
{
    asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_gpu(const uint32_t* addr)
{
    uint32_t v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(addr) : "memory");
    return v;
}

constexpr int COPY_PIECE = 32 * 1024;
constexpr int COPY_STAGES = 5;

// One elected thread of a copy CTA: remote shard -> smem ring -> local gathered buffer, chunk by chunk in the order
// the GEMM CTAs consume them. All data movement is done by the TMA engines (bulk async copies), so COPY_STAGES x 32 KB
// per CTA are in flight over NVLink without occupying registers or issue slots.
__device__ void copy_role(const CommArgs& ca, uint8_t* smem, uint64_t* bars, int copy_id, int n_chunks, int rot)
{
    const int rank = ca.rank;
    uint32_t seen = 1u << rank;
    uint32_t issued_in_stage[COPY_STAGES] = {0, 0, 0, 0, 0};
    int list_pos = 0;
    for (int i = 0; i < n_chunks; i++) {
        const int chunk = (i + rot) % n_chunks;
        const int peer = chunk / ca.chunks_per_peer;
        if (peer == rank) continue;  // my own shard is read in place
        if ((list_pos++ % ca.n_copy) != copy_id) continue;
        if (!(seen >> peer & 1)) {
            wait_flag(ca.flags, rank, ca.ch_in_ready, peer, ca.epoch);
            seen |= 1u << peer;
        }
        const uint8_t* src = reinterpret_cast<const uint8_t*>(ca.g_src[peer]) +
                             (int64_t)(chunk - peer * ca.chunks_per_peer) * ca.chunk_bytes;
        uint8_t* dst = ca.g_full + (int64_t)chunk * ca.chunk_bytes;
        if (ca.mode & 64) {  // debug: data already in place, only publish the flag
            __threadfence();
            st_release_gpu(ca.chunk_flags + chunk, ca.epoch);
            continue;
        }
        const int pieces = (int)((ca.chunk_bytes + COPY_PIECE - 1) / COPY_PIECE);
        auto piece_bytes = [&](int q) {
            const int64_t left = ca.chunk_bytes - (int64_t)q * COPY_PIECE;
            return (uint32_t)(left < COPY_PIECE ? left : COPY_PIECE);
        };
        auto load = [&](int q) {
            const int st = q % COPY_STAGES;
            mbar_expect_tx(&bars[st], piece_bytes(q));
            bulk_g2s(smem + st * COPY_PIECE, src + (int64_t)q * COPY_PIECE, piece_bytes(q), &bars[st]);
        };
        for (int q = 0; q < pieces && q < COPY_STAGES; q++) load(q);
        for (int q = 0; q < pieces; q++) {
            const int st = q % COPY_STAGES;
            mbar_wait(&bars[st], issued_in_stage[st] & 1);
            issued_in_stage[st]++;
            if (ca.mode & 32) {  // debug: generic-proxy copy out of shared memory instead of the bulk store
                const uint4* sp = reinterpret_cast<const uint4*>(smem + st * COPY_PIECE);
                uint4* gp = reinterpret_cast<uint4*>(dst + (int64_t)q * COPY_PIECE);
                for (uint32_t b = 0; b < piece_bytes(q) / 16; b++) gp[b] = sp[b];
            } else {
                bulk_s2g(dst + (int64_t)q * COPY_PIECE, smem + st * COPY_PIECE, piece_bytes(q));
            }
            if (q + COPY_STAGES < pieces) {
                asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");  // smem stage free again
                load(q + COPY_STAGES);
            }
        }
        asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");  // chunk completely written
        asm volatile("fence.proxy.async;" ::: "memory");
        __threadfence();
        st_release_gpu(ca.chunk_flags + chunk, ca.epoch);
    }
}

template <int BLOCK_N>
__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_comm_kernel(const __grid_constant__ MapTable maps_a, const __grid_constant__ MapTable maps_b,
                 const __grid_constant__ CommArgs ca, __nv_bfloat16* __restrict__ C, int M, int N, int K, int64_t ldc)
{
    constexpr int STAGE_BYTES = Cfg<BLOCK_N>::STAGE_BYTES;
    constexpr int STAGES = Cfg<BLOCK_N>::STAGES, TMEM_COLS = Cfg<BLOCK_N>::TMEM_COLS;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* tiles = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    Smem* sm = reinterpret_cast<Smem*>(tiles + STAGES * STAGE_BYTES);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int m_tiles = (M + BLOCK_M - 1) / BLOCK_M, n_tiles = (N + BLOCK_N - 1) / BLOCK_N;
    const int num_tiles = m_tiles * n_tiles;
    const int k_blocks = (K + BLOCK_K - 1) / BLOCK_K;
    const int mode = ca.mode, rank = ca.rank, world = ca.world;
    // start every rank on the tiles whose remote side is itself, then walk the ring: spreads NVLink traffic and lets
    // the first tiles run before any peer flag has to be observed
    const int m_rot = (mode & (MODE_AG_A | MODE_RS_M)) ? rank * (m_tiles / world) : 0;
    const int n_rot = (mode & (MODE_AG_B | MODE_SCATTER_N)) ? rank * (n_tiles / world) : 0;

    const int n_copy = ca.n_copy;
    const bool is_copy = (int)blockIdx.x < n_copy;
    const int gemm_cta = (int)blockIdx.x - n_copy, gemm_ctas = (int)gridDim.x - n_copy;
    if (threadIdx.x == 0) {
        for (int s = 0; s < MAX_STAGES; s++) {  // all of them: the copy CTAs use COPY_STAGES (> STAGES for wide tiles)
            mbar_init(&sm->full[s], 1);
            mbar_init(&sm->empty[s], 1);
        }
        for (int s = 0; s < ACC_STAGES; s++) {
            mbar_init(&sm->tmem_full[s], 1);
            mbar_init(&sm->tmem_empty[s], 4);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (is_copy) {
        __syncthreads();
        if (threadIdx.x == 0) {
            static_assert(COPY_STAGES <= MAX_STAGES, "copy ring uses the `full` barriers");
            if (mode & MODE_AG_B) copy_role(ca, tiles, sm->full, blockIdx.x, n_tiles, n_rot);
            else copy_role(ca, tiles, sm->full, blockIdx.x, m_tiles, m_rot);
        }
        __syncthreads();
    }
    if (!is_copy && warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&sm->tmem_base)),
                     "n"(TMEM_COLS)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = is_copy ? 0u : sm->tmem_base;

    if (is_copy) {
        // nothing else to do until the collective tail
    } else if (warp == 0) {
        // ================= TMA producer (local and gathered operand tiles) =================
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0, seen = 1u << rank;
            for (int t = gemm_cta; t < num_tiles; t += gemm_ctas) {
                const int m_blk = (t % m_tiles + m_rot) % m_tiles, n_blk = (t / m_tiles + n_rot) % n_tiles;
                int pa = 0, pb = 0, a_row = m_blk * BLOCK_M, b_row = n_blk * BLOCK_N;
                (void)seen;
                if (mode & MODE_AG_A) {
                    pa = a_row / ca.rows_per_peer_a;
                    a_row -= pa * ca.rows_per_peer_a;
                    if (pa != rank && !(mode & MODE_REUSE_FULL)) {  // gathered by the copy CTAs: wait for this tile-row block to land locally
                        while ((int32_t)(ld_acquire_gpu(ca.chunk_flags + m_blk) - ca.epoch) < 0) __nanosleep(32);
                        asm volatile("fence.proxy.async;" ::: "memory");
                    }
                }
                if (mode & MODE_AG_B) {
                    pb = b_row / ca.rows_per_peer_b;
                    b_row -= pb * ca.rows_per_peer_b;
                    if (pb != rank && !(mode & MODE_REUSE_FULL)) {
                        while ((int32_t)(ld_acquire_gpu(ca.chunk_flags + n_blk) - ca.epoch) < 0) __nanosleep(32);
                        asm volatile("fence.proxy.async;" ::: "memory");
                    }
                }
                const CUtensorMap* ma = &maps_a.m[pa];
                const CUtensorMap* mb = &maps_b.m[pb];
                for (int kb = 0; kb < k_blocks; kb++) {
                    mbar_wait(&sm->empty[stage], phase ^ 1);
                    uint8_t* a_dst = tiles + stage * STAGE_BYTES;
                    uint8_t* b_dst = a_dst + A_BYTES;
                    mbar_expect_tx(&sm->full[stage], STAGE_BYTES);
                    tma_load_2d(a_dst, ma, &sm->full[stage], kb * BLOCK_K, a_row);
                    tma_load_2d(b_dst, mb, &sm->full[stage], kb * BLOCK_K, b_row);
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
            for (int t = gemm_cta; t < num_tiles; t += gemm_ctas) {
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
                    for (int k = 0; k < BLOCK_K / UMMA_K; k++)
                        umma_bf16(tmem_d, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), idesc,
                                  (kb > 0 || k > 0) ? 1u : 0u);
                    umma_commit(&sm->empty[stage]);
                    if (++stage == STAGES) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
                umma_commit(&sm->tmem_full[acc]);
                if (++acc == ACC_STAGES) {
                    acc = 0;
                    acc_phase ^= 1;
                }
            }
        }
    } else {
        // ================= epilogue (warps 2..5): local or peer stores =================
        const int lane_group = warp & 3;
        int acc = 0;
        uint32_t acc_phase = 0, seen = (ca.reuse_epoch == 0) ? 0xffffffffu : (1u << rank);
        const int cols_per_peer = N / world, rows_per_peer = M / world;
        for (int t = gemm_cta; t < num_tiles; t += gemm_ctas) {
            const int m_blk = (t % m_tiles + m_rot) % m_tiles, n_blk = (t / m_tiles + n_rot) % n_tiles;
            mbar_wait(&sm->tmem_full[acc], acc_phase);
            tc_fence_after();
            const int row = m_blk * BLOCK_M + lane_group * 32 + lane;
            const int col_base = n_blk * BLOCK_N;
            __nv_bfloat16* crow;
            int peer = rank;
            if (mode & MODE_SCATTER_N) {
                peer = col_base / cols_per_peer;
                crow = reinterpret_cast<__nv_bfloat16*>(ca.c_ptrs[peer]) +
                       ((int64_t)rank * M + row) * cols_per_peer + (col_base - peer * cols_per_peer);
            } else if (mode & MODE_RS_M) {
                peer = (m_blk * BLOCK_M) / rows_per_peer;
                crow = reinterpret_cast<__nv_bfloat16*>(ca.c_ptrs[peer]) +
                       ((int64_t)rank * rows_per_peer + (row - peer * rows_per_peer)) * N + col_base;
            } else {
                crow = C + (int64_t)row * ldc + col_base;
            }
            if ((mode & (MODE_SCATTER_N | MODE_RS_M)) && !(seen >> peer & 1)) {
                // the destination buffer is reused: its owner must have consumed the previous contents
                if (lane == 0) wait_flag(ca.flags, rank, ca.ch_out_done, peer, ca.reuse_epoch);
                __syncwarp();
                seen |= 1u << peer;
            }
#pragma unroll 1
            for (int c = 0; c < BLOCK_N; c += 32) {
                uint32_t r[32];
                const uint32_t taddr = tmem_base + ((uint32_t)(lane_group * 32) << 16) + (uint32_t)(acc * BLOCK_N + c);
                tmem_ld_32x32(taddr, r);
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                if (row < M && col_base + c + 32 <= N) {
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
                } else if (row < M) {
                    for (int j = 0; j < 32; j++)
                        if (col_base + c + j < N) crow[c + j] = __float2bfloat16_rn(__uint_as_float(r[j]));
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
    if (!is_copy && warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(TMEM_COLS) : "memory");
    }

    // ================= collective tail =================
    __shared__ bool last;
    if (threadIdx.x == 0) {
        __threadfence_system();  // my stores into peer buffers / my reads of peer shards are complete
        last = (atomicInc(ca.tickets, gridDim.x - 1) == gridDim.x - 1);
    }
    __syncthreads();
    if (last) {
        if ((mode & (MODE_AG_A | MODE_AG_B)) && !(mode & MODE_REUSE_FULL))
            signal_all(ca.flags, rank, world, ca.ch_in_done, ca.epoch);
        if (mode & (MODE_SCATTER_N | MODE_RS_M)) signal_all(ca.flags, rank, world, ca.ch_out_ready, ca.epoch);
    }
    if (mode & (MODE_SCATTER_N | MODE_RS_M)) {
        // every CTA leaves only when all peers' contributions to MY buffer are visible: kernel completion then means
        // "output complete" for the consumer on this stream
        wait_all(ca.flags, rank, world, ca.ch_out_ready, ca.epoch);
    }
    if (mode & MODE_RS_M) {
        // phase 2: out[r, :] = sum_p staging[p][r, :] for my M/world rows (all CTAs, 16-byte vectors, fp32 sums)
        const int rows_per_peer = M / world;
        const int64_t nvec = (int64_t)rows_per_peer * N / 8;
        const __nv_bfloat16* stage = reinterpret_cast<const __nv_bfloat16*>(ca.c_ptrs[rank]);
        const int64_t slot = (int64_t)rows_per_peer * N;
        for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * blockDim.x) {
            float acc8[8] = {0, 0, 0, 0, 0, 0, 0, 0};
            for (int p = 0; p < world; p++) {
                const uint4 raw = *reinterpret_cast<const uint4*>(stage + p * slot + i * 8);
                const __nv_bfloat16* e = reinterpret_cast<const __nv_bfloat16*>(&raw);
#pragma unroll
                for (int k = 0; k < 8; k++) acc8[k] += __bfloat162float(e[k]);
            }
            dsb::f8 o;
#pragma unroll
            for (int k = 0; k < 8; k++) o.v[k] = acc8[k];
            const int64_t e0 = i * 8;
            dsb::store8(ca.rs_out + (e0 / N) * ca.rs_ld + (e0 % N), o);
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence();
            last = (atomicInc(ca.tickets + 1, gridDim.x - 1) == gridDim.x - 1);
        }
        __syncthreads();
        if (last) signal_all(ca.flags, rank, world, ca.ch_out_done, ca.epoch);  // my staging may be overwritten
    }
}

}  // namespace

extern "C" {

// a_ptrs / b_ptrs: `world` peer pointers when the operand is gathered (mode bit set), else element [0] is the local
// operand. M, N, K are the GLOBAL problem sizes. c_ptrs: per-peer output bases for SCATTER_N / RS_M (else C is used).
// Constraints: M % (world*128) == 0 for AG_A / RS_M; N % (world*block_n) == 0 for AG_B / SCATTER_N; K % 8 == 0.
// Returns the tile width used (> 0) or < 0 on error.
int dsb200_gemm_comm_bf16_nt(int mode, const uint64_t* a_ptrs, const uint64_t* b_ptrs, void* C,
                             const uint64_t* c_ptrs, void* rs_out, int M, int N, int K, int64_t lda, int64_t ldb,
                             int64_t ldc, const uint64_t* peer_flags, int rank, int world, int ch_in_ready,
                             int ch_in_done, int ch_out_ready, int ch_out_done, unsigned int epoch,
                             unsigned int reuse_epoch, unsigned int* tickets, void* g_full, void* chunk_flags,
                             int n_copy, int max_ctas, cudaStream_t stream)
{
    if (world < 1 || world > MAX_WORLD || M <= 0 || N <= 0 || K <= 0 || (lda % 8) || (ldb % 8)) {
        dsb200_set_error("gemm_comm: bad world size / shape / leading dimension");
        return -1;
    }
    // the tile width must divide the per-peer extent of whichever dimension is distributed
    int bn = 0;
    const int cands[3] = {256, 192, 128};
    for (int i = 0; i < 3 && bn == 0; i++) {
        const int c = cands[i];
        bool ok = true;
        if (mode & (MODE_AG_B | MODE_SCATTER_N)) ok = ok && (N % (world * c) == 0);
        else ok = ok && (N % c == 0 || c == 128);
        if (ok) bn = c;
    }
    if (bn == 0 || ((mode & (MODE_AG_A | MODE_RS_M)) && M % (world * BLOCK_M) != 0) ||
        ((mode & MODE_RS_M) && N % 8 != 0)) {
        dsb200_set_error("gemm_comm: distributed dimension is not a multiple of world * tile size");
        return -1;
    }
    MapTable ta, tb;
    memset(&ta, 0, sizeof(ta));
    memset(&tb, 0, sizeof(tb));
    if (mode & MODE_REUSE_FULL) n_copy = 0;  // the gathered operand is still resident in g_full from an earlier call
    if ((mode & (MODE_AG_A | MODE_AG_B)) &&
        (g_full == nullptr || (!(mode & MODE_REUSE_FULL) && (chunk_flags == nullptr || n_copy < 1)))) {
        dsb200_set_error("gemm_comm: all-gather modes need a local gather buffer, chunk flags and n_copy >= 1");
        return -1;
    }
    if (mode & MODE_AG_A) {
        if (lda != K) {
            dsb200_set_error("gemm_comm: the gathered operand must have contiguous rows (ld == K)");
            return -1;
        }
        for (int p = 0; p < world; p++) {
            const void* base = (p == rank) ? (const void*)a_ptrs[p]
                                           : (const void*)((const uint8_t*)g_full + (int64_t)p * (M / world) * K * 2);
            if (make_map(&ta.m[p], base, M / world, K, lda, BLOCK_M) != 0) return -2;
        }
    } else if (make_map(&ta.m[0], (const void*)a_ptrs[0], M, K, lda, BLOCK_M) != 0) {
        return -2;
    }
    if (mode & MODE_AG_B) {
        if (ldb != K) {
            dsb200_set_error("gemm_comm: the gathered operand must have contiguous rows (ld == K)");
            return -1;
        }
        for (int p = 0; p < world; p++) {
            const void* base = (p == rank) ? (const void*)b_ptrs[p]
                                           : (const void*)((const uint8_t*)g_full + (int64_t)p * (N / world) * K * 2);
            if (make_map(&tb.m[p], base, N / world, K, ldb, bn) != 0) return -2;
        }
    } else if (make_map(&tb.m[0], (const void*)b_ptrs[0], N, K, ldb, bn) != 0) {
        return -2;
    }
    CommArgs ca;
    memset(&ca, 0, sizeof(ca));
    ca.mode = mode;
    ca.rank = rank;
    ca.world = world;
    ca.rows_per_peer_a = M / world;
    ca.rows_per_peer_b = N / world;
    for (int p = 0; p < world; p++) ca.c_ptrs[p] = c_ptrs ? (void*)c_ptrs[p] : nullptr;
    ca.flags = make_peer_ptrs(peer_flags, world);
    ca.ch_in_ready = ch_in_ready;
    ca.ch_in_done = ch_in_done;
    ca.ch_out_ready = ch_out_ready;
    ca.ch_out_done = ch_out_done;
    ca.epoch = epoch;
    ca.reuse_epoch = reuse_epoch;
    ca.tickets = tickets;
    ca.rs_out = (__nv_bfloat16*)rs_out;
    ca.rs_ld = ldc;
    ca.n_copy = 0;
    if (mode & (MODE_AG_A | MODE_AG_B)) {
        const uint64_t* src = (mode & MODE_AG_A) ? a_ptrs : b_ptrs;
        const int rows = (mode & MODE_AG_A) ? BLOCK_M : bn;
        ca.n_copy = n_copy;
        ca.chunk_flags = (uint32_t*)chunk_flags;
        ca.g_full = (uint8_t*)g_full;
        ca.chunk_bytes = (int64_t)rows * K * 2;
        ca.chunks_per_peer = ((mode & MODE_AG_A) ? M : N) / world / rows;
        for (int p = 0; p < world; p++) ca.g_src[p] = (const void*)src[p];
    }
    const int m_tiles = (M + BLOCK_M - 1) / BLOCK_M;
    const int tiles = m_tiles * ((N + bn - 1) / bn);
    int grid = tiles < DSB_NUM_SMS - ca.n_copy ? tiles : DSB_NUM_SMS - ca.n_copy;
    if (max_ctas > 0 && grid > max_ctas) grid = max_ctas;
    grid += ca.n_copy;
#define COMM_LAUNCH(BN)                                                                                             \
    do {                                                                                                            \
        static bool attr_set[64] = {};                                                                               \
        int dev_ = 0;                                                                               \
        cudaGetDevice(&dev_);                                                                               \
        if (!attr_set[dev_ & 63]) {                                                                                            \
            DSB_CHECK(cudaFuncSetAttribute(gemm_comm_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize,       \
                                           Cfg<BN>::SMEM_BYTES));                                                  \
            attr_set[dev_ & 63] = true;                                                                                        \
        }                                                                                                           \
        gemm_comm_kernel<BN><<<grid, NUM_THREADS, Cfg<BN>::SMEM_BYTES, stream>>>(ta, tb, ca, (__nv_bfloat16*)C, M, \
                                                                                 N, K, ldc);                       \
    } while (0)
    if (bn == 256)
        COMM_LAUNCH(256);
    else if (bn == 192)
        COMM_LAUNCH(192);
    else
        COMM_LAUNCH(128);
#undef COMM_LAUNCH
    DSB_CHECK_LAUNCH("gemm_comm_bf16_nt");
    return bn;
}

}  // extern "C"

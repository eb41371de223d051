/* dsb200.h - C ABI of the deepspeed_b200 native libraries.
 *
 *   libdsb200.so      sm_100a CUDA kernels (csrc/ *.cu), built with nvcc -gencode arch=compute_100a,code=sm_100a
 *   libdsb200_cpu.so  host runtime (csrc/cpu/ *.cpp): AVX-512/OpenMP optimizers, asynchronous file I/O
 *
 * Conventions
 *   - every launcher takes raw device pointers, sizes and the cudaStream_t to launch on; nothing is synchronised;
 *   - return value: 0 on success, non-zero on failure; dsb200_last_error() returns the reason (thread local);
 *   - dtype codes: 0 = fp32, 1 = fp16, 2 = bf16;
 *   - unless said otherwise a launcher allocates nothing: scratch buffers are passed in by the caller.
 * Each declaration cites the reference (deepspeedai/DeepSpeed) code whose behaviour it covers.
 */
#ifndef DSB200_H
#define DSB200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#ifndef __DRIVER_TYPES_H__
typedef struct CUstream_st* cudaStream_t;
#endif

/* ---------------------------------------------------------------- runtime ------------------------------- */
const char* dsb200_last_error(void);
int dsb200_version(void);

/* ---------------------------------------------------------------- reductions / optimizers ---------------- */
/* fp32 sum of squares of x[n]. scratch: >= 1025 floats, word [1024] (ticket) zero before the first use; it resets
 * itself. Deterministic two-stage reduction in one launch. Reference: get_global_norm_of_tensors
 * (runtime/utils.py:878), multi_tensor_l2norm. */
int dsb200_sumsq(const void* x, int64_t n, int dt, float* scratch, float* out, cudaStream_t stream);

/* Same reduction for callers that cannot supply scratch: a process whose own pool already holds all device memory
 * (a raw cudaMalloc would fail or shrink that pool), a latency-critical loop that must avoid the device-wide
 * synchronisation of cudaMalloc/cudaFree, or a C program with no allocator to pass on. The partial sums live in a
 * __device__ array inside the library (about 4.3 KB on every GPU that loads it, kept until the process exits); the
 * call allocates nothing. No counterpart in the reference.
 *   - Limit: n <= DSB200_SUMSQ_NOALLOC_MAX_ELEMS (2^34 elements); larger inputs are refused with an error.
 *   - The first call on each device looks up the array's device address (cudaGetSymbolAddress, that device
 *     current); make that first call before starting any stream capture.
 *   - All calls on a GPU share the array, which keeps the previous call's contents; each call clears the one word
 *     that must start at zero. Calls that overlap on the same GPU (other streams or threads) race: serialise them. */
#define DSB200_SUMSQ_NOALLOC_MAX_ELEMS (1LL << 34)
int dsb200_sumsq_noalloc(const void* x, int64_t n, int dt, float* out, cudaStream_t stream);

/* One launch updates every chunk of a device-resident chunk table {p,g,m,v,lp,n,group,dtypes} (48 bytes/entry).
 * kind: 0 Adam/AdamW, 1 Lion, 2 Adagrad, 3 SGD-momentum. hyper_host: n_groups x 8 floats
 * (lr, beta1, beta2, eps, wd, bias_corr1, bias_corr2, mode). Fuses grad unscale/clip (grad_scale), the fp32 update and
 * the low-precision weight write-back. Reference: csrc/adam/multi_tensor_adam.cu:33-127,
 * csrc/lion/multi_tensor_lion.cu, multi_tensor_apply.cuh:38-125. */
int dsb200_multi_tensor_optim(int kind, const void* table, int n_chunks, const float* hyper_host, int n_groups,
                              float grad_scale, cudaStream_t stream);

/* LAMB on one tensor: phase 1 update + per-block norms (partials: >= 4096 floats), phase 2 trust ratio + apply.
 * coeff_out[0..2] = |w|, |update|, clamped ratio. Reference: csrc/lamb/fused_lamb_cuda_kernel.cu:357-453. */
int dsb200_lamb_step(float* p, float* g, float* m, float* v, void* lp, int64_t n, int lpdt, float* partials,
                     float* coeff_out, float lr, float b1, float b2, float eps, float wd, float bc1, float bc2,
                     int eps_mode, float max_coeff, float min_coeff, cudaStream_t stream);

/* ---------------------------------------------------------------- normalisation -------------------------- */
/* y = rmsnorm(x [+ residual]) * gamma; res_out (optional) receives x + residual; stats[rows] = rstd.
 * Reference: csrc/transformer/inference/csrc/rms_norm.cu:22,93 (forward only in the reference). */
int dsb200_rmsnorm_fwd(const void* x, const void* residual, const void* gamma, void* y, void* res_out, float* stats,
                       int rows, int cols, float eps, int dt, cudaStream_t stream);
/* scratch: fp32 [min(rows, 592) * cols] partial dgamma rows. dres (optional) is added to dx. */
int dsb200_rmsnorm_bwd(const void* dy, const void* x, const void* gamma, const float* stats, const void* dres,
                       void* dx, void* dgamma, float* scratch, int rows, int cols, int accumulate, int dt,
                       cudaStream_t stream);
/* LayerNorm; stats[rows*2] = (mean, rstd). Reference: csrc/transformer/normalize_kernels.cu. */
int dsb200_layernorm_fwd(const void* x, const void* residual, const void* gamma, const void* beta, void* y,
                         void* res_out, float* stats, int rows, int cols, float eps, int dt, cudaStream_t stream);
/* scratch: fp32 [2 * min(rows, 592) * cols]. */
int dsb200_layernorm_bwd(const void* dy, const void* x, const void* gamma, const float* stats, const void* dres,
                         void* dx, void* dgamma, void* dbeta, float* scratch, int rows, int cols, int accumulate,
                         int dt, cudaStream_t stream);

/* ---------------------------------------------------------------- elementwise ---------------------------- */
/* In-place half-rotation RoPE on x[tokens, heads, dim]; backward != 0 applies the transpose rotation.
 * Reference: csrc/transformer/inference/csrc/apply_rotary_pos_emb.cu:23-169. */
int dsb200_rope(void* x, const float* cos_tab, const float* sin_tab, const int* pos, int tokens, int heads, int dim,
                int64_t token_stride, int seq_len, int backward, int dt, cudaStream_t stream);
/* y = silu(gate) * up (inputs may be halves of one fused tensor: in_stride). Reference: swiglu_triton.py:36-70. */
int dsb200_swiglu_fwd(const void* gate, const void* up, void* y, int64_t rows, int inter, int64_t in_stride, int dt,
                      cudaStream_t stream);
int dsb200_swiglu_bwd(const void* dy, const void* gate, const void* up, void* dgate, void* dup, int64_t rows, int inter,
                      int64_t in_stride, int dt, cudaStream_t stream);
/* Row-wise softmax cross entropy; write_grad != 0 overwrites logits with (softmax - onehot) * grad_scale.
 * Reference: TiledFusedLogitsLoss (runtime/sequence_parallel/ulysses_sp.py:1065). */
int dsb200_ce_fwd(void* logits, const int64_t* target, float* loss, float* lse, int rows, int vocab, int64_t row_stride,
                  int ignore_index, int write_grad, float grad_scale, int dt, cudaStream_t stream);
/* as dsb200_ce_fwd with the gradient scale multiplied by a device scalar (e.g. 1 / #valid tokens without a host sync) */
int dsb200_ce_fwd_ex(void* logits, const int64_t* target, float* loss, float* lse, int rows, int vocab,
                     int64_t row_stride, int ignore_index, int write_grad, float grad_scale, const float* grad_scale_dev,
                     int dt, cudaStream_t stream);
/* x (fp16/bf16) *= *scale_dev, vectorised; a no-op when the device scalar is 1 (incoming loss gradient of 1) */
int dsb200_scale_by_dev(void* x, int64_t n, int dt, const float* scale_dev, cudaStream_t stream);

/* ---------------------------------------------------------------- fused collective + compute (symm_fused.cu) */
/* Reduce-scatter my gradient shard out of every peer's bucket and apply AdamW to the fp32 master / moments (+ bf16
 * copy) in the same kernel. seg_bounds[nseg+1]: element offsets (multiples of 8) of the parameter-group runs inside the
 * shard; hyper[nseg][8] = {lr, beta1, beta2, eps, wd, bias_corr1, bias_corr2, mode (1 = AdamW)}. Returns the grid size.
 * Reference: reduce_scatter_coalesced + a separate FusedAdam launch (coalesced_collectives.py:158, fused_adam.py:195). */
int dsb200_symm_rs_adam(const uint64_t* peer_bufs, const uint64_t* peer_flags, int rank, int world,
                        int64_t shard_numel, int in_dt, float grad_scale, float* p, float* m, float* v, void* lp,
                        int nseg, const int64_t* seg_bounds, const float* hyper, float* sumsq_partials,
                        unsigned int* ticket, int ch_ready, int ch_done, unsigned int epoch, int ctas,
                        cudaStream_t stream);
/* ZeRO++ qwZ all-gather in one kernel: block-scaled e4m3 quantize into the symmetric staging buffer, NVLink pull of the
 * peers' payload + scales, dequantize into `full`. Reference: partition_parameters.py:1248 + CUDAQuantizer. */
int64_t dsb200_symm_fp8_stage_bytes(int64_t shard_elems, int block);
int dsb200_symm_all_gather_fp8(const void* my_shard, int dt, const uint64_t* peer_stage, const uint64_t* peer_flags,
                               int rank, int world, int64_t shard_elems, int block, void* full,
                               unsigned int* tickets, int ch_ready, int ch_done, unsigned int epoch,
                               unsigned int reuse_epoch, int ctas, cudaStream_t stream);

/* ---------------------------------------------------------------- tcgen05 GEMM + collective (gemm_comm.cu) */
/* C = A * B^T (bf16, fp32 accumulate in TMEM) fused with an NVLink collective over symmetric memory. mode bits:
 *   1 AG_A  rows of A live on the peers   2 AG_B  rows of B live on the peers (ZeRO-3 weight all-gather -> GEMM):
 *           n_copy CTAs pull each remote tile-row block ONCE with bulk async copies into g_full and publish
 *           chunk_flags; the GEMM CTAs read them from there through per-peer TMA maps;
 *   4 SCATTER_N  column block j of C is stored to peer j's buffer c_ptrs[j] at row offset rank*M (GEMM -> all-to-all);
 *   8 RS_M  row block j goes to peer j's staging slot [rank], then every rank sums its slots into rs_out
 *           (GEMM -> reduce-scatter);   16 REUSE_FULL  the gathered operand is still resident in g_full.
 * M, N, K are global sizes; the distributed dimension must be a multiple of world * tile. tickets: 3 zeroed words.
 * Returns the tile width used (> 0). Reference: NCCL collective + cuBLAS GEMM as separate kernels
 * (partition_parameters.py:1301, sequence/layer.py:264, module_inject/layers.py:771). */
int dsb200_gemm_comm_bf16_nt(int mode, const uint64_t* a_ptrs, const uint64_t* b_ptrs, void* C,
                             const uint64_t* c_ptrs, void* rs_out, int M, int N, int K, int64_t lda, int64_t ldb,
                             int64_t ldc, const uint64_t* peer_flags, int rank, int world, int ch_in_ready,
                             int ch_in_done, int ch_out_ready, int ch_out_done, unsigned int epoch,
                             unsigned int reuse_epoch, unsigned int* tickets, void* g_full, void* chunk_flags,
                             int n_copy, int max_ctas, cudaStream_t stream);

/* ---------------------------------------------------------------- token movement ------------------------- */
/* scatter == 0: dst[i,:] = src[idx[i],:]; scatter != 0: dst[idx[i],:] = src[i,:]. idx is int32 (idx_bytes 4) or int64
 * (8); rows and strides are multiples of 16 bytes; indices outside [0, bound) are skipped.
 * Reference: csrc/random_ltd/gather_scatter.cu, MoE token permutation. */
int dsb200_move_rows(const void* src, void* dst, const void* idx, int idx_bytes, int64_t n_rows, int64_t row_bytes,
                     int64_t src_stride_bytes, int64_t dst_stride_bytes, int64_t bound, int scatter,
                     cudaStream_t stream);
/* dst[idx[i],:] += w[i] * src[i,:] (w may be NULL); strides in elements; atomic != 0 when indices may repeat. */
int dsb200_scatter_add_rows(const void* src, void* dst, const void* idx, int idx_bytes, const float* w, int64_t n_rows,
                            int cols, int64_t src_stride, int64_t dst_stride, int64_t bound, int atomic, int dt,
                            cudaStream_t stream);
/* Ascending sort of every row of data[n_rows, n] (n <= 8192). Reference: csrc/random_ltd/token_sort.cu. */
int dsb200_sort_rows_i32(int* data, int64_t n_rows, int n, cudaStream_t stream);

/* ---------------------------------------------------------------- tcgen05 GEMM --------------------------- */
/* C[M,N] = A[M,K] * B[N,K]^T, bf16 in/out, fp32 accumulation in TMEM (TMA + tcgen05.mma, persistent).
 * force_bn in {0 (auto), 128, 192, 256}. Reference: cuBLAS calls (csrc/transformer/cublas_wrappers.cu). */
int dsb200_gemm_bf16_nt(const void* A, const void* B, void* C, int M, int N, int K, int64_t lda, int64_t ldb,
                        int64_t ldc, int max_ctas, int force_bn, cudaStream_t stream);

/* ---------------------------------------------------------------- quantization --------------------------- */
/* Block-scaled FP8 (fmt 0 = e4m3, 1 = e5m2), one fp32 scale per `block` (multiple of 256) elements.
 * Reference: csrc/quantization/quantize.cu:23 (int8 groups) - this is the native ZeRO++ wire format here. */
int dsb200_fp8_block_quant(const void* x, int dt, void* q, float* scales, int64_t n, int block, int fmt,
                           cudaStream_t stream);
int dsb200_fp8_block_dequant(const void* q, const float* scales, void* y, int dt, int64_t n, int block, int fmt,
                             cudaStream_t stream);
/* Symmetric int8/int4 group quantization; swz_nodes*swz_devs > 0 re-orders partitions [node,dev]->[dev,node].
 * Reference: csrc/quantization/quantize.cu, swizzled_quantize.cu:22-83, dequantize.cu:12. */
int dsb200_int_group_quant(const void* x, int dt, void* q, float* scales, int64_t n, int group, int bits, int swz_nodes,
                           int swz_devs, cudaStream_t stream);
int dsb200_int_group_dequant(const void* q, const float* scales, void* y, int dt, int64_t n, int group, int bits,
                             cudaStream_t stream);
/* out = post * sum_c dequant(q[c]) over nchunks received chunks. Reference: csrc/quantization/quant_reduce.cu:21-133. */
int dsb200_dequant_reduce(const void* q, const float* scales, void* out, int dt, int64_t chunk_elems, int nchunks,
                          int group, int bits, float post, cudaStream_t stream);
/* Group-scaled minifloat rounding (exp_bits/man_bits), RNE or stochastic. Reference: csrc/fp_quantizer. */
int dsb200_minifloat_quant(const void* x, void* y, float* scales, int dt, int64_t n, int group, int exp_bits,
                           int man_bits, float fmt_max, int stochastic, unsigned int seed, cudaStream_t stream);

/* ---------------------------------------------------------------- NVLink collectives --------------------- */
/* peer_* : host arrays of 16 device addresses (one per rank, unused = 0) of symmetric buffers / flag buffers.
 * Epoch flags: channel ch_ready is signalled at kernel start, ch_done by the last CTA. See csrc/symm.cu.
 * Reference: coalesced_collectives.py:158-218, partition_parameters.py:1301-1343, sequence/layer.py:264,
 * csrc/compile/z3.cpp:95-113 (barrier + per-peer copy). */
int dsb200_symm_reduce_scatter(const uint64_t* peer_bufs, const uint64_t* peer_flags, uint64_t mc_ptr, int rank,
                               int world, int64_t shard_numel, int in_dt, void* out, int out_dt, int accumulate,
                               float scale, float* sumsq_partials, unsigned int* ticket, int ch_ready, int ch_done,
                               unsigned int epoch, int ctas, cudaStream_t stream);
int dsb200_symm_all_gather(const uint64_t* peer_shards, const uint64_t* peer_flags, uint64_t mc_full, int rank,
                           int world, int64_t shard_bytes, const void* my_shard, void* full, unsigned int* ticket,
                           int ch_ready, int ch_done, unsigned int epoch, unsigned int done_epoch, int ctas,
                           cudaStream_t stream);
/* All-gather on the copy engines (cudaMemcpyAsync pulls over NVLink on `n_helpers` <= 8 helper streams forked from /
 * joined into `stream`); occupies no SM. The caller orders it against the writers of the shards with
 * dsb200_symm_signal / dsb200_symm_wait. Reference: csrc/compile/z3.cpp:100-112 (symmetric-memory all-gather). */
int dsb200_symm_all_gather_ce(const uint64_t* peer_shards, int rank, int world, int64_t shard_bytes, void* full,
                              const cudaStream_t* helpers, int n_helpers, cudaStream_t stream);
int dsb200_symm_all_to_all(const uint64_t* peer_ins, const uint64_t* peer_flags, int rank, int world, int64_t rows,
                           int64_t row_bytes, int interleave, void* out, unsigned int* ticket, int ch_ready,
                           int ch_done, unsigned int epoch, int ctas, cudaStream_t stream);
int dsb200_symm_wait(const uint64_t* peer_flags, int rank, int world, int channel, unsigned int epoch,
                     cudaStream_t stream);
int dsb200_symm_signal(const uint64_t* peer_flags, int rank, int world, int channel, unsigned int epoch,
                       cudaStream_t stream);

/* ---------------------------------------------------------------- host library (libdsb200_cpu.so) -------- */
/* Reference: csrc/adam/cpu_adam_impl.cpp:35-357, csrc/adagrad/cpu_adagrad.cpp, csrc/lion/cpu_lion_impl.cpp. */
int dsb200_cpu_num_threads(void);
void dsb200_cpu_adam(float* p, const void* g, int g_dt, float* m, float* v, void* lp, int lp_dt, int64_t n, float lr,
                     float beta1, float beta2, float eps, float wd, int step, int bias_correction, int adamw,
                     float grad_scale);
void dsb200_cpu_adam_rollback(float* p, const void* g, int g_dt, float* m, float* v, int64_t n, float lr, float beta1,
                              float beta2, float eps, float wd, int step, int bias_correction, int adamw,
                              float grad_scale);
void dsb200_cpu_adagrad(float* p, const void* g, int g_dt, float* v, void* lp, int lp_dt, int64_t n, float lr,
                        float eps, float wd, float grad_scale);
void dsb200_cpu_lion(float* p, const void* g, int g_dt, float* m, void* lp, int lp_dt, int64_t n, float lr,
                     float beta1, float beta2, float wd, float grad_scale);
void dsb200_cpu_cast(const float* src, void* dst, int dst_dt, int64_t n);
double dsb200_cpu_sumsq(const float* x, int64_t n);
/* Reference: csrc/aio/py_lib/py_ds_aio.cpp:17-158 (aio_handle), csrc/aio/common/deepspeed_aio_common.cpp. */
void* dsb200_aio_create(int64_t block_size, int queue_depth, int threads, int use_direct);
void dsb200_aio_destroy(void* h);
int dsb200_aio_pread(void* h, void* buf, int64_t nbytes, const char* path, int64_t offset, int async);
int dsb200_aio_pwrite(void* h, const void* buf, int64_t nbytes, const char* path, int64_t offset, int async);
int64_t dsb200_aio_wait(void* h);
void* dsb200_aligned_alloc(int64_t nbytes, int64_t alignment);
void dsb200_aligned_free(void* p);

/* ---- flash attention on tcgen05 (csrc/attention_fwd.cu): bf16, head_dim 128, layout [batch, seq, heads, 128] with
 * element strides; lse [B, H, Sq] fp32 or null. Reference call sites: sequence/fpdt_layer.py:238-269 (flash-attn with
 * LSE), runtime/sequence_parallel/ulysses_sp.py:50. ---- */
int dsb200_flash_attn_fwd(const void* q, const void* k, const void* v, void* o, float* lse, int B, int Sq, int Skv,
                          int H, int Hkv, int64_t q_row_stride, int64_t q_batch_stride, int64_t k_row_stride,
                          int64_t k_batch_stride, int64_t v_row_stride, int64_t v_batch_stride, int64_t o_row_stride,
                          int64_t o_batch_stride, float scale, int causal, cudaStream_t stream);
/* ragged prefill over a blocked KV cache [num_blocks, 2, 128, Hkv, 128] (reference: ragged_ops/blocked_flash) */
/* Non-causal attention with additive score biases (Evoformer): score = scale * q.k + bias1[b, j] +
 * bias2[b / bias2_div, h, i, j]; biases bf16, either may be null, pointers / strides 16-byte aligned. */
int dsb200_flash_attn_fwd_bias(const void* q, const void* k, const void* v, void* o, float* lse, const void* bias1,
                               const void* bias2, int B, int Sq, int Skv, int H, int64_t q_row_stride,
                               int64_t q_batch_stride, int64_t k_row_stride, int64_t k_batch_stride,
                               int64_t v_row_stride, int64_t v_batch_stride, int64_t o_row_stride,
                               int64_t o_batch_stride, int64_t bias1_bstride, int64_t bias2_bstride,
                               int64_t bias2_hstride, int64_t bias2_rstride, int bias2_div, float scale,
                               cudaStream_t stream);
int dsb200_flash_attn_fwd_paged(const void* q, const void* kv, void* o, float* lse, const int* cu_q, const int* q_len,
                                const int* kv_len, const int* block_table, int n_seqs, int T_total, int max_q_len,
                                int H, int Hkv, int num_blocks, int max_blocks, int64_t q_row_stride,
                                int64_t o_row_stride, float scale, cudaStream_t stream);
/* backward (csrc/attention_bwd.cu): delta preprocess + transposed dK/dV kernel + dQ kernel, Sq == Skv multiple of 128;
 * dlse = gradient w.r.t. the LSE output or null; delta = fp32 scratch [B, H, S] */
int dsb200_flash_attn_bwd(const void* q, const void* k, const void* v, const void* o, const void* dout, const float* lse,
                          const float* dlse, float* delta, void* dq, void* dk, void* dv, int B, int S, int H, int Hkv,
                          int64_t q_row, int64_t q_batch, int64_t k_row, int64_t k_batch, int64_t v_row, int64_t v_batch,
                          int64_t o_row, int64_t o_batch, int64_t do_row, int64_t do_batch, int64_t dq_row,
                          int64_t dq_batch, int64_t dk_row, int64_t dk_batch, int64_t dv_row, int64_t dv_batch,
                          float scale, int causal, cudaStream_t stream);
/* decode (one token per sequence) over the blocked cache on the same kernel: query heads of a KV head are the tile rows,
 * split-KV partials merged by a second kernel (reference: ragged_ops/blocked_flash decode path) */
int dsb200_flash_decode_paged(const void* q, const void* kv, void* o, const int* kv_len, const int* block_table,
                              float* part_o, float* part_lse, int T, int Hq, int Hkv, int num_blocks, int max_blocks,
                              int splits, float scale, cudaStream_t stream);
/* ---- FastGen ragged ops (csrc/ragged_ops.cu); reference: inference/v2/kernels/ragged_ops/{linear_blocked_kv_rotary,
 * blocked_flash,atom_builder} ---- */
int dsb200_ragged_kv_append_rope(void* q, const void* k, const void* v, void* kv, const float* cos_tab,
                                 const float* sin_tab, const int* tok_seq, const int* tok_pos, const int* block_table,
                                 int T, int64_t q_stride, int64_t k_stride, int64_t v_stride, int Hq, int Hkv, int D,
                                 int rotary_dim, int block_size, int max_blocks, int dt, cudaStream_t stream);
int dsb200_ragged_paged_attention(const void* q, const void* kv, void* out, const int* tok_seq, const int* tok_pos,
                                  const int* block_table, const int* tok_row, float* part_acc, float* part_ml, int T,
                                  int64_t q_stride, int Hq, int Hkv, int D, int block_size, int max_blocks, int splits,
                                  float scale, cudaStream_t stream);

/* ---- weight-only quantized GEMM for decode (csrc/wo_gemm.cu): y[T, N] = x[T, K] * dequant(Wq[N, K])^T with Wq as
 * e4m3 bytes (bits 8) or e2m1 nibble pairs (bits 4) + fp32 scales per `group` elements of K ([N, K / group]).
 * Reference: inference/v2/kernels/core_ops/cuda_linear (FP6-LLM), cutlass_ops/mixed_gemm, csrc/fp_quantizer. ---- */
/* tcgen05 version, T <= 64, K % 64 == 0, group % 64 == 0; x16: scratch of T*K halves; y32: T*N floats (split-K) */
int dsb200_wo_gemm_tc(const void* x, const void* wq, const float* scales, const void* bias, void* y, float* y32, void* x16,
                      int T, int N, int K, int group, int bits, int64_t x_stride, int64_t y_stride, cudaStream_t stream);
/* SIMT reference version, M <= 8, K % 256 == 0, group % 16 == 0; y32: M*N floats */
int dsb200_wo_gemv(const void* x, const void* wq, const float* scales, const void* bias, void* y, float* y32, int M,
                   int N, int K, int group, int bits, int64_t x_stride, int64_t y_stride, cudaStream_t stream);
/* ---- expert-parallel MoE dispatch / combine over NVLink peer memory (csrc/moe_comm.cu); reference:
 * module_inject/auto_ep_layer.py:184-350,575-730 (_AllToAllV + permute / unpermute + weighted combine). segs: device
 * array of nseg {src_rank, src_start, dst_start, n} int32 records sorted by dst_start. ---- */
int dsb200_moe_dispatch(const uint64_t* peer_x, const uint64_t* peer_idx, const uint64_t* peer_scale,
                        const uint64_t* peer_flags, int rank, int world, const void* segs, int nseg, int n_rows,
                        int hidden, void* out, unsigned int* ticket, int ch_ready, int ch_done, unsigned int epoch,
                        int ctas, cudaStream_t stream);
int dsb200_moe_combine(const uint64_t* peer_y, const uint64_t* peer_flags, int rank, int world, const int* slot_rank,
                       const int* slot_row, const float* weights, int T, int k, int hidden, void* out, void* rows_out,
                       unsigned int* ticket, int ch_ready, int ch_done, unsigned int epoch, int ctas,
                       cudaStream_t stream);

/* ---- libdsb200_cpu.so: FastGen host runtime (csrc/cpu/ragged_host.cpp): blocked-KV allocator, sequence table and
 * the per-forward batch descriptor builder; reference: inference/v2/ragged/{blocked_allocator,manager,ragged_wrapper}.py,
 * ragged/csrc/fast_host_buffer.cu, kernels/ragged_ops/atom_builder ---- */
void* dsb200_ragged_create(int num_blocks, int block_size, int max_tracked_sequences);
void dsb200_ragged_destroy(void* h);
int dsb200_ragged_free_blocks(void* h);
int dsb200_ragged_tracked(void* h);
int64_t dsb200_ragged_seen(void* h, int64_t uid);
int dsb200_ragged_blocks_needed(void* h, int64_t uid, int new_tokens);
int dsb200_ragged_blocks(void* h, int64_t uid, int* out, int cap);
int dsb200_ragged_flush(void* h, int64_t uid);
/* buf layout and header fields: see ragged_host.cpp; 0 ok, -1 out of KV blocks, -2 too many sequences, -3 buffer small */
int dsb200_ragged_schedule(void* h, int n_seq, const int64_t* uids, const int* n_tokens, int prefill_min, int* buf,
                           int64_t cap, int* header, int commit);

/* ---- page-locked host allocations (csrc/cpu/pin_memory.cpp; libdsb200_cpu) --------------------------------------
 * page-aligned + mlock; *locked_out = 0 when RLIMIT_MEMLOCK refused the lock (memory still usable). */
void* dsb200_locked_alloc(int64_t bytes, int* locked_out);
int dsb200_locked_free(void* p);
int64_t dsb200_locked_bytes(void);

/* ---- shared-memory all-reduce between the ranks of one host (csrc/cpu/shm_comm.cpp; libdsb200_cpu) ---------------
 * create: rank 0 passes create != 0 first, the others attach afterwards (order them with a barrier); buf_bytes is the
 * chunk capacity. all_reduce: in-place sum, dtype 0 fp32 / 1 bf16 / 2 fp16; 0 ok, -1 bad argument, -2 peer timeout. */
void* dsb200_shm_create(const char* name, int rank, int world, int64_t buf_bytes, int create, double timeout_s);
void dsb200_shm_destroy(void* h);
int dsb200_shm_all_reduce(void* h, void* data, int64_t numel, int dtype);

/* ---- ZeRO-3 unit-timeline planner (csrc/cpu/dc_planner.cpp; libdsb200_cpu) -----------------------------------
 * fit_comm: least squares of us = latency + bytes / bandwidth; out[0] latency_us, out[1] bytes_per_us.
 * plan: see the comment above dsb200_dc_plan in dc_planner.cpp for every argument; outputs persist_out[n_units],
 * launch_step[n_steps], keep_after[n_steps], sim[8]. Returns 0, -1 empty input, -2 step_unit out of range. */
int dsb200_dc_fit_comm(int n, const int64_t* bytes, const double* us, double* out);
int dsb200_dc_plan(int n_units, const int64_t* unit_bytes, const uint8_t* persistent_in, int n_steps,
                   const int* step_unit, const double* step_us, const int64_t* step_peak, double latency_us,
                   double bytes_per_us, int64_t mem_budget, int64_t max_inflight, int64_t persist_headroom,
                   double reuse_window_us, double safety, uint8_t* persist_out, int* launch_step, int* keep_after,
                   double* sim);

/* LoCo error-feedback group quantization: q = Q(x + beta * err), err <- (x + beta * err) - DQ(q); err fp32 [n].
 * Reference: csrc/quantization/swizzled_quantize.cu:200-316. */
int dsb200_loco_group_quant(const void* x, int dt, float* err, void* q, float* scales, int64_t n, int group, int bits,
                            float beta, int swz_nodes, int swz_devs, cudaStream_t stream);
/* MoQ in-place fake quantization (sym / asym, nearest or stochastic rounding). Reference: fake_quantizer.cu. */
int dsb200_fake_quant(void* x, int dt, int64_t n, int group, int bits, int symmetric, int stochastic, unsigned int seed,
                      cudaStream_t stream);

/* FastGen / inference row kernels (csrc/fastgen_ops.cu). Reference: inference/v2/kernels/ragged_ops/{embed,logits_gather,
 * moe_scatter,moe_gather,top_k_gating}, core_ops/{bias_activations,gated_activations}. */
int dsb200_embed_rows(const void* table, const int64_t* ids, const void* pos_table, const int64_t* pos, int pos_offset,
                      void* out, int64_t rows, int cols, int dt, cudaStream_t stream);
int dsb200_bias_act(const void* x, const void* bias, void* y, int64_t rows, int cols, int act, int dt,
                    cudaStream_t stream);
int dsb200_gated_act(const void* x, const void* bias, void* y, int64_t rows, int inter, int act, int dt,
                     cudaStream_t stream);
int dsb200_topk_gating(const float* logits, int T, int E, int k, int renorm, float* scores, int* experts, int* counts,
                       int* slot, cudaStream_t stream);
int dsb200_moe_scatter(const void* x, const int* dst, int T, int k, void* out, int cols, int dt, cudaStream_t stream);
int dsb200_moe_gather(const void* expert_out, const int* src, const float* w, int T, int k, void* out, int cols, int dt,
                      cudaStream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* DSB200_H */

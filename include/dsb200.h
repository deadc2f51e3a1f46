/* dsb200.h -- C ABI of libdsb200.so (CUDA, sm_100a only) and libdsb200_host.so (CPU).
 *
 * Every function returns 0 on success; on failure dsb200_last_error() describes what went wrong (thread-local).
 * Pointers are device pointers unless stated otherwise; `stream` is a cudaStream_t.  dtype codes: 0 = bf16,
 * 1 = fp16, 2 = fp32.  Reference citations name the file of delock/DeepSpeedSYCLSupport whose role the entry takes.
 */
#ifndef DSB200_H_
#define DSB200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct CUstream_st* dsb200_stream_t;

const char* dsb200_last_error(void);
int dsb200_version(void);

/* ---- tcgen05 GEMM (reference: csrc/transformer/cublas_wrappers.cu:62,153; inference/v2/kernels/core_ops/blas_kernels/blas.h:35-138)
 * C[M,N] (=|+=) A*B, bf16 inputs, fp32 accumulate.  a_mn_major=0: A is [M][K]; 1: A stored [K][M].  b_mn_major=0: B is
 * [N][K]; 1: B stored [K][N].  out_mode 0 bf16 store, 1 fp32 store, 2 fp32 accumulate, 3 bf16 accumulate.
 * max_ctas caps the persistent grid (0 = all SMs). */
int dsb200_gemm_bf16(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb, int ldc, int a_mn_major,
                     int b_mn_major, int out_mode, int max_ctas, dsb200_stream_t stream);

/* ---- fused optimizers on flat fp32 shards (reference: csrc/adam/multi_tensor_adam.cu, csrc/lion/multi_tensor_lion.cu,
 * csrc/lamb/fused_lamb_cuda_kernel.cu, csrc/common/custom_cuda_kernel.cu).  grad_type 0 fp32 / 1 bf16 / 2 fp16;
 * lp = optional 16-bit parameter copy written in the same pass (lp_type 0 none / 1 bf16 / 2 fp16);
 * inv_scale, skip: optional DEVICE scalars (gradient multiplier; non-zero skip leaves all buffers untouched). */
int dsb200_fused_adam_flat(float* p, float* m, float* v, const void* g, int grad_type, void* lp, int lp_type, long long n,
                           float lr, float beta1, float beta2, float eps, float weight_decay, int adamw_mode,
                           int bias_correction, int step, const float* inv_scale, const int* skip, dsb200_stream_t stream);
int dsb200_fused_lion_flat(float* p, float* m, const void* g, int grad_type, void* lp, int lp_type, long long n, float lr,
                           float beta1, float beta2, float weight_decay, const float* inv_scale, const int* skip,
                           dsb200_stream_t stream);
int dsb200_fused_adagrad_flat(float* p, float* h, const void* g, int grad_type, void* lp, int lp_type, long long n, float lr,
                              float eps, float weight_decay, const float* inv_scale, const int* skip, dsb200_stream_t stream);
int dsb200_fused_lamb_scratch_floats(long long n);
int dsb200_fused_lamb_flat(float* p, float* m, float* v, const void* g, int grad_type, void* lp, int lp_type, float* u,
                           float* scratch, long long n, float lr, float beta1, float beta2, float eps, float weight_decay,
                           int bias_correction, int step, float max_coeff, float min_coeff, const float* inv_scale,
                           const int* skip, float* coeff_out, dsb200_stream_t stream);
/* list-of-tensors Adam: p/m/v/g/n are HOST arrays (of device pointers / element counts) */
int dsb200_multi_tensor_adam(int ntensors, void** p, void** m, void** v, const void** g, const long long* n, int grad_type,
                             float lr, float beta1, float beta2, float eps, float weight_decay, int adamw_mode,
                             int bias_correction, int step, const float* inv_scale, const int* skip, dsb200_stream_t stream);

/* ---- RMSNorm / LayerNorm (reference: inference/v2/kernels/core_ops/cuda_rms_norm/rms_norm.cu, cuda_layer_norm/layer_norm.cu,
 * csrc/transformer/normalize_kernels.cu).  residual/res_out/beta/mean/rstd/dx_add/dbeta may be NULL. */
int dsb200_norm_fwd(const void* x, const void* residual, const void* gamma, const void* beta, void* y, void* res_out,
                    float* mean, float* rstd, int rows, int hidden, float eps, int layer_norm, int dtype, dsb200_stream_t stream);
long long dsb200_norm_bwd_scratch_floats(int rows, int hidden, int layer_norm);
int dsb200_norm_bwd(const void* dy, const void* x, const void* gamma, const float* mean, const float* rstd, const void* dx_add,
                    void* dx, void* dgamma, void* dbeta, float* scratch, int rows, int hidden, int layer_norm, int dtype,
                    int dparam_f32, int accumulate, dsb200_stream_t stream);

/* RMSNorm backward that allocates nothing.  It has no counterpart in the reference.  For callers that cannot get device
 * memory when they run (a server whose KV-cache pool already owns all free HBM; a latency-critical loop that must avoid
 * the device-wide stall of cudaMalloc/cudaFree): the dgamma partials live in a fixed 4.6 MiB device array inside the
 * library (148 CTAs x hidden <= 8192 floats) that every GPU loading the library holds until exit.  hidden > 8192 is
 * rejected.  The first call on a device looks up the array's address -- make it before any stream capture.  All calls
 * share the array (it keeps the previous call's contents; nothing needs zeroing), so calls overlapping on the GPU race. */
int dsb200_rms_norm_bwd_noalloc(const void* dy, const void* x, const void* gamma, const float* rstd, const void* dx_add,
                                void* dx, void* dgamma, int rows, int hidden, int dtype, int dparam_f32, int accumulate,
                                dsb200_stream_t stream);

/* ---- element-wise transformer ops (reference: csrc/transformer/inference/csrc/apply_rotary_pos_emb.cu,
 * inference/v2/kernels/core_ops/gated_activations, core_ops/bias_activations, csrc/transformer/softmax_kernels.cu) */
int dsb200_rope_inplace(void* x, const int* positions, const float* inv_freq, long long tokens, int row_stride, int head0,
                        int nheads, int head_dim, int seq_len, int backward, int dtype, dsb200_stream_t stream);
int dsb200_gated_act_fwd(const void* in, const void* bias, void* out, long long rows, int F, int act, int interleaved, int dtype,
                         dsb200_stream_t stream);
int dsb200_gated_act_bwd(const void* in, const void* dout, void* din, long long rows, int F, int act, int interleaved, int dtype,
                         dsb200_stream_t stream);
int dsb200_bias_act(const void* in, const void* bias, void* out, long long rows, int cols, int act, int dtype,
                    dsb200_stream_t stream);
int dsb200_cross_entropy_fwd(const void* logits, const long long* labels, float* loss, float* lse, long long rows, int V,
                             long long ld, long long ignore_index, int dtype, dsb200_stream_t stream);
int dsb200_cross_entropy_bwd(void* logits, const long long* labels, const float* lse, const float* gscale, long long rows, int V,
                             long long ld, long long ignore_index, int dtype, dsb200_stream_t stream);
int dsb200_add(const void* a, const void* b, void* out, long long n, int dtype, dsb200_stream_t stream);

/* ---- host library (libdsb200_host.so; reference: csrc/adam/cpu_adam_impl.cpp, csrc/adagrad/cpu_adagrad.cpp,
 * csrc/lion/cpu_lion_impl.cpp).  All pointers are HOST pointers. */
int dsbh_simd_width(void);
int dsbh_adam_step(float* p, float* m, float* v, const void* g, int grad_type, void* lp, int lp_type, int64_t n, float lr,
                   float beta1, float beta2, float eps, float weight_decay, int adamw_mode, int bias_correction, int step,
                   float grad_scale);
int dsbh_adagrad_step(float* p, float* h, const void* g, int grad_type, void* lp, int lp_type, int64_t n, float lr, float eps,
                      float weight_decay, float grad_scale);
int dsbh_lion_step(float* p, float* m, const void* g, int grad_type, void* lp, int lp_type, int64_t n, float lr, float beta1,
                   float beta2, float weight_decay, float grad_scale);
int dsbh_cast_f32_lp(const float* src, void* dst, int lp_type, int64_t n);
double dsbh_sumsq(const void* g, int grad_type, int64_t n);

/* ==== generated by tools/gen_header.py from csrc/ -- do not edit below this line ==== */

/* ---- csrc/attn/evoformer.cu */
/* dtype: 0 bf16, 1 fp16.  q, k, v, o: [B][N][L][H][D] contiguous, D <= 64; bias1 [B][N][L] or NULL; bias2 [B][H][L][L] or NULL (same dtype);
 * lse: [B][N][H][L] fp32 (written). */
int dsb200_evoformer_attn_fwd(const void* q, const void* k, const void* v, const void* bias1, const void* bias2, void* o,
                              float* lse, int B, int N, int L, int H, int D, float scale, int dtype,
                              dsb200_stream_t stream);
/* Gradients: dq, dk, dv as q; dbias1 [B][N][L] / dbias2 [B][H][L][L] fp32, ZERO-INITIALISED by the caller (NULL = not wanted);
 * delta: [B][N][H][L] fp32 scratch. */
int dsb200_evoformer_attn_bwd(const void* q, const void* k, const void* v, const void* bias1, const void* bias2,
                              const void* o, const void* d_o, const float* lse, void* dq, void* dk, void* dv, float* delta,
                              float* dbias1, float* dbias2, int B, int N, int L, int H, int D, float scale, int dtype,
                              dsb200_stream_t stream);

/* ---- csrc/attn/flash_bwd.cu */
/* Gradient of dsb200_flash_attn_fwd.  dout/out: [B*S, hq*D] bf16; qkv, dqkv: packed [B*S, (hq+2*hkv)*D] bf16;
 * lse: [B, hq, S] from the forward.  Scratch supplied by the caller: dq_acc [B*S, hq*D] fp32 ZERO-INITIALISED,
 * delta [B, hq, S] fp32. */
int dsb200_flash_attn_bwd(const void* dout, const void* qkv, const void* out, const float* lse, void* dqkv, float* dq_acc,
                          float* delta, int B, int S, int hq, int hkv, int head_dim, int causal, float scale,
                          dsb200_stream_t stream);

/* ---- csrc/attn/flash_fwd.cu */
/* debug hook: device buffer (>= 8192 int64) receiving clock64 stamps of one CTA of the next forward launches; nullptr = off */
int dsb200_flash_fwd_trace(long long* buf);
/* out[B*S, hq*D] = softmax(Q K^T * scale [+ causal mask]) V with Q,K,V read from packed qkv [B*S, (hq+2*hkv)*D] (bf16).
 * lse: optional [B, hq, S] fp32 (natural-log sum-exp of the scaled scores) for the backward pass. */
int dsb200_flash_attn_fwd(const void* qkv, void* out, float* lse, int B, int S, int hq, int hkv, int head_dim, int causal,
                          float scale, dsb200_stream_t stream);

/* ---- csrc/attn/paged_flash.cu */
/* Bytes of scratch dsb200_paged_flash_attention needs: atom list + count, plus the split-KV partials when n_splits > 1. */
long long dsb200_paged_flash_scratch_bytes(int n_tokens, int hq, int head_dim, int max_atoms, int n_splits);
/* out[t][hq*D] = attention of ragged token t over keys [0, tok_pos[t]] of its sequence, KV read from the blocked cache
 * [num_blocks][2][block_size][hkv][D] (bf16) through the block table.  max_atoms >= number of atoms of the batch
 * (<= n_tokens; ceil(n_tokens / (128/group)) + number of sequences is enough).  n_splits >= 1 splits the KV range of every atom.
 * alibi_slopes: optional [hq] fp32 (device): score += slope[head] * key position; NULL = none.
 * window > 0: local attention, a token sees its last `window` keys (itself included); 0 = the whole history. */
int dsb200_paged_flash_attention(const void* qkv, const void* kv_cache, void* out, const int* tok_seq, const int* tok_pos,
                                 const int* block_table, int n_tokens, int hq, int hkv, int head_dim, int block_size,
                                 int max_blocks_per_seq, int num_blocks, float scale, int max_atoms, int n_splits,
                                 void* scratch, const float* alibi_slopes, int window, dsb200_stream_t stream);
/* Same contract as dsb200_paged_flash_attention for batches of small atoms (decode steps): atoms hold at most 16 / group tokens,
 * keys take the M dimension of the MMAs (see paged_decode_kernel).  head_dim 128, q:kv head ratio dividing 16.
 * max_atoms >= ceil(n_tokens / (16/group)) + number of sequences (<= n_tokens). */
int dsb200_paged_flash_decode(const void* qkv, const void* kv_cache, void* out, const int* tok_seq, const int* tok_pos,
                              const int* block_table, int n_tokens, int hq, int hkv, int head_dim, int block_size,
                              int max_blocks_per_seq, int num_blocks, float scale, int max_atoms, int n_splits,
                              void* scratch, const float* alibi_slopes, int window, dsb200_stream_t stream);

/* ---- csrc/comm/symm.cu */
/* ---- allocation / mapping (host pointers unless noted) */
int dsb200_symm_alloc(long long bytes, void** ptr, unsigned char* handle64);
int dsb200_symm_open(const unsigned char* handle64, void** ptr);
int dsb200_symm_close(void* ptr);
int dsb200_symm_free(void* ptr);
/* peer_ptrs: HOST array of `world` device pointers (this process's mappings of every rank's buffer, own included) */
int dsb200_symm_barrier(void* const* peer_ptrs, long long flag_off, int rank, int world, unsigned int epoch,
                        dsb200_stream_t stream);
/* flag_off: a pad of `world` uint32 used ONLY by device-epoch barriers (zero-initialised); epoch_ctr: local device uint32 */
int dsb200_symm_barrier_dev(void* const* peer_ptrs, long long flag_off, int rank, int world, unsigned int* epoch_ctr,
                            dsb200_stream_t stream);
int dsb200_p2p_allgather(void* const* peer_ptrs, long long src_off, void* dst, long long bytes_per_rank, int rank,
                         int world, int max_ctas, dsb200_stream_t stream);
/* dst[p * bytes_per_chunk ...] = chunk `rank` of peer p's buffer at src_off (equal-split all_to_all_single); dst is local memory. */
int dsb200_p2p_alltoall(void* const* peer_ptrs, long long src_off, void* dst, long long bytes_per_chunk, int rank,
                        int world, int max_ctas, dsb200_stream_t stream);
/* src dtype / dst dtype: 0 bf16, 2 fp32.  n = elements per rank (multiple of 8).  sumsq: optional device float (+= sum g^2). */
int dsb200_p2p_reduce_scatter(void* const* peer_ptrs, long long src_off, int src_dtype, void* dst, int dst_dtype,
                              long long n, int rank, int world, float scale, int accumulate, float* sumsq, int max_ctas,
                              dsb200_stream_t stream);
/* Range form: Adam on elements [elem_lo, elem_lo + n) of this rank's slice (slice length n_per_rank) of the W*n_per_rank
 * bf16 gradient buffer at src_off.  p/m/v/lp/g_out point at element elem_lo of the local shard.  mc_base != NULL: the sum
 * comes from the NVSwitch (multimem.ld_reduce); NULL: pulled from the peers.  g_out / sumsq optional (may be NULL). */
int dsb200_p2p_rs_adam_range(void* const* peer_ptrs, void* mc_base, long long src_off, long long n_per_rank,
                             long long elem_lo, long long n, float* p, float* m, float* v, void* lp, void* g_out,
                             float* sumsq, int rank, int world, float scale, float lr, float beta1, float beta2, float eps,
                             float weight_decay, int adamw_mode, int bias_correction, int step, int max_ctas,
                             dsb200_stream_t stream);
int dsb200_p2p_rs_adam(void* const* peer_ptrs, long long src_off, float* p, float* m, float* v, void* lp, long long n,
                       int rank, int world, float scale, float lr, float beta1, float beta2, float eps, float weight_decay,
                       int adamw_mode, int bias_correction, int step, int max_ctas, dsb200_stream_t stream);
int dsb200_p2p_allreduce_bf16(void* const* peer_ptrs, long long src_off, void* dst, long long n, int rank, int world,
                              int max_ctas, dsb200_stream_t stream);
/* All-gather on the copy engines: one peer -> local cudaMemcpyAsync per rank (NVLink DMA, zero SMs), staggered so that at any
 * moment the ranks read from different peers.  dst is local memory; the peers' shards sit at src_off of their symmetric buffers. */
int dsb200_p2p_allgather_ce(void* const* peer_ptrs, long long src_off, void* dst, long long bytes_per_rank, int rank,
                            int world, dsb200_stream_t stream);
/* ---- NVLS entry points.  `mc_base` = multicast mapping of the symmetric buffer (every rank bound), DEVICE pointer.
 * src dtype / dst dtype: 0 bf16, 2 fp32.  n = elements per rank (multiple of 8). */
int dsb200_nvls_reduce_scatter(void* mc_base, long long src_off, int src_dtype, void* dst, int dst_dtype, long long n,
                               int rank, int world, float scale, int accumulate, float* sumsq, int max_ctas,
                               dsb200_stream_t stream);
/* In-place bf16 all-reduce of n elements at `off` (n % (8*world) == 0).  The caller brackets it with dsb200_symm_barrier. */
int dsb200_nvls_allreduce_bf16(void* mc_base, long long off, long long n, int rank, int world, int max_ctas,
                               dsb200_stream_t stream);
int dsb200_nvls_allgather(void* mc_base, long long dst_off, const void* src, long long bytes, int rank, int max_ctas,
                          dsb200_stream_t stream);
/* qgZ consumer: q_off / s_off = byte offsets of the fp8 payload (W*n bytes) and the fp32 block scales (W*n/128 floats) in
 * every rank's symmetric buffer; n = elements per rank (multiple of 128); dst dtype 0 bf16 / 2 fp32. */
int dsb200_p2p_fp8_reduce_scatter(void* const* peer_ptrs, long long q_off, long long s_off, void* dst, int dst_dtype,
                                  long long n, int rank, int world, float scale, int accumulate, float* sumsq,
                                  int max_ctas, dsb200_stream_t stream);
/* ---- MoE expert-parallel dispatch / combine over peer memory.  buf_off: byte offset of the [W][n_local][capacity][M] bf16
 * receive buffer inside every rank's symmetric buffer; token_expert / token_slot: [S][K] int32 (expert < 0 = dropped token). */
int dsb200_moe_dispatch_push(void* const* peer_ptrs, long long buf_off, const void* x, const int* token_expert,
                             const int* token_slot, const float* row_scale, int S, int K, int M, int rank, int world,
                             int n_local, int capacity, dsb200_stream_t stream);
/* out [S][M] bf16 (may be NULL when only dot_out is wanted); weights [S][K] fp32 or NULL; dot_with [S][M] bf16 + dot_out [S][K]
 * fp32 (zero-initialised by the caller) optional. */
int dsb200_moe_combine_pull(void* const* peer_ptrs, long long buf_off, void* out, const int* token_expert,
                            const int* token_slot, const float* weights, const void* dot_with, float* dot_out, int S,
                            int K, int M, int rank, int world, int n_local, int capacity, dsb200_stream_t stream);
/* Ulysses head <-> sequence exchange over peer memory.  `local`: [B * S_local][src_ld] bf16 (source layout); the dest-layout buffer
 * [B * S_local * world][dst_ld] lives at buf_off of every rank's symmetric buffer.  segs: n_seg x {src_col, dst_col, heads,
 * heads_per_rank} (ints, host).  push = 1: local -> owners' buffers; push = 0: owners' buffers -> local. */
int dsb200_ulysses_heads_exchange(void* const* peer_ptrs, long long buf_off, void* local, int B, int S_local, int rank,
                                  int world, int D, int src_ld, int dst_ld, const int* segs, int n_seg, int push,
                                  int max_ctas, dsb200_stream_t stream);

/* ---- csrc/elementwise/elementwise.cu */
/* out[0] += sum(x[i]^2) over n elements (dtype 0 bf16 / 1 fp16 / 2 fp32).  x must be 16-byte aligned. */
int dsb200_sumsq(const void* x, long long n, int dtype, float* out, dsb200_stream_t stream);

/* ---- csrc/gemm/gemm_ag.cu */
/* C[M,N] (bf16) = A[M,K] * W^T with W[N,K] gathered from peers INSIDE the kernel.
 *   b_local     : [N,K] bf16 destination of the gather (kept for backward); also the GEMM's B operand
 *   peer_shards : HOST array of `world` device pointers -- each rank's shard region of the unit (this process's mappings)
 *   w_off       : flat element offset of W[0][0] inside the unit; shard_numel: elements per rank shard
 *   flags       : device scratch, >= ceil(N/256) + 1 ints (zeroed by this call): per n-block arrival counters + one work-claim counter */
int dsb200_gemm_ag_bf16(const void* A, void* b_local, void* C, int M, int N, int K, int lda, int ldc,
                        void* const* peer_shards, long long w_off, long long shard_numel, int* flags, int world,
                        dsb200_stream_t stream);

/* ---- csrc/gemm/gemm_ar.cu */
int dsb200_gemm_ar_bf16(const void* A, const void* B, void* const* peer_ptrs, long long part_off, long long flag_off,
                        unsigned int epoch, void* out, const void* residual, int M, int N, int K, int lda, int ldb,
                        int ldo, int ldr, int rank, int world, int max_ctas, dsb200_stream_t stream);
/* GEMM -> reduce-scatter over rows (sequence-parallel tensor parallelism): same kernel, but this rank reduces and writes only rows
 * [row_lo, row_hi) of the sum -- out and residual are [row_hi - row_lo][N].  Every rank still publishes its whole partial (the
 * owners of the other rows need it); the row ranges of the ranks must cover [0, M) and every rank must own at least one row. */
int dsb200_gemm_rs_bf16(const void* A, const void* B, void* const* peer_ptrs, long long part_off, long long flag_off,
                        unsigned int epoch, void* out, const void* residual, int M, int N, int K, int lda, int ldb,
                        int ldo, int ldr, int rank, int world, int row_lo, int row_hi, int max_ctas,
                        dsb200_stream_t stream);
/* Graph-capturable forms: the epoch lives in device memory (`epoch_ctr`, one uint32 of LOCAL memory per rank, all ranks start it at
 * the same value); every call bumps it with a one-thread kernel in front of the GEMM, and the GEMM picks the partial area
 * part_off0 / part_off1 by the bumped value's parity.  A captured call therefore replays correctly any number of times.
 * row_lo = 0, row_hi = M: all-reduce; a sub-range: reduce-scatter over rows. */
int dsb200_gemm_ar_bf16_dev(const void* A, const void* B, void* const* peer_ptrs, long long part_off0, long long part_off1,
                            long long flag_off, unsigned int* epoch_ctr, void* out, const void* residual, int M, int N,
                            int K, int lda, int ldb, int ldo, int ldr, int rank, int world, int row_lo, int row_hi,
                            int max_ctas, dsb200_stream_t stream);

/* ---- csrc/gemm/gemm_fp8.cu */
/* C[M,N] (bf16) = diag(sa) * (A * B^T) * diag(sb) with A [M][K], B [N][K] in e4m3 (one byte per element, K contiguous).
 * sa [M] / sb [N] fp32 scales, either may be null (= 1).  lda / ldb in elements (= bytes), multiples of 16; ldc multiple of 8. */
int dsb200_gemm_fp8(const void* A, const void* B, void* C, const float* sa, const float* sb, int M, int N, int K, int lda,
                    int ldb, int ldc, int max_ctas, dsb200_stream_t stream);

/* ---- csrc/gemm/gemm_fp8_block.cu */
/* C[M,N] (bf16) = sum over 128-wide K blocks of sa[m][kb] * sb[n][kb] * (A_kb * B_kb^T); A [M][K], B [N][K] e4m3 (K contiguous,
 * K % 128 == 0), sa [M][K/128], sb [N][K/128] fp32 (the layout dsb200_fp8_block_quantize writes for row-major operands). */
int dsb200_gemm_fp8_block(const void* A, const void* B, void* C, const float* sa, const float* sb, int M, int N, int K,
                          int lda, int ldb, int ldc, int max_ctas, dsb200_stream_t stream);

/* ---- csrc/gemm/gemm_grouped.cu */
/* C[r, :] = A[r, :] * W[expert(r)]^T for rows sorted by expert; `cum` (device, int64[E]) holds inclusive cumulative row counts.
 * Rows at or beyond cum[E-1] are left untouched.  N must be a multiple of 8; A, W, C 16-byte aligned with 16-byte row strides. */
int dsb200_moe_gemm_bf16(const void* A, const void* W, void* C, const long long* cum, int R, int N, int K, int E, int lda,
                         int ldw, int ldc, dsb200_stream_t stream);

/* ---- csrc/gemm/gemm_mxfp8.cu */
/* Bytes of the packed scale tensor for a [rows][K] operand (rows padded to a multiple of 256 so either operand role works). */
long long dsb200_mxfp8_sf_bytes(int rows, int K);
/* x [rows][K] (row stride ldx elements; dtype 0 bf16, 1 fp16, 2 fp32; K % 128 == 0) -> q [rows][K] e4m3 + sf (dsb200_mxfp8_sf_bytes,
 * ZERO-INITIALISED by the caller: rows beyond `rows` keep scale byte 0). */
int dsb200_mxfp8_quantize(const void* x, void* q, void* sf, int rows, int K, long long ldx, int dtype,
                          dsb200_stream_t stream);
/* C[M,N] (bf16) = dequant(A) * dequant(B)^T with MXFP8 operands (payload + packed UE8M0 scales from dsb200_mxfp8_quantize). */
int dsb200_gemm_mxfp8(const void* A, const void* B, void* C, const void* sfa, const void* sfb, int M, int N, int K,
                      int lda, int ldb, int ldc, int max_ctas, dsb200_stream_t stream);

/* ---- csrc/gemm/gemm_skinny.cu */
/* out[M][N] (bf16, row stride ldo) = x[M][K] w[N][K]^T, bf16 operands, M <= 128.  splits >= 1 cuts K; with splits > 1 the caller
 * passes a workspace of splits * M * N floats and N must be a multiple of 4.  K, ldx, ldw multiples of 8 (16-byte rows for TMA). */
int dsb200_gemm_skinny_bf16(const void* x, const void* w, void* out, void* workspace, int M, int N, int K, int ldx,
                            int ldw, int ldo, int splits, dsb200_stream_t stream);
/* out[M][N] (bf16) = x[M][K] (bf16) * dequant(wq)[N][K]^T, M <= 128.  bits = 8: wq int8 [N][K]; bits = 4: wq uint8 [N][K/2] (low nibble
 * first, value = nibble - 8).  scales fp32 [N][K / group], group a multiple of 64 dividing K; K a multiple of 64.
 * splits > 1 needs a workspace of splits * M * N floats and N, ldo multiples of 4. */
int dsb200_gemm_skinny_mixed(const void* x, const void* wq, const void* scales, void* out, void* workspace, int M, int N,
                             int K, int ldx, int ldo, int bits, int group, int splits, dsb200_stream_t stream);

/* ---- csrc/host/aio.cpp */
void* dsbh_aio_create(int64_t block_size, int queue_depth, int single_submit, int overlap_events, int num_threads);
void dsbh_aio_destroy(void* hp);
int64_t dsbh_aio_block_size(void* hp);
int dsbh_aio_queue_depth(void* hp);
int dsbh_aio_thread_count(void* hp);
/* Returns bytes waited for (>= 0) or -1 if any chunk failed since the last wait. */
int64_t dsbh_aio_wait(void* hp);
int dsbh_aio_pread(void* hp, void* buffer, int64_t nbytes, const char* path, int async, int64_t file_offset);
int dsbh_aio_pwrite(void* hp, const void* buffer, int64_t nbytes, const char* path, int async, int64_t file_offset);

/* ---- csrc/host/ragged_host.cpp */
void* dsbh_allocator_create(int num_blocks);
void dsbh_allocator_destroy(void* h);
int dsbh_allocator_free_blocks(void* h);
/* returns 0 on success, -1 when fewer than n blocks are free (nothing is allocated then) */
int dsbh_allocator_allocate(void* h, int n, int32_t* out);
/* returns 0 on success, -1 on an invalid / double free (allocator unchanged) */
int dsbh_allocator_free(void* h, const int32_t* blocks, int n);
/* seq_meta: per sequence (start token in batch, n new tokens, tokens already in the KV cache). Returns the atom count. */
int dsbh_build_atoms(const int32_t* seq_start, const int32_t* seq_ntok, const int32_t* seq_seen, int n_seqs, int q_block,
                     int kv_block, int32_t* atoms_out, int max_atoms);
/* Pack the per-token metadata of a ragged batch: tok_seq[t], tok_pos[t], last_tok[s]. */
int dsbh_pack_batch(const int32_t* seq_ntok, const int32_t* seq_seen, int n_seqs, int32_t* tok_seq, int32_t* tok_pos,
                    int32_t* last_tok);

/* ---- csrc/host/shm_comm.cpp */
/* Creates (rank 0) or attaches to the segment `/name`.  Returns an opaque handle or nullptr. */
void* dsbh_shm_open(const char* name, int rank, int world, long long chunk_bytes, double timeout_s);
/* In-place sum over all ranks of `numel` elements at `data`.  Returns 0, 1 (bad argument) or 2 (peer timeout). */
int dsbh_shm_allreduce(void* handle, void* data, long long numel, int dtype, double timeout_s);
int dsbh_shm_barrier(void* handle, double timeout_s);
void dsbh_shm_close(void* handle);

/* ---- csrc/host/sparse_lut.cpp */
/* Returns the number of non-zero blocks.  With out_h == nullptr only counts (and fills row_ptr if given). */
long long dsbh_sparse_lut(const int64_t* layout, int H, int M, int N, int64_t* out_h, int64_t* out_r, int64_t* out_c,
                          int64_t* row_ptr);

/* ---- csrc/quant/onebit.cu */
/* x[n] fp32 (read only), err[n] fp32 (error feedback, updated), packed[n/8], scale_io[2] fp32 device scratch:
 * on return scale_io[0] = ||x+err||_2 / sqrt(n).  n must be a multiple of 8; x/err 16-byte aligned. */
int dsb200_onebit_compress(void* x, void* err, void* packed, void* scale_io, long long n, dsb200_stream_t stream);
int dsb200_onebit_decompress(const void* packed, const void* scales, void* out, int rows, long long chunk_bytes, int mode,
                             dsb200_stream_t stream);

/* ---- csrc/quant/quantizer.cu */
/* x: n elements (n % 128 == 0) -> q: n bytes (e4m3), scales: n/128 floats */
int dsb200_fp8_block_quantize(const void* x, void* q, float* scales, long long n, int dtype, dsb200_stream_t stream);
int dsb200_fp8_block_dequantize(const void* q, const float* scales, void* out, long long n, int dtype,
                                dsb200_stream_t stream);
/* q: parts * n bytes, scales: parts * n/128 ; out: n elements = scale_out * sum over parts */
int dsb200_fp8_block_dequant_reduce(const void* q, const float* scales, void* out, long long n, int parts, float scale_out,
                                    int dtype, dsb200_stream_t stream);
/* groups * group_size elements; params: groups*2 floats (scale, offset); bits 8 or 4; q bytes = n * bits / 8 */
int dsb200_int_quantize(const void* x, void* q, float* params, long long groups, int group_size, int bits, int symmetric,
                        int dtype, dsb200_stream_t stream);
int dsb200_int_dequantize(const void* q, const float* params, void* out, long long groups, int group_size, int bits,
                          int symmetric, int dtype, dsb200_stream_t stream);
int dsb200_fake_quantize(void* x, long long groups, int group_size, int bits, int symmetric, int stochastic,
                         long long seed, int dtype, dsb200_stream_t stream);
/* q[rows][cols] (e4m3) and scales[rows] (fp32) with x ~= q * scale per row; x: [rows][cols] with row stride ldx (elements).
 * dtype: 0 bf16, 1 fp16, 2 fp32.  cols must be a multiple of 8. */
int dsb200_fp8_rowwise_quantize(const void* x, void* q, float* scales, int rows, int cols, long long ldx, int dtype,
                                dsb200_stream_t stream);

/* ---- csrc/ragged/ragged_ops.cu */
int dsb200_kv_rotary_append(void* qkv, void* kv_cache, const int* tok_seq, const int* tok_pos, const int* block_table,
                            const float* inv_freq, int n_tokens, int hq, int hkv, int head_dim, int block_size,
                            int max_blocks_per_seq, int rot_dim, dsb200_stream_t stream);
int dsb200_paged_attention(const void* qkv, const void* kv_cache, void* out, const int* tok_seq, const int* tok_pos,
                           const int* block_table, int n_tokens, int hq, int hkv, int head_dim, int block_size,
                           int max_blocks_per_seq, float scale, dsb200_stream_t stream);
int dsb200_ragged_embed(const long long* ids, const void* wte, const void* wpe, const int* tok_pos, void* out,
                        int n_tokens, int dim, int pos_offset, dsb200_stream_t stream);
int dsb200_logits_gather(const void* hidden, const int* last_tok, void* out, int n_seqs, int dim, dsb200_stream_t stream);
/* expert_counts must be zeroed by the caller.  K <= 8, E <= 2048. */
int dsb200_top_k_gating(const float* logits, int* expert_counts, float* scores, int* assignments, int* offsets,
                        int n_tokens, int n_experts, int top_k, int normalize, dsb200_stream_t stream);
int dsb200_moe_scatter(const void* x, const int* expert_counts, const int* assignments, const int* offsets,
                       long long* expert_cumsum, void* moe_in, int* mapped_slots, int n_tokens, int n_experts, int dim,
                       int top_k, dsb200_stream_t stream);
int dsb200_moe_gather(const void* moe_out, const float* scores, const int* mapped_slots, void* out, int n_tokens, int dim,
                      int top_k, dsb200_stream_t stream);

/* ---- csrc/random_ltd/token_ops.cu */
/* idx: int32 [rows, reserved] of unique positions in [0, seq); sorted ascending in place, row by row. */
int dsb200_token_sort(void* idx, int rows, int reserved, int seq, dsb200_stream_t stream);
/* part[B,R,row] (batch_first) or part[R,B,row]  <-  all[B,S,row] / all[S,B,row] at idx[B,R] */
int dsb200_token_gather(const void* all, void* part, const void* idx, int B, int S, int R, long long row_bytes,
                        int batch_first, dsb200_stream_t stream);
int dsb200_token_scatter(void* all, const void* part, const void* idx, int B, int S, int R, long long row_bytes,
                         int batch_first, dsb200_stream_t stream);

/* ---- csrc/spatial/bias_add.cu */
/* dtype: 0 = fp32, 1 = fp16, 2 = bf16.  other / other_bias may be null.  channels * sizeof(T) must be a multiple of 16. */
int dsb200_nhwc_bias_add(const void* act, const void* bias, const void* other, const void* other_bias, void* out,
                         long long total, int channels, int dtype, dsb200_stream_t stream);

/* ==== end of generated section ==== */

#ifdef __cplusplus
}
#endif
#endif /* DSB200_H_ */

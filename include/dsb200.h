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

#ifdef __cplusplus
}
#endif
#endif /* DSB200_H_ */

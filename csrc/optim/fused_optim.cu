// Fused optimizers for sm_100a: Adam/AdamW, Lion, LAMB, Adagrad on FLAT fp32 shards (the layout ZeRO keeps), with
// the low-precision parameter copy written in the same pass, gradient unscale/clip folded in through a device-side
// scale, and an overflow "skip" flag read on the device (no host sync between norm and step).  A multi-tensor entry
// point keeps API parity with list-of-tensors callers.
// Reference counterparts: csrc/adam/multi_tensor_adam.cu:23-166 + multi_tensor_apply.cuh:25-131 (chunked lists,
// <=320 blocks x 512 thr, ILP 4); csrc/lion/multi_tensor_lion.cu:23-126; csrc/lamb/fused_lamb_cuda_kernel.cu:346-460
// (3 launches); csrc/common/custom_cuda_kernel.cu:8-43 (fp32->fp16 copy, here fused).
#include "../common/host.h"
#include "../common/ptx.cuh"

namespace dsb {
namespace optim {

constexpr int THREADS = 512;
constexpr int VEC = 4;

enum GradType { G_F32 = 0, G_BF16 = 1, G_F16 = 2 };
enum LpType { LP_NONE = 0, LP_BF16 = 1, LP_F16 = 2 };

template <int GT>
__device__ __forceinline__ void load_grad4(const void* g, size_t i, float (&out)[4]) {
  if constexpr (GT == G_F32) {
    float4 v = *reinterpret_cast<const float4*>(static_cast<const float*>(g) + i);
    out[0] = v.x; out[1] = v.y; out[2] = v.z; out[3] = v.w;
  } else if constexpr (GT == G_BF16) {
    uint2 raw = *reinterpret_cast<const uint2*>(static_cast<const __nv_bfloat16*>(g) + i);
    __nv_bfloat162 a = *reinterpret_cast<__nv_bfloat162*>(&raw.x);
    __nv_bfloat162 b = *reinterpret_cast<__nv_bfloat162*>(&raw.y);
    float2 fa = __bfloat1622float2(a), fb = __bfloat1622float2(b);
    out[0] = fa.x; out[1] = fa.y; out[2] = fb.x; out[3] = fb.y;
  } else {
    uint2 raw = *reinterpret_cast<const uint2*>(static_cast<const __half*>(g) + i);
    __half2 a = *reinterpret_cast<__half2*>(&raw.x);
    __half2 b = *reinterpret_cast<__half2*>(&raw.y);
    float2 fa = __half22float2(a), fb = __half22float2(b);
    out[0] = fa.x; out[1] = fa.y; out[2] = fb.x; out[3] = fb.y;
  }
}
template <int GT>
__device__ __forceinline__ float load_grad1(const void* g, size_t i) {
  if constexpr (GT == G_F32) return static_cast<const float*>(g)[i];
  if constexpr (GT == G_BF16) return __bfloat162float(static_cast<const __nv_bfloat16*>(g)[i]);
  return __half2float(static_cast<const __half*>(g)[i]);
}
template <int LP>
__device__ __forceinline__ void store_lp4(void* lp, size_t i, const float (&p)[4]) {
  if constexpr (LP == LP_BF16) {
    uint2 raw;
    raw.x = pack_bf16x2(p[0], p[1]);
    raw.y = pack_bf16x2(p[2], p[3]);
    *reinterpret_cast<uint2*>(static_cast<__nv_bfloat16*>(lp) + i) = raw;
  } else if constexpr (LP == LP_F16) {
    __half2 a = __floats2half2_rn(p[0], p[1]), b = __floats2half2_rn(p[2], p[3]);
    uint2 raw;
    raw.x = *reinterpret_cast<uint32_t*>(&a);
    raw.y = *reinterpret_cast<uint32_t*>(&b);
    *reinterpret_cast<uint2*>(static_cast<__half*>(lp) + i) = raw;
  }
}
template <int LP>
__device__ __forceinline__ void store_lp1(void* lp, size_t i, float p) {
  if constexpr (LP == LP_BF16) static_cast<__nv_bfloat16*>(lp)[i] = __float2bfloat16_rn(p);
  if constexpr (LP == LP_F16) static_cast<__half*>(lp)[i] = __float2half_rn(p);
}

struct AdamArgs {
  float lr, beta1, beta2, eps, weight_decay;
  float bc1, bc2_rsqrt;  // 1/(1-b1^t) and 1/sqrt(1-b2^t) (1 when bias correction is off)
  int adamw;             // 1: decoupled decay, 0: L2 (added to the gradient)
  const float* inv_scale;  // device scalar multiplied into every gradient (1/loss_scale * clip coef); may be null
  const int* skip;         // device flag: non-zero -> leave everything untouched (fp16 overflow); may be null
};

__device__ __forceinline__ void adam_math(float& p, float& m, float& v, float g, const AdamArgs& a) {
  if (!a.adamw) g = fmaf(a.weight_decay, p, g);
  m = fmaf(a.beta1, m, (1.f - a.beta1) * g);
  v = fmaf(a.beta2, v, (1.f - a.beta2) * g * g);
  const float denom = sqrtf(v) * a.bc2_rsqrt + a.eps;
  float upd = (m * a.bc1) / denom;
  if (a.adamw) upd = fmaf(a.weight_decay, p, upd);
  p = fmaf(-a.lr, upd, p);
}

template <int GT, int LP>
__global__ void __launch_bounds__(THREADS)
adam_flat_kernel(float* __restrict__ p, float* __restrict__ m, float* __restrict__ v, const void* __restrict__ g,
                 void* __restrict__ lp, size_t n, AdamArgs a) {
  if (a.skip && *a.skip) return;
  const float gs = a.inv_scale ? *a.inv_scale : 1.f;
  const size_t nvec = n / VEC;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < nvec; i += stride) {
    const size_t e = i * VEC;
    float4 p4 = *reinterpret_cast<float4*>(p + e);
    float4 m4 = *reinterpret_cast<float4*>(m + e);
    float4 v4 = *reinterpret_cast<float4*>(v + e);
    float gg[4];
    load_grad4<GT>(g, e, gg);
    float pp[4] = {p4.x, p4.y, p4.z, p4.w}, mm[4] = {m4.x, m4.y, m4.z, m4.w}, vv[4] = {v4.x, v4.y, v4.z, v4.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) adam_math(pp[k], mm[k], vv[k], gg[k] * gs, a);
    *reinterpret_cast<float4*>(p + e) = make_float4(pp[0], pp[1], pp[2], pp[3]);
    *reinterpret_cast<float4*>(m + e) = make_float4(mm[0], mm[1], mm[2], mm[3]);
    *reinterpret_cast<float4*>(v + e) = make_float4(vv[0], vv[1], vv[2], vv[3]);
    store_lp4<LP>(lp, e, pp);
  }
  // tail (n not a multiple of 4)
  for (size_t e = nvec * VEC + static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; e < n; e += stride) {
    float pp = p[e], mm = m[e], vv = v[e];
    adam_math(pp, mm, vv, load_grad1<GT>(g, e) * gs, a);
    p[e] = pp; m[e] = mm; v[e] = vv;
    store_lp1<LP>(lp, e, pp);
  }
}

// ------------------------------------------------------------------ Lion
struct LionArgs {
  float lr, beta1, beta2, weight_decay;
  const float* inv_scale;
  const int* skip;
};
__device__ __forceinline__ void lion_math(float& p, float& m, float g, const LionArgs& a) {
  const float c = fmaf(a.beta1, m, (1.f - a.beta1) * g);
  const float s = (c > 0.f) ? 1.f : ((c < 0.f) ? -1.f : 0.f);
  p = p * (1.f - a.lr * a.weight_decay) - a.lr * s;
  m = fmaf(a.beta2, m, (1.f - a.beta2) * g);
}
template <int GT, int LP>
__global__ void __launch_bounds__(THREADS)
lion_flat_kernel(float* __restrict__ p, float* __restrict__ m, const void* __restrict__ g, void* __restrict__ lp,
                 size_t n, LionArgs a) {
  if (a.skip && *a.skip) return;
  const float gs = a.inv_scale ? *a.inv_scale : 1.f;
  const size_t nvec = n / VEC;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < nvec; i += stride) {
    const size_t e = i * VEC;
    float4 p4 = *reinterpret_cast<float4*>(p + e);
    float4 m4 = *reinterpret_cast<float4*>(m + e);
    float gg[4];
    load_grad4<GT>(g, e, gg);
    float pp[4] = {p4.x, p4.y, p4.z, p4.w}, mm[4] = {m4.x, m4.y, m4.z, m4.w};
#pragma unroll
    for (int k = 0; k < 4; ++k) lion_math(pp[k], mm[k], gg[k] * gs, a);
    *reinterpret_cast<float4*>(p + e) = make_float4(pp[0], pp[1], pp[2], pp[3]);
    *reinterpret_cast<float4*>(m + e) = make_float4(mm[0], mm[1], mm[2], mm[3]);
    store_lp4<LP>(lp, e, pp);
  }
  for (size_t e = nvec * VEC + static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; e < n; e += stride) {
    float pp = p[e], mm = m[e];
    lion_math(pp, mm, load_grad1<GT>(g, e) * gs, a);
    p[e] = pp; m[e] = mm;
    store_lp1<LP>(lp, e, pp);
  }
}

// ------------------------------------------------------------------ Adagrad
template <int GT, int LP>
__global__ void __launch_bounds__(THREADS)
adagrad_flat_kernel(float* __restrict__ p, float* __restrict__ h, const void* __restrict__ g, void* __restrict__ lp,
                    size_t n, float lr, float eps, float weight_decay, const float* inv_scale, const int* skip) {
  if (skip && *skip) return;
  const float gs = inv_scale ? *inv_scale : 1.f;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  for (size_t e = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; e < n; e += stride) {
    float pp = p[e];
    float gg = load_grad1<GT>(g, e) * gs;
    gg = fmaf(weight_decay, pp, gg);
    float hh = fmaf(gg, gg, h[e]);
    pp -= lr * gg / (sqrtf(hh) + eps);
    h[e] = hh;
    p[e] = pp;
    store_lp1<LP>(lp, e, pp);
  }
}

// ------------------------------------------------------------------ LAMB (per-tensor trust ratio)
// Phase 1: m,v update; update u = m_hat/(sqrt(v_hat)+eps) + wd*p written to `u`; per-block partial ||p||^2, ||u||^2.
// Phase 2: every block re-reduces the (<= grid) partials (deterministic order) and applies p -= lr * ratio * u.
// Scratch: 2 * grid floats.
struct LambArgs {
  float lr, beta1, beta2, eps, weight_decay, bc1, bc2, max_coeff, min_coeff;
  const float* inv_scale;
  const int* skip;
};
template <int GT>
__global__ void __launch_bounds__(THREADS)
lamb_phase1_kernel(const float* __restrict__ p, float* __restrict__ m, float* __restrict__ v,
                   const void* __restrict__ g, float* __restrict__ u, size_t n, LambArgs a,
                   float* __restrict__ partial) {
  __shared__ float red[2][THREADS / 32];
  float wn = 0.f, un = 0.f;
  const bool live = !(a.skip && *a.skip);
  if (live) {
    const float gs = a.inv_scale ? *a.inv_scale : 1.f;
    const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
    for (size_t e = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; e < n; e += stride) {
      const float gg = load_grad1<GT>(g, e) * gs;
      const float pp = p[e];
      const float mm = fmaf(a.beta1, m[e], (1.f - a.beta1) * gg);
      const float vv = fmaf(a.beta2, v[e], (1.f - a.beta2) * gg * gg);
      m[e] = mm;
      v[e] = vv;
      const float upd = (mm / a.bc1) / (sqrtf(vv / a.bc2) + a.eps) + a.weight_decay * pp;
      u[e] = upd;
      wn = fmaf(pp, pp, wn);
      un = fmaf(upd, upd, un);
    }
  }
  wn = warp_sum(wn);
  un = warp_sum(un);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  if (l == 0) { red[0][w] = wn; red[1][w] = un; }
  __syncthreads();
  if (w == 0) {
    wn = (l < THREADS / 32) ? red[0][l] : 0.f;
    un = (l < THREADS / 32) ? red[1][l] : 0.f;
    wn = warp_sum(wn);
    un = warp_sum(un);
    if (l == 0) { partial[blockIdx.x] = wn; partial[gridDim.x + blockIdx.x] = un; }
  }
}
template <int LP>
__global__ void __launch_bounds__(THREADS)
lamb_phase2_kernel(float* __restrict__ p, const float* __restrict__ u, void* __restrict__ lp, size_t n, LambArgs a,
                   const float* __restrict__ partial, int nparts, float* __restrict__ coeff_out) {
  if (a.skip && *a.skip) return;
  __shared__ float s_ratio;
  if (threadIdx.x < 32) {
    float wn = 0.f, un = 0.f;
    for (int i = threadIdx.x; i < nparts; i += 32) { wn += partial[i]; un += partial[nparts + i]; }
    wn = warp_sum(wn);
    un = warp_sum(un);
    if (threadIdx.x == 0) {
      const float w_norm = sqrtf(wn), u_norm = sqrtf(un);
      float r = 1.f;
      if (w_norm > 0.f && u_norm > 0.f) r = fminf(fmaxf(w_norm / u_norm, a.min_coeff), a.max_coeff);
      s_ratio = r;
      if (blockIdx.x == 0 && coeff_out) *coeff_out = r;
    }
  }
  __syncthreads();
  const float step = a.lr * s_ratio;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  for (size_t e = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; e < n; e += stride) {
    const float pp = fmaf(-step, u[e], p[e]);
    p[e] = pp;
    store_lp1<LP>(lp, e, pp);
  }
}

// ------------------------------------------------------------------ multi-tensor Adam (list API parity)
constexpr int MT_MAX_TENSORS = 24;
constexpr int MT_MAX_BLOCKS = 320;
constexpr int MT_CHUNK = 2048 * 32;
struct MtMeta {
  void* p[MT_MAX_TENSORS];
  void* m[MT_MAX_TENSORS];
  void* v[MT_MAX_TENSORS];
  const void* g[MT_MAX_TENSORS];
  long long n[MT_MAX_TENSORS];
  unsigned char block_tensor[MT_MAX_BLOCKS];
  int block_chunk[MT_MAX_BLOCKS];
};
// p/m/v fp32; gradients of type GT; params updated in place (the list API has no separate low-precision copy)
template <int GT>
__global__ void __launch_bounds__(THREADS) adam_multi_kernel(MtMeta meta, AdamArgs a) {
  if (a.skip && *a.skip) return;
  const float gs = a.inv_scale ? *a.inv_scale : 1.f;
  const int t = meta.block_tensor[blockIdx.x];
  const long long off = static_cast<long long>(meta.block_chunk[blockIdx.x]) * MT_CHUNK;
  long long n = meta.n[t] - off;
  if (n > MT_CHUNK) n = MT_CHUNK;
  float* p = static_cast<float*>(meta.p[t]) + off;
  float* m = static_cast<float*>(meta.m[t]) + off;
  float* v = static_cast<float*>(meta.v[t]) + off;
  for (long long e = threadIdx.x; e < n; e += blockDim.x) {
    float pp = p[e], mm = m[e], vv = v[e];
    adam_math(pp, mm, vv, load_grad1<GT>(meta.g[t], off + e) * gs, a);
    p[e] = pp; m[e] = mm; v[e] = vv;
  }
}

static int grid_for(size_t n) {
  size_t blocks = (n / VEC + THREADS - 1) / THREADS;
  size_t cap = static_cast<size_t>(num_sms()) * 4;
  if (blocks < 1) blocks = 1;
  return static_cast<int>(blocks < cap ? blocks : cap);
}

}  // namespace optim
}  // namespace dsb

using namespace dsb;
using namespace dsb::optim;

#define DSB_DISPATCH_GT_LP(GT, LP, ...)                                              \
  [&] {                                                                              \
    if (GT == G_F32 && LP == LP_NONE) { constexpr int kG = G_F32, kL = LP_NONE; __VA_ARGS__; }   \
    else if (GT == G_F32 && LP == LP_BF16) { constexpr int kG = G_F32, kL = LP_BF16; __VA_ARGS__; } \
    else if (GT == G_F32 && LP == LP_F16) { constexpr int kG = G_F32, kL = LP_F16; __VA_ARGS__; }   \
    else if (GT == G_BF16 && LP == LP_NONE) { constexpr int kG = G_BF16, kL = LP_NONE; __VA_ARGS__; } \
    else if (GT == G_BF16 && LP == LP_BF16) { constexpr int kG = G_BF16, kL = LP_BF16; __VA_ARGS__; } \
    else if (GT == G_F16 && LP == LP_NONE) { constexpr int kG = G_F16, kL = LP_NONE; __VA_ARGS__; }   \
    else if (GT == G_F16 && LP == LP_F16) { constexpr int kG = G_F16, kL = LP_F16; __VA_ARGS__; }     \
    else { set_last_error("unsupported grad/lp dtype combination %d/%d", GT, LP); return DSB_ERR_UNSUPPORTED; } \
    return DSB_OK;                                                                   \
  }()

// Adam / AdamW on a flat fp32 shard. grad_type: 0 fp32, 1 bf16, 2 fp16. lp_type: 0 none, 1 bf16, 2 fp16 (lp = low-precision
// parameter copy written in the same pass). step >= 1. inv_scale / skip: optional DEVICE scalars (may be null).
DSB_EXPORT int dsb200_fused_adam_flat(float* p, float* m, float* v, const void* g, int grad_type, void* lp, int lp_type,
                                      long long n, float lr, float beta1, float beta2, float eps, float weight_decay,
                                      int adamw_mode, int bias_correction, int step, const float* inv_scale,
                                      const int* skip, cudaStream_t stream) {
  DSB_REQUIRE(n >= 0, "adam: negative size");
  if (n == 0) return DSB_OK;
  DSB_REQUIRE((reinterpret_cast<uintptr_t>(p) & 15) == 0 && (reinterpret_cast<uintptr_t>(m) & 15) == 0 &&
                  (reinterpret_cast<uintptr_t>(v) & 15) == 0 && (reinterpret_cast<uintptr_t>(g) & 7) == 0 &&
                  (reinterpret_cast<uintptr_t>(lp) & 7) == 0,
              "adam: buffers must be 16-byte aligned (8 for 16-bit buffers)");
  AdamArgs a;
  a.lr = lr; a.beta1 = beta1; a.beta2 = beta2; a.eps = eps; a.weight_decay = weight_decay; a.adamw = adamw_mode;
  a.bc1 = bias_correction ? 1.f / (1.f - powf(beta1, (float)step)) : 1.f;
  a.bc2_rsqrt = bias_correction ? 1.f / sqrtf(1.f - powf(beta2, (float)step)) : 1.f;
  a.inv_scale = inv_scale; a.skip = skip;
  const int grid = grid_for((size_t)n);
  int rc = DSB_DISPATCH_GT_LP(grad_type, lp_type,
                              (adam_flat_kernel<kG, kL><<<grid, THREADS, 0, stream>>>(p, m, v, g, lp, (size_t)n, a)));
  if (rc) return rc;
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

DSB_EXPORT int dsb200_fused_lion_flat(float* p, float* m, const void* g, int grad_type, void* lp, int lp_type,
                                      long long n, float lr, float beta1, float beta2, float weight_decay,
                                      const float* inv_scale, const int* skip, cudaStream_t stream) {
  if (n == 0) return DSB_OK;
  LionArgs a{lr, beta1, beta2, weight_decay, inv_scale, skip};
  const int grid = grid_for((size_t)n);
  int rc = DSB_DISPATCH_GT_LP(grad_type, lp_type,
                              (lion_flat_kernel<kG, kL><<<grid, THREADS, 0, stream>>>(p, m, g, lp, (size_t)n, a)));
  if (rc) return rc;
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

DSB_EXPORT int dsb200_fused_adagrad_flat(float* p, float* h, const void* g, int grad_type, void* lp, int lp_type,
                                         long long n, float lr, float eps, float weight_decay, const float* inv_scale,
                                         const int* skip, cudaStream_t stream) {
  if (n == 0) return DSB_OK;
  const int grid = grid_for((size_t)n);
  int rc = DSB_DISPATCH_GT_LP(grad_type, lp_type,
                              (adagrad_flat_kernel<kG, kL><<<grid, THREADS, 0, stream>>>(p, h, g, lp, (size_t)n, lr, eps,
                                                                                         weight_decay, inv_scale, skip)));
  if (rc) return rc;
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

// Number of scratch floats dsb200_fused_lamb_flat needs (u buffer excluded): 2 * lamb grid.
DSB_EXPORT int dsb200_fused_lamb_scratch_floats(long long n) { return 2 * grid_for((size_t)(n > 0 ? n : 1)); }

// LAMB on one tensor (flat fp32 p/m/v). `u` is an n-float update buffer, `scratch` holds the per-block norm partials
// (dsb200_fused_lamb_scratch_floats(n) floats). coeff_out (optional, device) receives the trust ratio.
DSB_EXPORT int dsb200_fused_lamb_flat(float* p, float* m, float* v, const void* g, int grad_type, void* lp, int lp_type,
                                      float* u, float* scratch, long long n, float lr, float beta1, float beta2, float eps,
                                      float weight_decay, int bias_correction, int step, float max_coeff, float min_coeff,
                                      const float* inv_scale, const int* skip, float* coeff_out, cudaStream_t stream) {
  if (n == 0) return DSB_OK;
  LambArgs a;
  a.lr = lr; a.beta1 = beta1; a.beta2 = beta2; a.eps = eps; a.weight_decay = weight_decay;
  a.bc1 = bias_correction ? 1.f - powf(beta1, (float)step) : 1.f;
  a.bc2 = bias_correction ? 1.f - powf(beta2, (float)step) : 1.f;
  a.max_coeff = max_coeff; a.min_coeff = min_coeff; a.inv_scale = inv_scale; a.skip = skip;
  const int grid = grid_for((size_t)n);
  if (grad_type == G_F32) lamb_phase1_kernel<G_F32><<<grid, THREADS, 0, stream>>>(p, m, v, g, u, (size_t)n, a, scratch);
  else if (grad_type == G_BF16) lamb_phase1_kernel<G_BF16><<<grid, THREADS, 0, stream>>>(p, m, v, g, u, (size_t)n, a, scratch);
  else lamb_phase1_kernel<G_F16><<<grid, THREADS, 0, stream>>>(p, m, v, g, u, (size_t)n, a, scratch);
  if (lp_type == LP_NONE) lamb_phase2_kernel<LP_NONE><<<grid, THREADS, 0, stream>>>(p, u, lp, (size_t)n, a, scratch, grid, coeff_out);
  else if (lp_type == LP_BF16) lamb_phase2_kernel<LP_BF16><<<grid, THREADS, 0, stream>>>(p, u, lp, (size_t)n, a, scratch, grid, coeff_out);
  else lamb_phase2_kernel<LP_F16><<<grid, THREADS, 0, stream>>>(p, u, lp, (size_t)n, a, scratch, grid, coeff_out);
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

// List-of-tensors Adam (fp32 p/m/v, gradients of one dtype). Pointers are HOST arrays of device pointers.
DSB_EXPORT int dsb200_multi_tensor_adam(int ntensors, void** p, void** m, void** v, const void** g, const long long* n,
                                        int grad_type, float lr, float beta1, float beta2, float eps, float weight_decay,
                                        int adamw_mode, int bias_correction, int step, const float* inv_scale,
                                        const int* skip, cudaStream_t stream) {
  AdamArgs a;
  a.lr = lr; a.beta1 = beta1; a.beta2 = beta2; a.eps = eps; a.weight_decay = weight_decay; a.adamw = adamw_mode;
  a.bc1 = bias_correction ? 1.f / (1.f - powf(beta1, (float)step)) : 1.f;
  a.bc2_rsqrt = bias_correction ? 1.f / sqrtf(1.f - powf(beta2, (float)step)) : 1.f;
  a.inv_scale = inv_scale; a.skip = skip;
  MtMeta meta;
  int nt = 0, nb = 0;
  auto flush = [&]() -> int {
    if (nb == 0) return DSB_OK;
    if (grad_type == G_F32) adam_multi_kernel<G_F32><<<nb, THREADS, 0, stream>>>(meta, a);
    else if (grad_type == G_BF16) adam_multi_kernel<G_BF16><<<nb, THREADS, 0, stream>>>(meta, a);
    else adam_multi_kernel<G_F16><<<nb, THREADS, 0, stream>>>(meta, a);
    nt = 0; nb = 0;
    DSB_CHECK_CUDA(cudaGetLastError());
    return DSB_OK;
  };
  for (int t = 0; t < ntensors; ++t) {
    if (n[t] == 0) continue;
    long long chunks = (n[t] + MT_CHUNK - 1) / MT_CHUNK;
    long long c = 0;
    while (c < chunks) {
      if (nt == MT_MAX_TENSORS || nb == MT_MAX_BLOCKS) { int rc = flush(); if (rc) return rc; }
      // (re)register tensor t in the current table
      meta.p[nt] = p[t]; meta.m[nt] = m[t]; meta.v[nt] = v[t]; meta.g[nt] = g[t]; meta.n[nt] = n[t];
      while (c < chunks && nb < MT_MAX_BLOCKS) {
        meta.block_tensor[nb] = (unsigned char)nt;
        meta.block_chunk[nb] = (int)c;
        ++nb; ++c;
      }
      ++nt;
    }
  }
  return flush();
}

// Element-wise / row-wise fused ops for the transformer path (sm_100a): RoPE (in place on packed QKV, forward and
// backward), gated activations (SiGLU / GEGLU / ReGLU, fwd + bwd), bias + activation, fused softmax-cross-entropy
// (forward loss + in-place dlogits), residual add.
// Reference counterparts: csrc/transformer/inference/csrc/apply_rotary_pos_emb.cu (rotate-half RoPE),
// deepspeed/inference/v2/kernels/core_ops/gated_activations/gated_activation_kernels.cu:104-107,
// core_ops/bias_activations/bias_activation.cu, csrc/transformer/gelu_kernels.cu, softmax_kernels.cu.
#include "../common/host.h"
#include "../common/ptx.cuh"

namespace dsb {
namespace ew {

template <typename T> __device__ __forceinline__ float to_f(T v);
template <> __device__ __forceinline__ float to_f<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <> __device__ __forceinline__ float to_f<__half>(__half v) { return __half2float(v); }
template <> __device__ __forceinline__ float to_f<float>(float v) { return v; }
template <typename T> __device__ __forceinline__ T from_f(float v);
template <> __device__ __forceinline__ __nv_bfloat16 from_f<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }
template <> __device__ __forceinline__ __half from_f<__half>(float v) { return __float2half_rn(v); }
template <> __device__ __forceinline__ float from_f<float>(float v) { return v; }

// ------------------------------------------------------------------ RoPE (rotate-half / "neox" pairing: (i, i + D/2))
// x: [tokens, row_stride] ; heads [head0, head0 + nheads) of size D start at column head0*D. pos[token] (int32) or
// token % seq_len when pos == null. inv_freq[D/2] fp32. sign = +1 forward, -1 backward (inverse rotation).
template <typename T>
__global__ void rope_kernel(T* __restrict__ x, const int* __restrict__ pos, const float* __restrict__ inv_freq,
                            long long tokens, int row_stride, int head0, int nheads, int D, int seq_len, float sign) {
  const int half = D / 2;
  const long long pairs_per_token = static_cast<long long>(nheads) * half;
  const long long total = tokens * pairs_per_token;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long t = i / pairs_per_token;
    const int r = static_cast<int>(i - t * pairs_per_token);
    const int h = r / half, j = r - h * half;
    const float p = pos ? static_cast<float>(pos[t]) : static_cast<float>(t % seq_len);
    float s, c;
    sincosf(p * inv_freq[j], &s, &c);
    s *= sign;
    T* base = x + t * row_stride + static_cast<long long>(head0 + h) * D;
    const float a = to_f(base[j]), b = to_f(base[j + half]);
    base[j] = from_f<T>(a * c - b * s);
    base[j + half] = from_f<T>(b * c + a * s);
  }
}

// Vectorised bf16/fp16 RoPE: one thread owns 8 consecutive frequencies j..j+7 of one token, computes their sin/cos
// ONCE (the angle does not depend on the head) and then walks the heads of its head group: per head two 16-byte loads
// (first / second half of the head) and two 16-byte stores.  sincosf is evaluated tokens*half*head_groups times instead
// of tokens*half*nheads, and every access is a full 16-byte vector.  Requires D % 16 == 0 and 16-byte aligned rows.
template <typename T>
__global__ void rope_vec_kernel(T* __restrict__ x, const int* __restrict__ pos, const float* __restrict__ inv_freq,
                                long long tokens, int row_stride, int head0, int nheads, int D, int seq_len, float sign,
                                int heads_per_group) {
  const int half = D / 2;
  const int vec_per_token = half / 8;
  const long long total = tokens * vec_per_token;
  const int hbeg = blockIdx.y * heads_per_group;
  const int hend = min(nheads, hbeg + heads_per_group);
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long t = i / vec_per_token;
    const int j = static_cast<int>(i - t * vec_per_token) * 8;
    const float p = pos ? static_cast<float>(pos[t]) : static_cast<float>(t % seq_len);
    float sn[8], cs[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      sincosf(p * inv_freq[j + k], &sn[k], &cs[k]);
      sn[k] *= sign;
    }
    T* row = x + t * row_stride + static_cast<long long>(head0) * D + j;
    for (int h = hbeg; h < hend; ++h) {
      T* lo = row + static_cast<long long>(h) * D;
      T* hi = lo + half;
      T a[8], b[8], ra[8], rb[8];
      *reinterpret_cast<uint4*>(a) = *reinterpret_cast<const uint4*>(lo);
      *reinterpret_cast<uint4*>(b) = *reinterpret_cast<const uint4*>(hi);
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const float af = to_f(a[k]), bf = to_f(b[k]);
        ra[k] = from_f<T>(af * cs[k] - bf * sn[k]);
        rb[k] = from_f<T>(bf * cs[k] + af * sn[k]);
      }
      *reinterpret_cast<uint4*>(lo) = *reinterpret_cast<const uint4*>(ra);
      *reinterpret_cast<uint4*>(hi) = *reinterpret_cast<const uint4*>(rb);
    }
  }
}

// ------------------------------------------------------------------ sum of squares (gradient norm), accumulated into *out
template <typename T, int VEC>
__global__ void sumsq_kernel(const T* __restrict__ x, long long n, float* __restrict__ out) {
  float acc = 0.f;
  const long long nvec = n / VEC;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long v = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; v < nvec; v += stride) {
    T e[VEC];
    *reinterpret_cast<uint4*>(e) = *reinterpret_cast<const uint4*>(x + v * VEC);
#pragma unroll
    for (int k = 0; k < VEC; ++k) {
      const float f = to_f(e[k]);
      acc += f * f;
    }
  }
  if (blockIdx.x == 0 && threadIdx.x == 0)
    for (long long i = nvec * VEC; i < n; ++i) {
      const float f = to_f(x[i]);
      acc += f * f;
    }
  acc = warp_sum(acc);
  __shared__ float ws[32];
  if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    acc = threadIdx.x < (blockDim.x >> 5) ? ws[threadIdx.x] : 0.f;
    acc = warp_sum(acc);
    if (threadIdx.x == 0) atomicAdd(out, acc);
  }
}

// ------------------------------------------------------------------ gated activation: out = act(gate) * up
// in: [rows, 2F] with gate = in[:, :F], up = in[:, F:]  (interleaved=1: gate = even cols, up = odd cols)
enum Act { ACT_IDENTITY = 0, ACT_RELU = 1, ACT_GELU = 2, ACT_SILU = 3 };
__device__ __forceinline__ float act_fwd(float x, int act) {
  switch (act) {
    case ACT_RELU: return fmaxf(x, 0.f);
    case ACT_GELU: return 0.5f * x * (1.f + tanhf(0.7978845608f * (x + 0.044715f * x * x * x)));
    case ACT_SILU: return x / (1.f + __expf(-x));
    default: return x;
  }
}
__device__ __forceinline__ float act_bwd(float x, int act) {  // d act / dx
  switch (act) {
    case ACT_RELU: return x > 0.f ? 1.f : 0.f;
    case ACT_GELU: {
      const float u = 0.7978845608f * (x + 0.044715f * x * x * x);
      const float t = tanhf(u);
      return 0.5f * (1.f + t) + 0.5f * x * (1.f - t * t) * 0.7978845608f * (1.f + 3.f * 0.044715f * x * x);
    }
    case ACT_SILU: {
      const float s = 1.f / (1.f + __expf(-x));
      return s * (1.f + x * (1.f - s));
    }
    default: return 1.f;
  }
}
template <typename T>
__global__ void gated_act_fwd_kernel(const T* __restrict__ in, const T* __restrict__ bias, T* __restrict__ out,
                                     long long rows, int F, int act, int interleaved) {
  const long long total = rows * F;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long r = i / F;
    const int c = static_cast<int>(i - r * F);
    const int gi = interleaved ? 2 * c : c, ui = interleaved ? 2 * c + 1 : F + c;
    float g = to_f(in[r * 2 * F + gi]), u = to_f(in[r * 2 * F + ui]);
    if (bias) { g += to_f(bias[gi]); u += to_f(bias[ui]); }
    out[i] = from_f<T>(act_fwd(g, act) * u);
  }
}
template <typename T>
__global__ void gated_act_bwd_kernel(const T* __restrict__ in, const T* __restrict__ dout, T* __restrict__ din,
                                     long long rows, int F, int act, int interleaved) {
  const long long total = rows * F;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long r = i / F;
    const int c = static_cast<int>(i - r * F);
    const int gi = interleaved ? 2 * c : c, ui = interleaved ? 2 * c + 1 : F + c;
    const float g = to_f(in[r * 2 * F + gi]), u = to_f(in[r * 2 * F + ui]);
    const float d = to_f(dout[i]);
    din[r * 2 * F + gi] = from_f<T>(d * u * act_bwd(g, act));
    din[r * 2 * F + ui] = from_f<T>(d * act_fwd(g, act));
  }
}

// 8-wide vectorised SiGLU (non-interleaved, no bias) -- the Llama MLP hot path
__global__ void swiglu_fwd_vec_kernel(const __nv_bfloat16* __restrict__ in, __nv_bfloat16* __restrict__ out,
                                      long long rows, int F) {
  const int fv = F / 8;
  const long long total = rows * fv;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long r = i / fv;
    const int c = static_cast<int>(i - r * fv) * 8;
    uint4 gv = *reinterpret_cast<const uint4*>(in + r * 2 * F + c);
    uint4 uv = *reinterpret_cast<const uint4*>(in + r * 2 * F + F + c);
    const __nv_bfloat162* g2 = reinterpret_cast<const __nv_bfloat162*>(&gv);
    const __nv_bfloat162* u2 = reinterpret_cast<const __nv_bfloat162*>(&uv);
    uint4 o;
    uint32_t* o32 = reinterpret_cast<uint32_t*>(&o);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      float2 g = __bfloat1622float2(g2[k]), u = __bfloat1622float2(u2[k]);
      o32[k] = pack_bf16x2(g.x / (1.f + __expf(-g.x)) * u.x, g.y / (1.f + __expf(-g.y)) * u.y);
    }
    *reinterpret_cast<uint4*>(out + r * F + c) = o;
  }
}
__global__ void swiglu_bwd_vec_kernel(const __nv_bfloat16* __restrict__ in, const __nv_bfloat16* __restrict__ dout,
                                      __nv_bfloat16* __restrict__ din, long long rows, int F) {
  const int fv = F / 8;
  const long long total = rows * fv;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long r = i / fv;
    const int c = static_cast<int>(i - r * fv) * 8;
    uint4 gv = *reinterpret_cast<const uint4*>(in + r * 2 * F + c);
    uint4 uv = *reinterpret_cast<const uint4*>(in + r * 2 * F + F + c);
    uint4 dv = *reinterpret_cast<const uint4*>(dout + r * F + c);
    const __nv_bfloat162* g2 = reinterpret_cast<const __nv_bfloat162*>(&gv);
    const __nv_bfloat162* u2 = reinterpret_cast<const __nv_bfloat162*>(&uv);
    const __nv_bfloat162* d2 = reinterpret_cast<const __nv_bfloat162*>(&dv);
    uint4 og, ou;
    uint32_t* og32 = reinterpret_cast<uint32_t*>(&og);
    uint32_t* ou32 = reinterpret_cast<uint32_t*>(&ou);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      float2 g = __bfloat1622float2(g2[k]), u = __bfloat1622float2(u2[k]), d = __bfloat1622float2(d2[k]);
      const float sx = 1.f / (1.f + __expf(-g.x)), sy = 1.f / (1.f + __expf(-g.y));
      og32[k] = pack_bf16x2(d.x * u.x * sx * (1.f + g.x * (1.f - sx)), d.y * u.y * sy * (1.f + g.y * (1.f - sy)));
      ou32[k] = pack_bf16x2(d.x * g.x * sx, d.y * g.y * sy);
    }
    *reinterpret_cast<uint4*>(din + r * 2 * F + c) = og;
    *reinterpret_cast<uint4*>(din + r * 2 * F + F + c) = ou;
  }
}

// ------------------------------------------------------------------ bias + activation (in place allowed)
template <typename T>
__global__ void bias_act_kernel(const T* __restrict__ in, const T* __restrict__ bias, T* __restrict__ out, long long rows,
                                int cols, int act) {
  const long long total = rows * cols;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(i % cols);
    float v = to_f(in[i]);
    if (bias) v += to_f(bias[c]);
    out[i] = from_f<T>(act_fwd(v, act));
  }
}

// ------------------------------------------------------------------ fused softmax cross-entropy
// One CTA per row. Forward: loss[row] = lse - logit[label] (0 for ignored rows), lse[row] saved.
// Backward (in place): logits <- (softmax - onehot) * gscale[0] * (label valid).  gscale is a device scalar
// (upstream grad / n_valid), so no host sync is needed between forward and backward.
constexpr int CE_THREADS = 512;
template <typename T>
__global__ void __launch_bounds__(CE_THREADS)
ce_fwd_kernel(const T* __restrict__ logits, const long long* __restrict__ labels, float* __restrict__ loss,
              float* __restrict__ lse_out, int V, long long ld, long long ignore_index) {
  __shared__ float red_m[CE_THREADS / 32], red_s[CE_THREADS / 32];
  const long long row = blockIdx.x;
  const T* lr = logits + row * ld;
  float m = -INFINITY, s = 0.f;
  for (int c = threadIdx.x; c < V; c += CE_THREADS) {
    const float v = to_f(lr[c]);
    const float nm = fmaxf(m, v);
    s = s * __expf(m - nm) + __expf(v - nm);
    m = nm;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float om = __shfl_xor_sync(0xffffffffu, m, o), os = __shfl_xor_sync(0xffffffffu, s, o);
    const float nm = fmaxf(m, om);
    s = s * __expf(m - nm) + os * __expf(om - nm);
    m = nm;
  }
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  if (l == 0) { red_m[w] = m; red_s[w] = s; }
  __syncthreads();
  if (w == 0) {
    m = l < CE_THREADS / 32 ? red_m[l] : -INFINITY;
    s = l < CE_THREADS / 32 ? red_s[l] : 0.f;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float om = __shfl_xor_sync(0xffffffffu, m, o), os = __shfl_xor_sync(0xffffffffu, s, o);
      const float nm = fmaxf(m, om);
      s = s * __expf(m - nm) + os * __expf(om - nm);
      m = nm;
    }
    if (l == 0) {
      const float lse = m + __logf(s);
      lse_out[row] = lse;
      const long long lab = labels[row];
      loss[row] = (lab == ignore_index || lab < 0 || lab >= V) ? 0.f : lse - to_f(lr[lab]);
    }
  }
}
template <typename T>
__global__ void __launch_bounds__(CE_THREADS)
ce_bwd_kernel(T* __restrict__ logits, const long long* __restrict__ labels, const float* __restrict__ lse,
              const float* __restrict__ gscale, int V, long long ld, long long ignore_index) {
  const long long row = blockIdx.x;
  T* lr = logits + row * ld;
  const long long lab = labels[row];
  const bool valid = !(lab == ignore_index || lab < 0 || lab >= V);
  const float g = valid ? *gscale : 0.f;
  const float l = lse[row];
  for (int c = threadIdx.x; c < V; c += CE_THREADS) {
    float p = __expf(to_f(lr[c]) - l);
    if (c == lab) p -= 1.f;
    lr[c] = from_f<T>(p * g);
  }
}

template <typename T>
__global__ void add_kernel(const T* __restrict__ a, const T* __restrict__ b, T* __restrict__ out, long long n) {
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x)
    out[i] = from_f<T>(to_f(a[i]) + to_f(b[i]));
}

static int ew_grid(long long work, int threads) {
  long long b = (work + threads - 1) / threads;
  long long cap = static_cast<long long>(num_sms()) * 16;
  if (b < 1) b = 1;
  return static_cast<int>(b < cap ? b : cap);
}

}  // namespace ew
}  // namespace dsb

using namespace dsb;
using namespace dsb::ew;

#define DSB_EW_DISPATCH(dtype, ...)                                            \
  do {                                                                         \
    if (dtype == 0) { using T = __nv_bfloat16; __VA_ARGS__; }                  \
    else if (dtype == 1) { using T = __half; __VA_ARGS__; }                    \
    else if (dtype == 2) { using T = float; __VA_ARGS__; }                     \
    else { set_last_error("bad dtype %d", dtype); return DSB_ERR_ARG; }        \
  } while (0)

DSB_EXPORT int dsb200_rope_inplace(void* x, const int* positions, const float* inv_freq, long long tokens, int row_stride,
                                   int head0, int nheads, int head_dim, int seq_len, int backward, int dtype,
                                   cudaStream_t stream) {
  DSB_REQUIRE(head_dim % 2 == 0, "rope: odd head_dim");
  const long long work = tokens * nheads * (head_dim / 2);
  if (work == 0) return DSB_OK;
  const float sign = backward ? -1.f : 1.f;
  const bool vec_ok = dtype != 2 && head_dim % 16 == 0 && row_stride % 8 == 0 && (reinterpret_cast<uintptr_t>(x) & 15) == 0;
  if (vec_ok) {
    // enough head groups to fill the machine, few enough to keep the sincos amortisation
    const long long vecs = tokens * (head_dim / 16);
    int groups = 1;
    while (groups < nheads && vecs * groups < static_cast<long long>(num_sms()) * 2048) groups *= 2;
    const int hpg = (nheads + groups - 1) / groups;
    dim3 grid(ew_grid(vecs, 128), (nheads + hpg - 1) / hpg);
    if (dtype == 0)
      rope_vec_kernel<__nv_bfloat16><<<grid, 128, 0, stream>>>((__nv_bfloat16*)x, positions, inv_freq, tokens, row_stride, head0,
                                                              nheads, head_dim, seq_len, sign, hpg);
    else
      rope_vec_kernel<__half><<<grid, 128, 0, stream>>>((__half*)x, positions, inv_freq, tokens, row_stride, head0, nheads,
                                                       head_dim, seq_len, sign, hpg);
  } else {
    DSB_EW_DISPATCH(dtype, (rope_kernel<T><<<ew_grid(work, 256), 256, 0, stream>>>((T*)x, positions, inv_freq, tokens, row_stride,
                                                                                head0, nheads, head_dim, seq_len, sign)));
  }
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

// out[0] += sum(x[i]^2) over n elements (dtype 0 bf16 / 1 fp16 / 2 fp32).  x must be 16-byte aligned.
DSB_EXPORT int dsb200_sumsq(const void* x, long long n, int dtype, float* out, cudaStream_t stream) {
  if (n == 0) return DSB_OK;
  DSB_REQUIRE((reinterpret_cast<uintptr_t>(x) & 15) == 0, "sumsq: x must be 16-byte aligned");
  const int threads = 512;
  long long want = (n / 8 + threads - 1) / threads;
  const long long cap = static_cast<long long>(num_sms()) * 4;
  const int grid = static_cast<int>(want < cap ? (want > 0 ? want : 1) : cap);
  if (dtype == 0)
    sumsq_kernel<__nv_bfloat16, 8><<<grid, threads, 0, stream>>>((const __nv_bfloat16*)x, n, out);
  else if (dtype == 1)
    sumsq_kernel<__half, 8><<<grid, threads, 0, stream>>>((const __half*)x, n, out);
  else
    sumsq_kernel<float, 4><<<grid, threads, 0, stream>>>((const float*)x, n, out);
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

DSB_EXPORT int dsb200_gated_act_fwd(const void* in, const void* bias, void* out, long long rows, int F, int act,
                                    int interleaved, int dtype, cudaStream_t stream) {
  if (rows * F == 0) return DSB_OK;
  if (dtype == 0 && act == ACT_SILU && !interleaved && !bias && F % 8 == 0) {
    swiglu_fwd_vec_kernel<<<ew_grid(rows * (F / 8), 256), 256, 0, stream>>>((const __nv_bfloat16*)in, (__nv_bfloat16*)out, rows, F);
  } else {
    DSB_EW_DISPATCH(dtype, (gated_act_fwd_kernel<T><<<ew_grid(rows * F, 256), 256, 0, stream>>>((const T*)in, (const T*)bias,
                                                                                            (T*)out, rows, F, act, interleaved)));
  }
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

DSB_EXPORT int dsb200_gated_act_bwd(const void* in, const void* dout, void* din, long long rows, int F, int act,
                                    int interleaved, int dtype, cudaStream_t stream) {
  if (rows * F == 0) return DSB_OK;
  if (dtype == 0 && act == ACT_SILU && !interleaved && F % 8 == 0) {
    swiglu_bwd_vec_kernel<<<ew_grid(rows * (F / 8), 256), 256, 0, stream>>>((const __nv_bfloat16*)in, (const __nv_bfloat16*)dout,
                                                                         (__nv_bfloat16*)din, rows, F);
  } else {
    DSB_EW_DISPATCH(dtype, (gated_act_bwd_kernel<T><<<ew_grid(rows * F, 256), 256, 0, stream>>>((const T*)in, (const T*)dout,
                                                                                            (T*)din, rows, F, act, interleaved)));
  }
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

DSB_EXPORT int dsb200_bias_act(const void* in, const void* bias, void* out, long long rows, int cols, int act, int dtype,
                               cudaStream_t stream) {
  if (rows * cols == 0) return DSB_OK;
  DSB_EW_DISPATCH(dtype, (bias_act_kernel<T><<<ew_grid(rows * cols, 256), 256, 0, stream>>>((const T*)in, (const T*)bias, (T*)out,
                                                                                        rows, cols, act)));
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

DSB_EXPORT int dsb200_cross_entropy_fwd(const void* logits, const long long* labels, float* loss, float* lse, long long rows,
                                        int V, long long ld, long long ignore_index, int dtype, cudaStream_t stream) {
  if (rows == 0) return DSB_OK;
  DSB_EW_DISPATCH(dtype, (ce_fwd_kernel<T><<<(unsigned)rows, CE_THREADS, 0, stream>>>((const T*)logits, labels, loss, lse, V, ld,
                                                                                  ignore_index)));
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

DSB_EXPORT int dsb200_cross_entropy_bwd(void* logits, const long long* labels, const float* lse, const float* gscale,
                                        long long rows, int V, long long ld, long long ignore_index, int dtype,
                                        cudaStream_t stream) {
  if (rows == 0) return DSB_OK;
  DSB_EW_DISPATCH(dtype, (ce_bwd_kernel<T><<<(unsigned)rows, CE_THREADS, 0, stream>>>((T*)logits, labels, lse, gscale, V, ld,
                                                                                  ignore_index)));
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

DSB_EXPORT int dsb200_add(const void* a, const void* b, void* out, long long n, int dtype, cudaStream_t stream) {
  if (n == 0) return DSB_OK;
  DSB_EW_DISPATCH(dtype, (add_kernel<T><<<ew_grid(n, 256), 256, 0, stream>>>((const T*)a, (const T*)b, (T*)out, n)));
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

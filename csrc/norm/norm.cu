// RMSNorm / LayerNorm forward + backward for sm_100a (bf16 / fp16 / fp32 activations, fp32 statistics).
// One CTA per row, 16-byte vector access, the row stays in registers between the statistics pass and the
// normalisation pass; optional fused residual add ("pre-norm": res_out = x + residual; y = norm(res_out)).
// Backward: dx per row + per-CTA partial dgamma/dbeta kept in registers over a persistent row loop, written to a
// [grid, hidden] fp32 scratch and column-reduced by a second tiny kernel (deterministic, no atomics).
// Reference counterparts: deepspeed/inference/v2/kernels/core_ops/cuda_rms_norm/rms_norm.cu:15,203-211,
// cuda_layer_norm/layer_norm.cu:147-155, csrc/transformer/normalize_kernels.cu (fwd/bwd),
// csrc/transformer/inference/csrc/{rms_norm,layer_norm}.cu.
#include "../common/host.h"
#include "../common/ptx.cuh"

namespace dsb {
namespace norm {

constexpr int THREADS = 256;
constexpr int MAXV_LIMIT = 8;  // up to 8 x 8-element vectors per thread in registers (hidden <= 256*8*8 = 16384)

template <typename T> struct Vec;  // 16-byte vector of T
template <> struct Vec<__nv_bfloat16> {
  static constexpr int N = 8;
  static __device__ __forceinline__ void load(const __nv_bfloat16* p, float (&f)[8]) {
    uint4 r = *reinterpret_cast<const uint4*>(p);
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&r);
#pragma unroll
    for (int i = 0; i < 4; ++i) { float2 t = __bfloat1622float2(h[i]); f[2 * i] = t.x; f[2 * i + 1] = t.y; }
  }
  static __device__ __forceinline__ void store(__nv_bfloat16* p, const float (&f)[8]) {
    uint4 r;
    r.x = pack_bf16x2(f[0], f[1]); r.y = pack_bf16x2(f[2], f[3]); r.z = pack_bf16x2(f[4], f[5]); r.w = pack_bf16x2(f[6], f[7]);
    *reinterpret_cast<uint4*>(p) = r;
  }
};
template <> struct Vec<__half> {
  static constexpr int N = 8;
  static __device__ __forceinline__ void load(const __half* p, float (&f)[8]) {
    uint4 r = *reinterpret_cast<const uint4*>(p);
    const __half2* h = reinterpret_cast<const __half2*>(&r);
#pragma unroll
    for (int i = 0; i < 4; ++i) { float2 t = __half22float2(h[i]); f[2 * i] = t.x; f[2 * i + 1] = t.y; }
  }
  static __device__ __forceinline__ void store(__half* p, const float (&f)[8]) {
    uint4 r;
    __half2* h = reinterpret_cast<__half2*>(&r);
#pragma unroll
    for (int i = 0; i < 4; ++i) h[i] = __floats2half2_rn(f[2 * i], f[2 * i + 1]);
    *reinterpret_cast<uint4*>(p) = r;
  }
};
template <> struct Vec<float> {
  static constexpr int N = 8;  // two float4
  static __device__ __forceinline__ void load(const float* p, float (&f)[8]) {
    float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
    f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
  }
  static __device__ __forceinline__ void store(float* p, const float (&f)[8]) {
    *reinterpret_cast<float4*>(p) = make_float4(f[0], f[1], f[2], f[3]);
    *reinterpret_cast<float4*>(p + 4) = make_float4(f[4], f[5], f[6], f[7]);
  }
};

__device__ __forceinline__ float block_sum(float v, float* red) {
  v = warp_sum(v);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  __syncthreads();
  if (l == 0) red[w] = v;
  __syncthreads();
  float t = (l < THREADS / 32) ? red[l] : 0.f;
  return warp_sum(t);
}

// ---------------------------------------------------------------- forward
// LN = false: RMSNorm (beta ignored). residual / res_out / stats may be null.
template <typename T, bool LN, int MAXV>
__global__ void __launch_bounds__(THREADS)
norm_fwd_kernel(const T* __restrict__ x, const T* __restrict__ residual, const T* __restrict__ gamma,
                const T* __restrict__ beta, T* __restrict__ y, T* __restrict__ res_out, float* __restrict__ mean_out,
                float* __restrict__ rstd_out, int rows, int hidden, float eps) {
  __shared__ float red[THREADS / 32];
  const int nvec = hidden / 8;
  for (int row = blockIdx.x; row < rows; row += gridDim.x) {
    const T* xr = x + static_cast<size_t>(row) * hidden;
    float v[MAXV][8];
    float s = 0.f, ss = 0.f;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      const int c = threadIdx.x + i * THREADS;
      if (c < nvec) {
        Vec<T>::load(xr + c * 8, v[i]);
        if (residual) {
          float r[8];
          Vec<T>::load(residual + static_cast<size_t>(row) * hidden + c * 8, r);
#pragma unroll
          for (int k = 0; k < 8; ++k) v[i][k] += r[k];
          if (res_out) Vec<T>::store(res_out + static_cast<size_t>(row) * hidden + c * 8, v[i]);
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) { s += v[i][k]; ss = fmaf(v[i][k], v[i][k], ss); }
      }
    }
    float mean = 0.f, rstd;
    if constexpr (LN) {
      mean = block_sum(s, red) / hidden;
      float var = 0.f;
#pragma unroll
      for (int i = 0; i < MAXV; ++i) {
        const int c = threadIdx.x + i * THREADS;
        if (c < nvec) {
#pragma unroll
          for (int k = 0; k < 8; ++k) { const float d = v[i][k] - mean; var = fmaf(d, d, var); }
        }
      }
      rstd = rsqrtf(block_sum(var, red) / hidden + eps);
    } else {
      rstd = rsqrtf(block_sum(ss, red) / hidden + eps);
    }
    if (threadIdx.x == 0) {
      if (rstd_out) rstd_out[row] = rstd;
      if (LN && mean_out) mean_out[row] = mean;
    }
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      const int c = threadIdx.x + i * THREADS;
      if (c < nvec) {
        float g[8], o[8];
        Vec<T>::load(gamma + c * 8, g);
        if constexpr (LN) {
          float b[8];
          if (beta) Vec<T>::load(beta + c * 8, b);
#pragma unroll
          for (int k = 0; k < 8; ++k) o[k] = (v[i][k] - mean) * rstd * g[k] + (beta ? b[k] : 0.f);
        } else {
#pragma unroll
          for (int k = 0; k < 8; ++k) o[k] = v[i][k] * rstd * g[k];
        }
        Vec<T>::store(y + static_cast<size_t>(row) * hidden + c * 8, o);
      }
    }
  }
}

// ---------------------------------------------------------------- backward
// dx = rstd * (dy*g - xhat * mean(dy*g*xhat) [- mean(dy*g) for LN]);  partial dgamma/dbeta per CTA -> scratch.
// dx_add (optional): gradient flowing into the same tensor from the residual branch, added to dx (fused add).
template <typename T, bool LN, int MAXV>
__global__ void __launch_bounds__(THREADS)
norm_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ x, const T* __restrict__ gamma,
                const float* __restrict__ mean_in, const float* __restrict__ rstd_in, const T* __restrict__ dx_add,
                T* __restrict__ dx, float* __restrict__ dg_part, float* __restrict__ db_part, int rows, int hidden) {
  __shared__ float red[THREADS / 32];
  const int nvec = hidden / 8;
  float dg[MAXV][8], db[MAXV][8];
#pragma unroll
  for (int i = 0; i < MAXV; ++i)
#pragma unroll
    for (int k = 0; k < 8; ++k) { dg[i][k] = 0.f; db[i][k] = 0.f; }
  for (int row = blockIdx.x; row < rows; row += gridDim.x) {
    const size_t base = static_cast<size_t>(row) * hidden;
    const float rstd = rstd_in[row];
    const float mean = LN ? mean_in[row] : 0.f;
    float xh[MAXV][8], wdy[MAXV][8];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      const int c = threadIdx.x + i * THREADS;
      if (c < nvec) {
        float g[8], d[8];
        Vec<T>::load(x + base + c * 8, xh[i]);
        Vec<T>::load(dy + base + c * 8, d);
        Vec<T>::load(gamma + c * 8, g);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          xh[i][k] = (xh[i][k] - mean) * rstd;
          wdy[i][k] = d[k] * g[k];
          s1 = fmaf(wdy[i][k], xh[i][k], s1);
          s2 += wdy[i][k];
          dg[i][k] = fmaf(d[k], xh[i][k], dg[i][k]);
          if constexpr (LN) db[i][k] += d[k];
        }
      }
    }
    s1 = block_sum(s1, red) / hidden;
    if constexpr (LN) s2 = block_sum(s2, red) / hidden; else s2 = 0.f;
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      const int c = threadIdx.x + i * THREADS;
      if (c < nvec) {
        float o[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) o[k] = rstd * (wdy[i][k] - xh[i][k] * s1 - s2);
        if (dx_add) {
          float a[8];
          Vec<T>::load(dx_add + base + c * 8, a);
#pragma unroll
          for (int k = 0; k < 8; ++k) o[k] += a[k];
        }
        Vec<T>::store(dx + base + c * 8, o);
      }
    }
  }
#pragma unroll
  for (int i = 0; i < MAXV; ++i) {
    const int c = threadIdx.x + i * THREADS;
    if (c < nvec) {
      float* dst = dg_part + static_cast<size_t>(blockIdx.x) * hidden + c * 8;
      *reinterpret_cast<float4*>(dst) = make_float4(dg[i][0], dg[i][1], dg[i][2], dg[i][3]);
      *reinterpret_cast<float4*>(dst + 4) = make_float4(dg[i][4], dg[i][5], dg[i][6], dg[i][7]);
      if constexpr (LN) {
        float* dstb = db_part + static_cast<size_t>(blockIdx.x) * hidden + c * 8;
        *reinterpret_cast<float4*>(dstb) = make_float4(db[i][0], db[i][1], db[i][2], db[i][3]);
        *reinterpret_cast<float4*>(dstb + 4) = make_float4(db[i][4], db[i][5], db[i][6], db[i][7]);
      }
    }
  }
}

// out[c] (=|+=) sum_p part[p][c]   -- out dtype fp32 (main grad) or T
template <typename TO>
__global__ void colreduce_kernel(const float* __restrict__ part, TO* __restrict__ out, int nparts, int hidden,
                                 int accumulate) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= hidden) return;
  float s = 0.f;
  for (int p = 0; p < nparts; ++p) s += part[static_cast<size_t>(p) * hidden + c];
  if constexpr (sizeof(TO) == 4) {
    out[c] = accumulate ? out[c] + s : s;
  } else {
    float prev = accumulate ? static_cast<float>(out[c]) : 0.f;
    out[c] = static_cast<TO>(prev + s);
  }
}

static int bwd_grid(int rows) {
  int g = num_sms() * 2;
  return rows < g ? rows : g;
}

}  // namespace norm
}  // namespace dsb

using namespace dsb;
using namespace dsb::norm;

// dtype: 0 bf16, 1 fp16, 2 fp32
#define DSB_NORM_DISPATCH_T(dtype, LNFLAG, NVV, ...)                                              \
  do {                                                                                            \
    constexpr int kNV = NVV;                                                                      \
    if (dtype == 0) { using T = __nv_bfloat16; constexpr bool kLN = LNFLAG; __VA_ARGS__; }        \
    else if (dtype == 1) { using T = __half; constexpr bool kLN = LNFLAG; __VA_ARGS__; }          \
    else if (dtype == 2) { using T = float; constexpr bool kLN = LNFLAG; __VA_ARGS__; }           \
    else { set_last_error("norm: bad dtype %d", dtype); return DSB_ERR_ARG; }                     \
  } while (0)
#define DSB_NORM_DISPATCH(dtype, LNFLAG, ...)                                                     \
  do {                                                                                            \
    const int _nv = (hidden / 8 + THREADS - 1) / THREADS;                                         \
    if (_nv <= 1) DSB_NORM_DISPATCH_T(dtype, LNFLAG, 1, __VA_ARGS__);                             \
    else if (_nv <= 2) DSB_NORM_DISPATCH_T(dtype, LNFLAG, 2, __VA_ARGS__);                        \
    else if (_nv <= 4) DSB_NORM_DISPATCH_T(dtype, LNFLAG, 4, __VA_ARGS__);                        \
    else DSB_NORM_DISPATCH_T(dtype, LNFLAG, 8, __VA_ARGS__);                                      \
  } while (0)

// y = norm(x [+ residual]) * gamma [+ beta].  layer_norm: 0 RMSNorm, 1 LayerNorm.  residual/res_out/beta/mean/rstd optional.
DSB_EXPORT int dsb200_norm_fwd(const void* x, const void* residual, const void* gamma, const void* beta, void* y,
                               void* res_out, float* mean, float* rstd, int rows, int hidden, float eps, int layer_norm,
                               int dtype, cudaStream_t stream) {
  DSB_REQUIRE(hidden % 8 == 0 && hidden <= THREADS * MAXV_LIMIT * 8, "norm: hidden=%d must be a multiple of 8 and <= %d", hidden,
              THREADS * MAXV_LIMIT * 8);
  if (rows == 0) return DSB_OK;
  const int grid = rows < num_sms() * 8 ? rows : num_sms() * 8;
  if (layer_norm)
    DSB_NORM_DISPATCH(dtype, true, (norm_fwd_kernel<T, kLN, kNV><<<grid, THREADS, 0, stream>>>(
                                       (const T*)x, (const T*)residual, (const T*)gamma, (const T*)beta, (T*)y, (T*)res_out,
                                       mean, rstd, rows, hidden, eps)));
  else
    DSB_NORM_DISPATCH(dtype, false, (norm_fwd_kernel<T, kLN, kNV><<<grid, THREADS, 0, stream>>>(
                                        (const T*)x, (const T*)residual, (const T*)gamma, (const T*)beta, (T*)y, (T*)res_out,
                                        mean, rstd, rows, hidden, eps)));
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

// Scratch floats needed by dsb200_norm_bwd: parts * hidden (x2 for LayerNorm).
DSB_EXPORT long long dsb200_norm_bwd_scratch_floats(int rows, int hidden, int layer_norm) {
  return static_cast<long long>(bwd_grid(rows)) * hidden * (layer_norm ? 2 : 1);
}

// dgamma/dbeta: fp32 if dparam_f32 else same dtype as activations; accumulate: += into dgamma/dbeta.
DSB_EXPORT int dsb200_norm_bwd(const void* dy, const void* x, const void* gamma, const float* mean, const float* rstd,
                               const void* dx_add, void* dx, void* dgamma, void* dbeta, float* scratch, int rows, int hidden,
                               int layer_norm, int dtype, int dparam_f32, int accumulate, cudaStream_t stream) {
  DSB_REQUIRE(hidden % 8 == 0 && hidden <= THREADS * MAXV_LIMIT * 8, "norm: hidden=%d unsupported", hidden);
  if (rows == 0) return DSB_OK;
  const int grid = bwd_grid(rows);
  float* dg_part = scratch;
  float* db_part = scratch + static_cast<size_t>(grid) * hidden;
  if (layer_norm)
    DSB_NORM_DISPATCH(dtype, true, (norm_bwd_kernel<T, kLN, kNV><<<grid, THREADS, 0, stream>>>(
                                       (const T*)dy, (const T*)x, (const T*)gamma, mean, rstd, (const T*)dx_add, (T*)dx,
                                       dg_part, db_part, rows, hidden)));
  else
    DSB_NORM_DISPATCH(dtype, false, (norm_bwd_kernel<T, kLN, kNV><<<grid, THREADS, 0, stream>>>(
                                        (const T*)dy, (const T*)x, (const T*)gamma, mean, rstd, (const T*)dx_add, (T*)dx,
                                        dg_part, db_part, rows, hidden)));
  const int cb = (hidden + 255) / 256;
  if (dparam_f32 || dtype == 2) {
    colreduce_kernel<float><<<cb, 256, 0, stream>>>(dg_part, (float*)dgamma, grid, hidden, accumulate);
    if (layer_norm && dbeta) colreduce_kernel<float><<<cb, 256, 0, stream>>>(db_part, (float*)dbeta, grid, hidden, accumulate);
  } else if (dtype == 0) {
    colreduce_kernel<__nv_bfloat16><<<cb, 256, 0, stream>>>(dg_part, (__nv_bfloat16*)dgamma, grid, hidden, accumulate);
    if (layer_norm && dbeta) colreduce_kernel<__nv_bfloat16><<<cb, 256, 0, stream>>>(db_part, (__nv_bfloat16*)dbeta, grid, hidden, accumulate);
  } else {
    colreduce_kernel<__half><<<cb, 256, 0, stream>>>(dg_part, (__half*)dgamma, grid, hidden, accumulate);
    if (layer_norm && dbeta) colreduce_kernel<__half><<<cb, 256, 0, stream>>>(db_part, (__half*)dbeta, grid, hidden, accumulate);
  }
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// RMSNorm backward without any allocation.
//
// Some callers of this library cannot obtain device memory at the moment they need a kernel: a serving process that
// handed all free HBM to its paged KV-cache pool at start-up gets "out of memory" from any later cudaMalloc, and a
// latency-critical loop cannot afford the device-wide serialisation that cudaMalloc / cudaFree imply.  For them the
// per-CTA dgamma partials of the backward pass live in the fixed array below instead of a caller-supplied workspace.
//
//   * Capacity: DSB_NOALLOC_NORM_MAX_FLOATS floats (148 CTAs x hidden <= 8192, 4.6 MiB).  Every GPU that loads the
//     library keeps this array resident until the process exits.  Larger problems are rejected with an error.
//   * The first call on a device resolves the array's device address (cudaGetSymbolAddress, with that device
//     current).  Make that first call before starting any stream capture.
//   * All calls share the one array and it keeps whatever the previous call left in it.  Nothing has to start at
//     zero: each CTA overwrites its whole row of partials before the column reduction reads it.  Two calls that
//     overlap on the GPU (different streams) race on the array -- serialise them.
//   * No cudaMalloc / cudaFree / allocating library routine is called.
#define DSB_NOALLOC_NORM_MAX_PARTS 148
#define DSB_NOALLOC_NORM_MAX_HIDDEN 8192
#define DSB_NOALLOC_NORM_MAX_FLOATS (DSB_NOALLOC_NORM_MAX_PARTS * DSB_NOALLOC_NORM_MAX_HIDDEN)
static __device__ __align__(256) float g_noalloc_norm_scratch[DSB_NOALLOC_NORM_MAX_FLOATS];

DSB_EXPORT int dsb200_rms_norm_bwd_noalloc(const void* dy, const void* x, const void* gamma, const float* rstd,
                                           const void* dx_add, void* dx, void* dgamma, int rows, int hidden, int dtype,
                                           int dparam_f32, int accumulate, cudaStream_t stream) {
  static float* scratch_addr[64] = {nullptr};  // per device; in host code the symbol itself is NOT a device address
  DSB_REQUIRE(hidden % 8 == 0 && hidden > 0, "rms_norm_bwd_noalloc: hidden=%d must be a positive multiple of 8", hidden);
  DSB_REQUIRE(hidden <= DSB_NOALLOC_NORM_MAX_HIDDEN,
              "rms_norm_bwd_noalloc: hidden=%d exceeds the built-in scratch (max hidden %d, %d floats); use dsb200_norm_bwd "
              "with a caller-provided workspace",
              hidden, DSB_NOALLOC_NORM_MAX_HIDDEN, DSB_NOALLOC_NORM_MAX_FLOATS);
  if (rows == 0) return DSB_OK;
  int dev = 0;
  DSB_CHECK_CUDA(cudaGetDevice(&dev));
  DSB_REQUIRE(dev >= 0 && dev < 64, "rms_norm_bwd_noalloc: device index %d out of range", dev);
  if (scratch_addr[dev] == nullptr) {
    float* p = nullptr;
    DSB_CHECK_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_noalloc_norm_scratch));
    scratch_addr[dev] = p;
  }
  float* scratch = scratch_addr[dev];
  const int grid = rows < DSB_NOALLOC_NORM_MAX_PARTS ? rows : DSB_NOALLOC_NORM_MAX_PARTS;
  DSB_REQUIRE(static_cast<long long>(grid) * hidden <= DSB_NOALLOC_NORM_MAX_FLOATS, "rms_norm_bwd_noalloc: scratch overflow");
  DSB_NORM_DISPATCH(dtype, false, (norm_bwd_kernel<T, kLN, kNV><<<grid, THREADS, 0, stream>>>(
                                      (const T*)dy, (const T*)x, (const T*)gamma, nullptr, rstd, (const T*)dx_add, (T*)dx,
                                      scratch, nullptr, rows, hidden)));
  const int cb = (hidden + 255) / 256;
  if (dparam_f32 || dtype == 2)
    colreduce_kernel<float><<<cb, 256, 0, stream>>>(scratch, (float*)dgamma, grid, hidden, accumulate);
  else if (dtype == 0)
    colreduce_kernel<__nv_bfloat16><<<cb, 256, 0, stream>>>(scratch, (__nv_bfloat16*)dgamma, grid, hidden, accumulate);
  else
    colreduce_kernel<__half><<<cb, 256, 0, stream>>>(scratch, (__half*)dgamma, grid, hidden, accumulate);
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

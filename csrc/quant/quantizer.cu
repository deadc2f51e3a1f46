// Quantizer suite for sm_100a:
//   * block-scaled fp8 (e4m3 payload + one fp32 scale per 128 elements) quantize / dequantize / dequant-reduce --
//     the wire format of ZeRO++ qwZ (weights) and qgZ (gradients) in this framework;
//   * int8 / int4 group-wise symmetric / asymmetric quantize + dequantize (API parity with the reference);
//   * fake quantization (quantize-dequantize in place) with nearest or stochastic rounding for MoQ / QAT.
// Reference counterparts: csrc/quantization/quantize.cu:92-151 (cached_quantization), dequantize.cu:36-42,
// swizzled_quantize.cu:117-157, quant_reduce.cu:147-263 (dequant_reduce), fake_quantizer.cu (fake_quantize_kernel*,
// sr_fake_quantize_kernel*), quantize_intX.cu.
#include <cuda_fp8.h>

#include "../common/host.h"
#include "../common/ptx.cuh"

namespace dsb {
namespace quant {

constexpr int FP8_BLOCK = 128;
constexpr float FP8_MAX = 448.f;

template <typename T> __device__ __forceinline__ float ldf(const T* p, size_t i);
template <> __device__ __forceinline__ float ldf<__nv_bfloat16>(const __nv_bfloat16* p, size_t i) { return __bfloat162float(p[i]); }
template <> __device__ __forceinline__ float ldf<__half>(const __half* p, size_t i) { return __half2float(p[i]); }
template <> __device__ __forceinline__ float ldf<float>(const float* p, size_t i) { return p[i]; }
template <typename T> __device__ __forceinline__ void stf(T* p, size_t i, float v);
template <> __device__ __forceinline__ void stf<__nv_bfloat16>(__nv_bfloat16* p, size_t i, float v) { p[i] = __float2bfloat16_rn(v); }
template <> __device__ __forceinline__ void stf<__half>(__half* p, size_t i, float v) { p[i] = __float2half_rn(v); }
template <> __device__ __forceinline__ void stf<float>(float* p, size_t i, float v) { p[i] = v; }

// one warp per 128-element block: 4 elements per lane
template <typename T>
__global__ void fp8_quant_kernel(const T* __restrict__ x, uint8_t* __restrict__ q, float* __restrict__ scales, size_t nblocks) {
  const size_t warp = (static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= nblocks) return;
  const size_t base = warp * FP8_BLOCK + lane * 4;
  float v[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) v[k] = ldf<T>(x, base + k);
  float amax = fmaxf(fmaxf(fabsf(v[0]), fabsf(v[1])), fmaxf(fabsf(v[2]), fabsf(v[3])));
  amax = warp_max(amax);
  const float scale = amax > 0.f ? amax / FP8_MAX : 1.f;
  const float inv = 1.f / scale;
  uint32_t packed = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    __nv_fp8_e4m3 f(v[k] * inv);
    packed |= static_cast<uint32_t>(f.__x) << (8 * k);
  }
  *reinterpret_cast<uint32_t*>(q + base) = packed;
  if (lane == 0) scales[warp] = scale;
}

template <typename T>
__global__ void fp8_dequant_kernel(const uint8_t* __restrict__ q, const float* __restrict__ scales, T* __restrict__ out,
                                   size_t nblocks) {
  const size_t warp = (static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= nblocks) return;
  const size_t base = warp * FP8_BLOCK + lane * 4;
  const uint32_t packed = *reinterpret_cast<const uint32_t*>(q + base);
  const float s = scales[warp];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    __nv_fp8_e4m3 f;
    f.__x = static_cast<__nv_fp8_storage_t>((packed >> (8 * k)) & 0xff);
    stf<T>(out, base + k, static_cast<float>(f) * s);
  }
}

// out[i] = scale_out * sum_{p < parts} dequant(q[p*n + i])  : the consumer side of the quantized reduce-scatter (qgZ)
template <typename T>
__global__ void fp8_dequant_reduce_kernel(const uint8_t* __restrict__ q, const float* __restrict__ scales, T* __restrict__ out,
                                          size_t nblocks_per_part, int parts, float scale_out) {
  const size_t warp = (static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= nblocks_per_part) return;
  float acc[4] = {0.f, 0.f, 0.f, 0.f};
  for (int p = 0; p < parts; ++p) {
    const size_t blk = static_cast<size_t>(p) * nblocks_per_part + warp;
    const uint32_t packed = *reinterpret_cast<const uint32_t*>(q + blk * FP8_BLOCK + lane * 4);
    const float s = scales[blk];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      __nv_fp8_e4m3 f;
      f.__x = static_cast<__nv_fp8_storage_t>((packed >> (8 * k)) & 0xff);
      acc[k] += static_cast<float>(f) * s;
    }
  }
  const size_t base = warp * FP8_BLOCK + lane * 4;
#pragma unroll
  for (int k = 0; k < 4; ++k) stf<T>(out, base + k, acc[k] * scale_out);
}

// ---------------------------------------------------------------- per-row (per-token) fp8 quantisation for the W8A8 linear
// One warp per row: the row is read once into registers (up to 32 x 32 = 8192 elements... longer rows take a second read),
// amax by warp shuffle, scale = amax / 448, payload e4m3.  Replaces an abs / amax / div / clamp / cast chain of five eager
// passes over the activation in front of every fp8 GEMM.
template <typename T>
__global__ void fp8_rowwise_quant_kernel(const T* __restrict__ x, uint8_t* __restrict__ q, float* __restrict__ scales, int rows, int cols,
                                         long long ldx) {
  const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (row >= rows) return;
  const T* xr = x + static_cast<size_t>(row) * ldx;
  uint8_t* qr = q + static_cast<size_t>(row) * cols;
  constexpr int MAXV = 8;  // vectors of 8 elements cached per lane: rows up to 32 * 8 * 8 = 2048 elements stay in registers
  const int nvec = cols / 8;
  float cache[MAXV][8];
  float amax = 0.f;
#pragma unroll
  for (int j = 0; j < MAXV; ++j) {
    const int v = lane + j * 32;
    if (v < nvec) {
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        cache[j][k] = ldf<T>(xr, static_cast<size_t>(v) * 8 + k);
        amax = fmaxf(amax, fabsf(cache[j][k]));
      }
    }
  }
  for (int v = lane + MAXV * 32; v < nvec; v += 32)
#pragma unroll
    for (int k = 0; k < 8; ++k) amax = fmaxf(amax, fabsf(ldf<T>(xr, static_cast<size_t>(v) * 8 + k)));
  amax = warp_max(amax);
  const float scale = fmaxf(amax, 1e-12f) / FP8_MAX;
  const float inv = 1.f / scale;
  auto pack8 = [&](const float* f) {
    uint2 o;
    uint32_t w[2] = {0, 0};
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      __nv_fp8_e4m3 e(fminf(fmaxf(f[k] * inv, -FP8_MAX), FP8_MAX));
      w[k >> 2] |= static_cast<uint32_t>(e.__x) << (8 * (k & 3));
    }
    o.x = w[0]; o.y = w[1];
    return o;
  };
#pragma unroll
  for (int j = 0; j < MAXV; ++j) {
    const int v = lane + j * 32;
    if (v < nvec) *reinterpret_cast<uint2*>(qr + static_cast<size_t>(v) * 8) = pack8(cache[j]);
  }
  for (int v = lane + MAXV * 32; v < nvec; v += 32) {
    float f[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) f[k] = ldf<T>(xr, static_cast<size_t>(v) * 8 + k);
    *reinterpret_cast<uint2*>(qr + static_cast<size_t>(v) * 8) = pack8(f);
  }
  if (lane == 0) scales[row] = scale;
}

// ---------------------------------------------------------------- int8 / int4 group quantization (one CTA per group)
__device__ __forceinline__ float block_reduce(float v, float* red, bool is_max) {
  v = is_max ? warp_max(v) : -warp_max(-v);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  __syncthreads();
  if (l == 0) red[w] = v;
  __syncthreads();
  float t = (l < (blockDim.x >> 5)) ? red[l] : (is_max ? -INFINITY : INFINITY);
  t = is_max ? warp_max(t) : -warp_max(-t);
  return __shfl_sync(0xffffffffu, t, 0);
}

// params: [groups][2] = (scale, offset).  bits = 8 or 4 (int4: two values per byte, low nibble first).
template <typename T>
__global__ void int_quant_kernel(const T* __restrict__ x, int8_t* __restrict__ q, float* __restrict__ params, int group_size,
                                 int bits, int symmetric) {
  __shared__ float red[32];
  const size_t g = blockIdx.x;
  const T* xg = x + g * group_size;
  float mx = -INFINITY, mn = INFINITY;
  for (int i = threadIdx.x; i < group_size; i += blockDim.x) {
    const float v = ldf<T>(xg, i);
    mx = fmaxf(mx, v);
    mn = fminf(mn, v);
  }
  mx = block_reduce(mx, red, true);
  mn = block_reduce(mn, red, false);
  const float qmax = static_cast<float>((1 << (bits - 1)) - 1);
  float scale, offset;
  if (symmetric) {
    const float amax = fmaxf(fabsf(mx), fabsf(mn));
    scale = amax > 0.f ? amax / qmax : 1.f;
    offset = 0.f;
  } else {
    const float range = mx - mn;
    scale = range > 0.f ? range / static_cast<float>((1 << bits) - 1) : 1.f;
    offset = mn;
  }
  if (threadIdx.x == 0) { params[2 * g] = scale; params[2 * g + 1] = offset; }
  const float inv = 1.f / scale;
  const int qlo = symmetric ? -static_cast<int>(qmax) - 1 : 0;
  const int qhi = symmetric ? static_cast<int>(qmax) : (1 << bits) - 1;
  if (bits == 8) {
    for (int i = threadIdx.x; i < group_size; i += blockDim.x) {
      int v = __float2int_rn((ldf<T>(xg, i) - offset) * inv);
      v = max(qlo, min(qhi, v));
      q[g * group_size + i] = static_cast<int8_t>(symmetric ? v : v - 128);
    }
  } else {
    for (int i = threadIdx.x; i < group_size / 2; i += blockDim.x) {
      int a = __float2int_rn((ldf<T>(xg, 2 * i) - offset) * inv);
      int b = __float2int_rn((ldf<T>(xg, 2 * i + 1) - offset) * inv);
      a = max(qlo, min(qhi, a));
      b = max(qlo, min(qhi, b));
      if (!symmetric) { a -= 8; b -= 8; }
      q[g * (group_size / 2) + i] = static_cast<int8_t>((a & 0xf) | ((b & 0xf) << 4));
    }
  }
}

template <typename T>
__global__ void int_dequant_kernel(const int8_t* __restrict__ q, const float* __restrict__ params, T* __restrict__ out,
                                   int group_size, int bits, int symmetric) {
  const size_t g = blockIdx.x;
  const float scale = params[2 * g], offset = params[2 * g + 1];
  if (bits == 8) {
    for (int i = threadIdx.x; i < group_size; i += blockDim.x) {
      int v = q[g * group_size + i];
      if (!symmetric) v += 128;
      stf<T>(out, g * group_size + i, v * scale + offset);
    }
  } else {
    for (int i = threadIdx.x; i < group_size / 2; i += blockDim.x) {
      const int byte = static_cast<int>(static_cast<uint8_t>(q[g * (group_size / 2) + i]));
      int a = ((byte & 0xf) ^ 8) - 8;         // sign-extend the low nibble
      int b = (((byte >> 4) & 0xf) ^ 8) - 8;  // sign-extend the high nibble
      if (!symmetric) { a += 8; b += 8; }
      stf<T>(out, g * group_size + 2 * i, a * scale + offset);
      stf<T>(out, g * group_size + 2 * i + 1, b * scale + offset);
    }
  }
}

// ---------------------------------------------------------------- fake quantization (in place), optional stochastic rounding
__device__ __forceinline__ float uniform01(uint64_t seed, uint64_t idx) {
  uint64_t z = seed + idx * 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  z ^= z >> 31;
  return static_cast<float>(z >> 40) * (1.f / 16777216.f);
}
template <typename T>
__global__ void fake_quant_kernel(T* __restrict__ x, int group_size, int bits, int symmetric, int stochastic, uint64_t seed) {
  __shared__ float red[32];
  const size_t g = blockIdx.x;
  T* xg = x + g * group_size;
  float mx = -INFINITY, mn = INFINITY;
  for (int i = threadIdx.x; i < group_size; i += blockDim.x) {
    const float v = ldf<T>(xg, i);
    mx = fmaxf(mx, v);
    mn = fminf(mn, v);
  }
  mx = block_reduce(mx, red, true);
  mn = block_reduce(mn, red, false);
  const float levels = static_cast<float>(1 << bits);
  float scale, offset, qlo, qhi;
  if (symmetric) {
    const float amax = fmaxf(fabsf(mx), fabsf(mn));
    scale = amax > 0.f ? 2.f * amax / levels : 1.f;
    offset = 0.f;
    qlo = -levels / 2.f;
    qhi = levels / 2.f - 1.f;
  } else {
    scale = (mx - mn) > 0.f ? (mx - mn) / (levels - 1.f) : 1.f;
    offset = mn;
    qlo = 0.f;
    qhi = levels - 1.f;
  }
  const float inv = 1.f / scale;
  for (int i = threadIdx.x; i < group_size; i += blockDim.x) {
    const float t = (ldf<T>(xg, i) - offset) * inv;
    float r;
    if (stochastic) {
      const float fl = floorf(t);
      r = fl + ((t - fl) > uniform01(seed, g * group_size + i) ? 1.f : 0.f);
    } else {
      r = rintf(t);
    }
    r = fminf(fmaxf(r, qlo), qhi);
    stf<T>(xg, i, r * scale + offset);
  }
}

}  // namespace quant
}  // namespace dsb

using namespace dsb;
using namespace dsb::quant;

#define DSB_Q_DISPATCH(dtype, ...)                                              \
  do {                                                                          \
    if (dtype == 0) { using T = __nv_bfloat16; __VA_ARGS__; }                   \
    else if (dtype == 1) { using T = __half; __VA_ARGS__; }                     \
    else if (dtype == 2) { using T = float; __VA_ARGS__; }                      \
    else { set_last_error("quantizer: bad dtype %d", dtype); return DSB_ERR_ARG; } \
  } while (0)

// x: n elements (n % 128 == 0) -> q: n bytes (e4m3), scales: n/128 floats
DSB_EXPORT int dsb200_fp8_block_quantize(const void* x, void* q, float* scales, long long n, int dtype, cudaStream_t stream) {
  DSB_REQUIRE(n % FP8_BLOCK == 0, "fp8 quantize: n=%lld must be a multiple of %d", n, FP8_BLOCK);
  if (n == 0) return DSB_OK;
  const size_t nb = static_cast<size_t>(n) / FP8_BLOCK;
  const unsigned grid = static_cast<unsigned>((nb * 32 + 255) / 256);
  DSB_Q_DISPATCH(dtype, (fp8_quant_kernel<T><<<grid, 256, 0, stream>>>((const T*)x, (uint8_t*)q, scales, nb)));
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}
DSB_EXPORT int dsb200_fp8_block_dequantize(const void* q, const float* scales, void* out, long long n, int dtype,
                                           cudaStream_t stream) {
  DSB_REQUIRE(n % FP8_BLOCK == 0, "fp8 dequantize: n=%lld must be a multiple of %d", n, FP8_BLOCK);
  if (n == 0) return DSB_OK;
  const size_t nb = static_cast<size_t>(n) / FP8_BLOCK;
  const unsigned grid = static_cast<unsigned>((nb * 32 + 255) / 256);
  DSB_Q_DISPATCH(dtype, (fp8_dequant_kernel<T><<<grid, 256, 0, stream>>>((const uint8_t*)q, scales, (T*)out, nb)));
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}
// q: parts * n bytes, scales: parts * n/128 ; out: n elements = scale_out * sum over parts
DSB_EXPORT int dsb200_fp8_block_dequant_reduce(const void* q, const float* scales, void* out, long long n, int parts,
                                               float scale_out, int dtype, cudaStream_t stream) {
  DSB_REQUIRE(n % FP8_BLOCK == 0, "fp8 dequant-reduce: n=%lld must be a multiple of %d", n, FP8_BLOCK);
  if (n == 0) return DSB_OK;
  const size_t nb = static_cast<size_t>(n) / FP8_BLOCK;
  const unsigned grid = static_cast<unsigned>((nb * 32 + 255) / 256);
  DSB_Q_DISPATCH(dtype, (fp8_dequant_reduce_kernel<T><<<grid, 256, 0, stream>>>((const uint8_t*)q, scales, (T*)out, nb, parts, scale_out)));
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}
// groups * group_size elements; params: groups*2 floats (scale, offset); bits 8 or 4; q bytes = n * bits / 8
DSB_EXPORT int dsb200_int_quantize(const void* x, void* q, float* params, long long groups, int group_size, int bits,
                                   int symmetric, int dtype, cudaStream_t stream) {
  DSB_REQUIRE(bits == 8 || bits == 4, "int quantize: bits=%d (8 or 4)", bits);
  DSB_REQUIRE(group_size % 2 == 0, "int quantize: odd group size");
  if (groups == 0) return DSB_OK;
  DSB_Q_DISPATCH(dtype, (int_quant_kernel<T><<<static_cast<unsigned>(groups), 256, 0, stream>>>((const T*)x, (int8_t*)q, params,
                                                                                                group_size, bits, symmetric)));
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}
DSB_EXPORT int dsb200_int_dequantize(const void* q, const float* params, void* out, long long groups, int group_size, int bits,
                                     int symmetric, int dtype, cudaStream_t stream) {
  DSB_REQUIRE(bits == 8 || bits == 4, "int dequantize: bits=%d (8 or 4)", bits);
  if (groups == 0) return DSB_OK;
  DSB_Q_DISPATCH(dtype, (int_dequant_kernel<T><<<static_cast<unsigned>(groups), 256, 0, stream>>>((const int8_t*)q, params, (T*)out,
                                                                                                  group_size, bits, symmetric)));
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}
DSB_EXPORT int dsb200_fake_quantize(void* x, long long groups, int group_size, int bits, int symmetric, int stochastic,
                                    long long seed, int dtype, cudaStream_t stream) {
  if (groups == 0) return DSB_OK;
  DSB_Q_DISPATCH(dtype, (fake_quant_kernel<T><<<static_cast<unsigned>(groups), 256, 0, stream>>>((T*)x, group_size, bits, symmetric,
                                                                                                 stochastic, (uint64_t)seed)));
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

// q[rows][cols] (e4m3) and scales[rows] (fp32) with x ~= q * scale per row; x: [rows][cols] with row stride ldx (elements).
// dtype: 0 bf16, 1 fp16, 2 fp32.  cols must be a multiple of 8.
DSB_EXPORT int dsb200_fp8_rowwise_quantize(const void* x, void* q, float* scales, int rows, int cols, long long ldx, int dtype,
                                           cudaStream_t stream) {
  DSB_REQUIRE(cols % 8 == 0 && rows >= 0 && ldx >= cols, "fp8_rowwise_quantize: cols must be a multiple of 8");
  if (rows == 0) return DSB_OK;
  const unsigned grid = static_cast<unsigned>((static_cast<long long>(rows) * 32 + 255) / 256);
  uint8_t* qq = static_cast<uint8_t*>(q);
  if (dtype == 0) quant::fp8_rowwise_quant_kernel<__nv_bfloat16><<<grid, 256, 0, stream>>>(static_cast<const __nv_bfloat16*>(x), qq, scales, rows, cols, ldx);
  else if (dtype == 1) quant::fp8_rowwise_quant_kernel<__half><<<grid, 256, 0, stream>>>(static_cast<const __half*>(x), qq, scales, rows, cols, ldx);
  else if (dtype == 2) quant::fp8_rowwise_quant_kernel<float><<<grid, 256, 0, stream>>>(static_cast<const float*>(x), qq, scales, rows, cols, ldx);
  else {
    set_last_error("fp8_rowwise_quantize: unsupported dtype %d", dtype);
    return DSB_ERR_UNSUPPORTED;
  }
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

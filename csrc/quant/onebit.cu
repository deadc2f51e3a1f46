// 1-bit (sign + one scale) compression with error feedback for the 1-bit Adam / 1-bit LAMB / 0/1 Adam optimizers.
// The reference does this with cupy packbits and a chain of ~10 torch elementwise ops per call
// (deepspeed/runtime/comm/nccl.py:52-170, deepspeed/runtime/compression/cupy.py); here each direction is two launches:
//   compress  : (1) sum((x+e)^2)  (2) scale = sqrt(sum/n); v = x+e; bit = v>=0; e = v - scale*(+-1); pack 8 signs/byte
//   decompress: signs x per-row scale -> either the server average over rows or the expanded full vector
// Everything is HBM-bound streaming; 8 elements per thread keep loads at 2 x 16 B and the byte stores coalesced.
#include "../common/host.h"

namespace dsb {
namespace onebit {

__global__ void sumsq_kernel(const float* __restrict__ x, const float* __restrict__ e, float* __restrict__ acc, long n) {
  float s = 0.f;
  long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  const long stride = (long)gridDim.x * blockDim.x;
  for (long j = i * 4; j < n; j += stride * 4) {
    float4 a = *reinterpret_cast<const float4*>(x + j);
    float4 b = *reinterpret_cast<const float4*>(e + j);
    float v0 = a.x + b.x, v1 = a.y + b.y, v2 = a.z + b.z, v3 = a.w + b.w;
    s += v0 * v0 + v1 * v1 + v2 * v2 + v3 * v3;
  }
#pragma unroll
  for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  __shared__ float ws[32];
  if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x < 32) {
    s = threadIdx.x < (blockDim.x >> 5) ? ws[threadIdx.x] : 0.f;
#pragma unroll
    for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (threadIdx.x == 0) atomicAdd(acc, s);
  }
}

// scale_io[0] <- scale, scale_io[1] holds the sum of squares
__global__ void pack_kernel(float* __restrict__ x, float* __restrict__ e, unsigned char* __restrict__ packed,
                            float* __restrict__ scale_io, long n) {
  const float scale = sqrtf(scale_io[1] / (float)n);
  long t = (long)blockIdx.x * blockDim.x + threadIdx.x;
  const long stride = (long)gridDim.x * blockDim.x;
  for (long b = t; b * 8 < n; b += stride) {
    float4 a0 = *reinterpret_cast<const float4*>(x + b * 8), a1 = *reinterpret_cast<const float4*>(x + b * 8 + 4);
    float4 e0 = *reinterpret_cast<const float4*>(e + b * 8), e1 = *reinterpret_cast<const float4*>(e + b * 8 + 4);
    float v[8] = {a0.x + e0.x, a0.y + e0.y, a0.z + e0.z, a0.w + e0.w, a1.x + e1.x, a1.y + e1.y, a1.z + e1.z, a1.w + e1.w};
    unsigned bits = 0;
    float r[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const bool pos = v[k] >= 0.f;  // sign(0) counts as +1, as (sign+1).bool() does after the upstream add_(1)
      bits |= (pos ? 1u : 0u) << (7 - k);
      r[k] = v[k] - (pos ? scale : -scale);
    }
    *reinterpret_cast<float4*>(e + b * 8) = make_float4(r[0], r[1], r[2], r[3]);
    *reinterpret_cast<float4*>(e + b * 8 + 4) = make_float4(r[4], r[5], r[6], r[7]);
    packed[b] = (unsigned char)bits;
  }
  if (t == 0) scale_io[0] = scale;
}

// mode 0: out[c*8+k] = (1/rows) * sum_r sign(r,c,k) * scales[r]          (server-side average of the workers' chunks)
// mode 1: out[(r*chunk_bytes + c)*8+k] = sign(r,c,k) * scales[r]         (expand the gathered server chunks)
__global__ void unpack_kernel(const unsigned char* __restrict__ packed, const float* __restrict__ scales,
                              float* __restrict__ out, int rows, long chunk_bytes, int mode) {
  long t = (long)blockIdx.x * blockDim.x + threadIdx.x;
  const long stride = (long)gridDim.x * blockDim.x;
  if (mode == 0) {
    const float inv = 1.f / (float)rows;
    for (long c = t; c < chunk_bytes; c += stride) {
      float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
      for (int r = 0; r < rows; ++r) {
        const unsigned bits = packed[(long)r * chunk_bytes + c];
        const float s = scales[r] * inv;
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[k] += ((bits >> (7 - k)) & 1u) ? s : -s;
      }
      *reinterpret_cast<float4*>(out + c * 8) = make_float4(acc[0], acc[1], acc[2], acc[3]);
      *reinterpret_cast<float4*>(out + c * 8 + 4) = make_float4(acc[4], acc[5], acc[6], acc[7]);
    }
  } else {
    const long total = (long)rows * chunk_bytes;
    for (long c = t; c < total; c += stride) {
      const unsigned bits = packed[c];
      const float s = scales[c / chunk_bytes];
      float v[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) v[k] = ((bits >> (7 - k)) & 1u) ? s : -s;
      *reinterpret_cast<float4*>(out + c * 8) = make_float4(v[0], v[1], v[2], v[3]);
      *reinterpret_cast<float4*>(out + c * 8 + 4) = make_float4(v[4], v[5], v[6], v[7]);
    }
  }
}

}  // namespace onebit
}  // namespace dsb

// x[n] fp32 (read only), err[n] fp32 (error feedback, updated), packed[n/8], scale_io[2] fp32 device scratch:
// on return scale_io[0] = ||x+err||_2 / sqrt(n).  n must be a multiple of 8; x/err 16-byte aligned.
DSB_EXPORT int dsb200_onebit_compress(void* x, void* err, void* packed, void* scale_io, long long n, cudaStream_t stream) {
  using namespace dsb;
  DSB_REQUIRE(n > 0 && n % 8 == 0, "onebit_compress: n=%lld must be a positive multiple of 8", n);
  DSB_REQUIRE(((uintptr_t)x & 15) == 0 && ((uintptr_t)err & 15) == 0, "onebit_compress: 16-byte alignment required");
  DSB_CHECK_CUDA(cudaMemsetAsync(static_cast<float*>(scale_io) + 1, 0, sizeof(float), stream));
  const int threads = 256;
  long want = (n / 4 + threads - 1) / threads;
  int grid = (int)(want < (long)num_sms() * 8 ? want : (long)num_sms() * 8);
  onebit::sumsq_kernel<<<grid, threads, 0, stream>>>(static_cast<const float*>(x), static_cast<const float*>(err),
                                                     static_cast<float*>(scale_io) + 1, n);
  want = (n / 8 + threads - 1) / threads;
  grid = (int)(want < (long)num_sms() * 8 ? want : (long)num_sms() * 8);
  onebit::pack_kernel<<<grid, threads, 0, stream>>>(static_cast<float*>(x), static_cast<float*>(err),
                                                    static_cast<unsigned char*>(packed), static_cast<float*>(scale_io), n);
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

DSB_EXPORT int dsb200_onebit_decompress(const void* packed, const void* scales, void* out, int rows, long long chunk_bytes,
                                        int mode, cudaStream_t stream) {
  using namespace dsb;
  DSB_REQUIRE(rows > 0 && chunk_bytes > 0 && (mode == 0 || mode == 1), "onebit_decompress: bad arguments");
  DSB_REQUIRE(((uintptr_t)out & 15) == 0, "onebit_decompress: out must be 16-byte aligned");
  const int threads = 256;
  const long work = mode == 0 ? chunk_bytes : chunk_bytes * rows;
  long want = (work + threads - 1) / threads;
  int grid = (int)(want < (long)num_sms() * 8 ? want : (long)num_sms() * 8);
  onebit::unpack_kernel<<<grid, threads, 0, stream>>>(static_cast<const unsigned char*>(packed),
                                                      static_cast<const float*>(scales), static_cast<float*>(out), rows,
                                                      chunk_bytes, mode);
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

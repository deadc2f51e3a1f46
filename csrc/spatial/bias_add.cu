// Channels-last (NHWC) bias-add family used by the diffusers UNet/VAE wrappers:
//   out = act + bias[c]                                  (nhwc_bias_add)
//   out = act + bias[c] + other                          (nhwc_bias_add_add)
//   out = act + bias[c] + other + other_bias[c]          (nhwc_bias_add_bias_add)
// Reference: csrc/spatial/csrc/opt_bias_add.cu (fp16 only, one kernel per variant).  Here one templated kernel handles
// fp16 / bf16 / fp32 with 16-byte vectors; the channel index of a vector is (flat element index % channels), so any
// tensor whose innermost dimension is the channel works ([N,H,W,C] or [tokens, C]).
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include "../common/host.h"

namespace dsb {
namespace spatial {

template <typename T>
struct Cvt;
template <>
struct Cvt<float> {
  static __device__ float to(float v) { return v; }
  static __device__ float from(float v) { return v; }
};
template <>
struct Cvt<__half> {
  static __device__ float to(__half v) { return __half2float(v); }
  static __device__ __half from(float v) { return __float2half(v); }
};
template <>
struct Cvt<__nv_bfloat16> {
  static __device__ float to(__nv_bfloat16 v) { return __bfloat162float(v); }
  static __device__ __nv_bfloat16 from(float v) { return __float2bfloat16(v); }
};

template <typename T, int VEC>
__global__ void bias_add_kernel(const T* __restrict__ act, const T* __restrict__ bias, const T* __restrict__ other,
                                const T* __restrict__ other_bias, T* __restrict__ out, long long total, int channels) {
  const long long nvec = total / VEC;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long v = (long long)blockIdx.x * blockDim.x + threadIdx.x; v < nvec; v += stride) {
    const long long base = v * VEC;
    const int c0 = (int)(base % channels);  // channels % VEC == 0, so a vector never straddles a row
    T a[VEC], b[VEC], o[VEC], ob[VEC], r[VEC];
    *reinterpret_cast<uint4*>(a) = *reinterpret_cast<const uint4*>(act + base);
    *reinterpret_cast<uint4*>(b) = *reinterpret_cast<const uint4*>(bias + c0);
    if (other) *reinterpret_cast<uint4*>(o) = *reinterpret_cast<const uint4*>(other + base);
    if (other_bias) *reinterpret_cast<uint4*>(ob) = *reinterpret_cast<const uint4*>(other_bias + c0);
#pragma unroll
    for (int k = 0; k < VEC; ++k) {
      float s = Cvt<T>::to(a[k]) + Cvt<T>::to(b[k]);
      if (other) s += Cvt<T>::to(o[k]);
      if (other_bias) s += Cvt<T>::to(ob[k]);
      r[k] = Cvt<T>::from(s);
    }
    *reinterpret_cast<uint4*>(out + base) = *reinterpret_cast<const uint4*>(r);
  }
}

template <typename T>
static int launch(const void* act, const void* bias, const void* other, const void* other_bias, void* out, long long total,
                  int channels, cudaStream_t stream) {
  constexpr int VEC = 16 / sizeof(T);
  const int threads = 256;
  long long want = (total / VEC + threads - 1) / threads;
  const long long cap = (long long)num_sms() * 16;
  const int grid = (int)(want < cap ? (want > 0 ? want : 1) : cap);
  bias_add_kernel<T, VEC><<<grid, threads, 0, stream>>>(static_cast<const T*>(act), static_cast<const T*>(bias),
                                                       static_cast<const T*>(other), static_cast<const T*>(other_bias),
                                                       static_cast<T*>(out), total, channels);
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

}  // namespace spatial
}  // namespace dsb

// dtype: 0 = fp32, 1 = fp16, 2 = bf16.  other / other_bias may be null.  channels * sizeof(T) must be a multiple of 16.
DSB_EXPORT int dsb200_nhwc_bias_add(const void* act, const void* bias, const void* other, const void* other_bias, void* out,
                                    long long total, int channels, int dtype, cudaStream_t stream) {
  using namespace dsb;
  DSB_REQUIRE(total > 0 && channels > 0 && total % channels == 0, "nhwc_bias_add: total=%lld channels=%d", total, channels);
  const int esz = dtype == 0 ? 4 : 2;
  DSB_REQUIRE((channels * esz) % 16 == 0, "nhwc_bias_add: channels*elem_size must be a multiple of 16 bytes");
  DSB_REQUIRE(other_bias == nullptr || other != nullptr, "nhwc_bias_add: other_bias needs other");
  switch (dtype) {
    case 0: return spatial::launch<float>(act, bias, other, other_bias, out, total, channels, stream);
    case 1: return spatial::launch<__half>(act, bias, other, other_bias, out, total, channels, stream);
    case 2: return spatial::launch<__nv_bfloat16>(act, bias, other, other_bias, out, total, channels, stream);
  }
  set_last_error("nhwc_bias_add: bad dtype");
  return DSB_ERR_ARG;
}

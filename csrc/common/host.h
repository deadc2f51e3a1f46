// Host-side helpers shared by the launchers: error codes, TMA tensor-map encoding (driver entry point
// resolved at run time so the library links without libcuda on the build box), device properties.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#define DSB_OK 0
#define DSB_ERR_ARG 1
#define DSB_ERR_CUDA 2
#define DSB_ERR_UNSUPPORTED 3

#define DSB_EXPORT extern "C" __attribute__((visibility("default")))

namespace dsb {

void set_last_error(const char* fmt, ...);

#define DSB_CHECK_CUDA(expr)                                                                   \
  do {                                                                                         \
    cudaError_t _e = (expr);                                                                   \
    if (_e != cudaSuccess) {                                                                   \
      ::dsb::set_last_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
      return DSB_ERR_CUDA;                                                                     \
    }                                                                                          \
  } while (0)

#define DSB_REQUIRE(cond, ...)              \
  do {                                      \
    if (!(cond)) {                          \
      ::dsb::set_last_error(__VA_ARGS__);   \
      return DSB_ERR_ARG;                   \
    }                                       \
  } while (0)

int num_sms();

enum TmaDtype { TMA_BF16 = 0, TMA_F32 = 1, TMA_U8 = 2, TMA_F16 = 3 };
enum TmaSwizzle { TMA_SW_NONE = 0, TMA_SW_128 = 3 };

// Encodes a rank-`rank` tiled tensor map. dims/box innermost first; strides in BYTES for dims 1..rank-1.
// Returns 0 on success.
int make_tensor_map(CUtensorMap* out, TmaDtype dt, int rank, const void* base, const uint64_t* dims,
                    const uint64_t* strides_bytes, const uint32_t* box, TmaSwizzle sw);

// Convenience: row-major 2-D matrix [rows, cols] with leading dimension ld (elements), box [box_rows, box_cols].
int make_tensor_map_2d(CUtensorMap* out, TmaDtype dt, const void* base, uint64_t rows, uint64_t cols, uint64_t ld,
                       uint32_t box_rows, uint32_t box_cols, TmaSwizzle sw);

}  // namespace dsb

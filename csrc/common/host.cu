#include "host.h"

#include <stdarg.h>
#include <string.h>

#include <mutex>

namespace dsb {

static thread_local char g_last_error[512] = "";

void set_last_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_last_error, sizeof(g_last_error), fmt, ap);
  va_end(ap);
}

int num_sms() {
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cached[dev] = n;
  }
  return cached[dev];
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

int make_tensor_map(CUtensorMap* out, TmaDtype dt, int rank, const void* base, const uint64_t* dims,
                    const uint64_t* strides_bytes, const uint32_t* box, TmaSwizzle sw) {
  EncodeTiledFn enc = get_encode();
  if (!enc) {
    set_last_error("cuTensorMapEncodeTiled entry point unavailable");
    return DSB_ERR_CUDA;
  }
  CUtensorMapDataType cdt = CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  if (dt == TMA_F32) cdt = CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
  if (dt == TMA_U8) cdt = CU_TENSOR_MAP_DATA_TYPE_UINT8;
  if (dt == TMA_F16) cdt = CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
  cuuint64_t gdims[5];
  cuuint64_t gstr[4];
  cuuint32_t gbox[5];
  cuuint32_t estr[5];
  for (int i = 0; i < rank; ++i) {
    gdims[i] = dims[i];
    gbox[i] = box[i];
    estr[i] = 1;
    if (i > 0) gstr[i - 1] = strides_bytes[i - 1];
  }
  auto encode = [&] {
    return enc(out, cdt, static_cast<cuuint32_t>(rank), const_cast<void*>(base), gdims, gstr, gbox, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
               sw == TMA_SW_128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  };
  CUresult r = encode();
  if (r == CUDA_ERROR_INVALID_CONTEXT) {
    // a thread that has made no runtime call yet (an autograd worker running its first backward) has no current context:
    // bind the primary context of the device that owns `base`, then retry
    cudaPointerAttributes attr;
    if (cudaPointerGetAttributes(&attr, base) == cudaSuccess && attr.type == cudaMemoryTypeDevice) cudaSetDevice(attr.device);
    cudaFree(nullptr);
    (void)cudaGetLastError();
    r = encode();
  }
  if (r != CUDA_SUCCESS) {
    set_last_error("cuTensorMapEncodeTiled failed (%d): rank=%d base=%p dims=[%llu,%llu,..] stride1=%llu box=[%u,%u]",
                   (int)r, rank, base, (unsigned long long)dims[0], (unsigned long long)(rank > 1 ? dims[1] : 0),
                   (unsigned long long)(rank > 1 ? strides_bytes[0] : 0), box[0], rank > 1 ? box[1] : 0);
    return DSB_ERR_CUDA;
  }
  return DSB_OK;
}

int make_tensor_map_2d(CUtensorMap* out, TmaDtype dt, const void* base, uint64_t rows, uint64_t cols, uint64_t ld,
                       uint32_t box_rows, uint32_t box_cols, TmaSwizzle sw) {
  uint64_t esz = (dt == TMA_F32) ? 4 : (dt == TMA_U8 ? 1 : 2);
  uint64_t dims[2] = {cols, rows};
  uint64_t strides[1] = {ld * esz};
  uint32_t box[2] = {box_cols, box_rows};
  return make_tensor_map(out, dt, 2, base, dims, strides, box, sw);
}

}  // namespace dsb

DSB_EXPORT const char* dsb200_last_error() { return dsb::g_last_error; }

DSB_EXPORT int dsb200_version() { return 100; }

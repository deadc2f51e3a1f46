// bf16 GEMM for sm_100a: TMA -> 128B-swizzled smem ring -> tcgen05.mma (single issuing thread) -> fp32 accumulators in
// TMEM (double buffered, 2 x BN columns) -> tcgen05.ld epilogue -> swizzled smem staging -> TMA store / TMA reduce-add.
// Persistent grid (one CTA per SM), warp-specialised: warp 0 = TMA producer, warp 1 = MMA issuer + TMEM owner,
// warps 2..5 = epilogue (one TMEM lane quarter each).
//
// Operand layouts (all row-major in global memory):
//   A "K-major"  : A[M][K]   (activations X, dY for dgrad)        A "MN-major": stored as [K][M] (dY^T for wgrad)
//   B "K-major"  : B[N][K]   (weights W[out][in] for forward)     B "MN-major": stored as [K][N] (W for dgrad, X for wgrad)
// so forward  Y = X W^T        -> (A K-major, B K-major)
//    dgrad    dX = dY W        -> (A K-major, B MN-major)
//    wgrad    dW = dY^T X      -> (A MN-major, B MN-major)        (accumulates straight into fp32 main grads
//                                                                   with TMA reduce-add: out_mode 2)
// Replaces the cublasGemmEx call sites of the reference (csrc/transformer/cublas_wrappers.cu:62,153;
// deepspeed/inference/v2/kernels/core_ops/blas_kernels/blas.h:35-138) and F.linear under ZeRO-3
// (deepspeed/runtime/zero/linear.py:43-112).
#include "gemm_2cta.cuh"

namespace dsb {
namespace gemm {

template <int BN, int STAGES, bool A_MN, bool B_MN, bool OUT_F32>
static int launch(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb, int ldc,
                  int reduce_add, int max_ctas, cudaStream_t stream) {
  using S = Smem<BN, STAGES>;
  CUtensorMap tmA, tmB, tmC;
  int rc;
  if (!A_MN)
    rc = make_tensor_map_2d(&tmA, TMA_BF16, A, M, K, lda, BM, BK, TMA_SW_128);
  else
    rc = make_tensor_map_2d(&tmA, TMA_BF16, A, K, M, lda, BK, 64, TMA_SW_128);
  if (rc) return rc;
  if (!B_MN)
    rc = make_tensor_map_2d(&tmB, TMA_BF16, B, N, K, ldb, BN, BK, TMA_SW_128);
  else
    rc = make_tensor_map_2d(&tmB, TMA_BF16, B, K, N, ldb, BK, 64, TMA_SW_128);
  if (rc) return rc;
  rc = make_tensor_map_2d(&tmC, OUT_F32 ? TMA_F32 : TMA_BF16, C, M, N, ldc, BM, OUT_F32 ? 32 : 64, TMA_SW_128);
  if (rc) return rc;

  auto kern = gemm_bf16_kernel<BN, STAGES, A_MN, B_MN, OUT_F32>;
  static bool attr_set = false;
  if (!attr_set) {
    DSB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, S::TOTAL));
    attr_set = true;
  }
  Params p;
  p.M = M;
  p.N = N;
  p.K = K;
  p.tiles_m = (M + BM - 1) / BM;
  p.tiles_n = (N + BN - 1) / BN;
  p.reduce_add = reduce_add;
  int grid = p.tiles_m * p.tiles_n;
  int cap = num_sms();
  if (max_ctas > 0 && max_ctas < cap) cap = max_ctas;
  if (grid > cap) grid = cap;
  kern<<<grid, NUM_THREADS, S::TOTAL, stream>>>(tmA, tmB, tmC, p);
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

// CTA-pair kernel (256 x 256 tiles, see gemm_2cta.cuh)
template <bool A_MN, bool B_MN, bool OUT_F32>
static int launch2(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb, int ldc, int reduce_add,
                   int max_ctas, cudaStream_t stream) {
  constexpr int STAGES = 6;
  using S = Smem2<STAGES>;
  CUtensorMap tmA, tmB, tmC;
  int rc;
  if (!A_MN)
    rc = make_tensor_map_2d(&tmA, TMA_BF16, A, M, K, lda, BM, BK, TMA_SW_128);
  else
    rc = make_tensor_map_2d(&tmA, TMA_BF16, A, K, M, lda, BK, 64, TMA_SW_128);
  if (rc) return rc;
  if (!B_MN)
    rc = make_tensor_map_2d(&tmB, TMA_BF16, B, N, K, ldb, BN2_CTA, BK, TMA_SW_128);
  else
    rc = make_tensor_map_2d(&tmB, TMA_BF16, B, K, N, ldb, BK, 64, TMA_SW_128);
  if (rc) return rc;
  rc = make_tensor_map_2d(&tmC, OUT_F32 ? TMA_F32 : TMA_BF16, C, M, N, ldc, BM, OUT_F32 ? 32 : 64, TMA_SW_128);
  if (rc) return rc;
  auto kern = gemm2_bf16_kernel<STAGES, A_MN, B_MN, OUT_F32>;
  static bool attr_set = false;
  if (!attr_set) {
    DSB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, S::TOTAL));
    attr_set = true;
  }
  Params p;
  p.M = M;
  p.N = N;
  p.K = K;
  p.tiles_m = (M + BM2 - 1) / BM2;
  p.tiles_n = (N + BN2 - 1) / BN2;
  p.reduce_add = reduce_add;
  int clusters = p.tiles_m * p.tiles_n;
  int cap = num_sms();
  if (max_ctas > 0 && max_ctas < cap) cap = max_ctas;
  if (clusters > cap / 2) clusters = cap / 2;
  if (clusters < 1) clusters = 1;
  kern<<<2 * clusters, NUM_THREADS, S::TOTAL, stream>>>(tmA, tmB, tmC, p);  // cluster size is a kernel attribute
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

static int dispatch2(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb, int ldc, int a_mn,
                     int b_mn, int out_mode, int max_ctas, cudaStream_t stream) {
  const bool f32 = (out_mode == 1 || out_mode == 2);
  const int red = (out_mode == 2 || out_mode == 3) ? 1 : 0;
#define DSB_GEMM2_CASE(AM, BMj)                                                                      \
  if (a_mn == AM && b_mn == BMj) {                                                                   \
    if (f32) return launch2<AM, BMj, true>(A, B, C, M, N, K, lda, ldb, ldc, red, max_ctas, stream);  \
    return launch2<AM, BMj, false>(A, B, C, M, N, K, lda, ldb, ldc, red, max_ctas, stream);          \
  }
  DSB_GEMM2_CASE(0, 0)
  DSB_GEMM2_CASE(0, 1)
  DSB_GEMM2_CASE(1, 0)
  DSB_GEMM2_CASE(1, 1)
#undef DSB_GEMM2_CASE
  set_last_error("gemm: bad layout flags");
  return DSB_ERR_ARG;
}

template <int BN, int STAGES>
static int dispatch(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb, int ldc, int a_mn,
                    int b_mn, int out_mode, int max_ctas, cudaStream_t stream) {
  const bool f32 = (out_mode == 1 || out_mode == 2);
  const int red = (out_mode == 2 || out_mode == 3) ? 1 : 0;
#define DSB_GEMM_CASE(AM, BMj)                                                                             \
  if (a_mn == AM && b_mn == BMj) {                                                                         \
    if (f32) return launch<BN, STAGES, AM, BMj, true>(A, B, C, M, N, K, lda, ldb, ldc, red, max_ctas, stream); \
    return launch<BN, STAGES, AM, BMj, false>(A, B, C, M, N, K, lda, ldb, ldc, red, max_ctas, stream);         \
  }
  DSB_GEMM_CASE(0, 0)
  DSB_GEMM_CASE(0, 1)
  DSB_GEMM_CASE(1, 0)
  DSB_GEMM_CASE(1, 1)
#undef DSB_GEMM_CASE
  set_last_error("gemm: bad layout flags");
  return DSB_ERR_ARG;
}

}  // namespace gemm
}  // namespace dsb

// C[M,N] (=|+=) A * B, bf16 inputs, fp32 accumulation on tcgen05 tensor cores.
//   a_mn_major = 0: A is [M][K] (lda = row stride) ; 1: A is stored [K][M]
//   b_mn_major = 0: B is [N][K] (ldb = row stride) ; 1: B is stored [K][N]
//   out_mode   = 0: bf16 store, 1: fp32 store, 2: fp32 accumulate (C += AB), 3: bf16 accumulate
//   max_ctas   = 0: use every SM ; >0 caps the persistent grid (leaves SMs for concurrently running copy kernels)
DSB_EXPORT int dsb200_gemm_bf16(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb, int ldc,
                                int a_mn_major, int b_mn_major, int out_mode, int max_ctas, cudaStream_t stream) {
  using namespace dsb;
  DSB_REQUIRE(M > 0 && N > 0 && K > 0, "gemm: empty problem M=%d N=%d K=%d", M, N, K);
  DSB_REQUIRE((reinterpret_cast<uintptr_t>(A) & 15) == 0 && (reinterpret_cast<uintptr_t>(B) & 15) == 0 &&
                  (reinterpret_cast<uintptr_t>(C) & 15) == 0,
              "gemm: operands must be 16-byte aligned");
  DSB_REQUIRE(lda % 8 == 0 && ldb % 8 == 0, "gemm: lda/ldb must be multiples of 8 elements (TMA 16-byte strides)");
  const int c_esz = (out_mode == 1 || out_mode == 2) ? 4 : 2;
  DSB_REQUIRE((ldc * c_esz) % 16 == 0, "gemm: ldc*elem must be a multiple of 16 bytes");
  DSB_REQUIRE(out_mode >= 0 && out_mode <= 3, "gemm: out_mode %d", out_mode);
  // variant: 0 = auto, 1 = force single-CTA tiles, 2 = force the CTA-pair kernel (env DSB200_GEMM_VARIANT, read once)
  static const int variant = [] {
    const char* e = getenv("DSB200_GEMM_VARIANT");
    return e ? atoi(e) : 0;
  }();
  const long long tiles2 = static_cast<long long>((M + 255) / 256) * ((N + 255) / 256);
  const bool pair_ok = M > 128 && N > 128;
  if (pair_ok && (variant == 2 || (variant == 0 && tiles2 >= 37)))
    return gemm::dispatch2(A, B, C, M, N, K, lda, ldb, ldc, a_mn_major, b_mn_major, out_mode, max_ctas, stream);
  if (N <= 128) return gemm::dispatch<128, 6>(A, B, C, M, N, K, lda, ldb, ldc, a_mn_major, b_mn_major, out_mode, max_ctas, stream);
  return gemm::dispatch<256, 4>(A, B, C, M, N, K, lda, ldb, ldc, a_mn_major, b_mn_major, out_mode, max_ctas, stream);
}

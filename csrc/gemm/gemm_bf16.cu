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
#include "../common/host.h"
#include "../common/ptx.cuh"

namespace dsb {
namespace gemm {

constexpr int BM = 128;
constexpr int BK = 64;  // 64 x bf16 = one 128-byte swizzle row
constexpr int UMMA_K = 16;
constexpr int NUM_THREADS = 192;
constexpr int EPI_THREADS = 128;
constexpr int EPI_BUF_BYTES = 128 * 128;  // 128 rows x 128 B
constexpr int GROUP_M = 8;

template <int BN, int STAGES>
struct Smem {
  static constexpr int A_BYTES = BM * BK * 2;
  static constexpr int B_BYTES = BN * BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int BAR_OFFSET = STAGES * STAGE_BYTES + 2 * EPI_BUF_BYTES;
  static constexpr int TOTAL = BAR_OFFSET + 256 + 1024;  // + barriers + 1024-B alignment slack
};

struct Params {
  int M, N, K;
  int tiles_m, tiles_n;
  int reduce_add;  // 0: C = AB ; 1: C += AB (TMA reduce-add)
};

__device__ __forceinline__ void tile_coords(int tile, int tiles_m, int tiles_n, int& m_blk, int& n_blk) {
  const int group = GROUP_M * tiles_n;
  const int gid = tile / group;
  const int first_m = gid * GROUP_M;
  const int gsz = min(tiles_m - first_m, GROUP_M);
  const int r = tile - gid * group;
  m_blk = first_m + r % gsz;
  n_blk = r / gsz;
}

template <int BN, int STAGES, bool A_MN, bool B_MN, bool OUT_F32>
__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                 const __grid_constant__ CUtensorMap tmC, const Params p) {
  using S = Smem<BN, STAGES>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* epi_buf = smem + STAGES * S::STAGE_BYTES;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + S::BAR_OFFSET);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full = empty_bar + STAGES;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int num_tiles = p.tiles_m * p.tiles_n;
  const int num_kb = (p.K + BK - 1) / BK;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    tma_prefetch_desc(&tmC);
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tmem_full[i], 1);
      mbar_init(&tmem_empty[i], EPI_THREADS);
    }
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc<1>(tmem_ptr, 2 * BN);
    tmem_relinquish<1>();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  if (warp == 0) {
    // ------------------------------------------------------------- TMA producer
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        int m_blk, n_blk;
        tile_coords(tile, p.tiles_m, p.tiles_n, m_blk, n_blk);
        const int m0 = m_blk * BM, n0 = n_blk * BN;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * S::STAGE_BYTES;
          uint8_t* sb = sa + S::A_BYTES;
          mbar_expect_tx(&full_bar[stage], S::STAGE_BYTES);
          const int k0 = kb * BK;
          if constexpr (!A_MN) {
            tma_load_2d(&tmA, &full_bar[stage], sa, k0, m0);
          } else {
#pragma unroll
            for (int c = 0; c < BM / 64; ++c) tma_load_2d(&tmA, &full_bar[stage], sa + c * (BK * 128), m0 + c * 64, k0);
          }
          if constexpr (!B_MN) {
            tma_load_2d(&tmB, &full_bar[stage], sb, k0, n0);
          } else {
#pragma unroll
            for (int c = 0; c < BN / 64; ++c) tma_load_2d(&tmB, &full_bar[stage], sb + c * (BK * 128), n0 + c * 64, k0);
          }
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------- MMA issuer
    if (lane == 0) {
      constexpr uint32_t idesc = umma_idesc(BM, BN, /*bf16*/ 1, 1, A_MN ? 1 : 0, B_MN ? 1 : 0);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + stage * S::STAGE_BYTES);
          const uint32_t sb = sa + S::A_BYTES;
#pragma unroll
          for (int kk = 0; kk < BK / UMMA_K; ++kk) {
            const uint64_t adesc = A_MN ? umma_smem_desc(sa + kk * (UMMA_K * 128), BK * 128, 1024)
                                        : umma_smem_desc(sa + kk * (UMMA_K * 2), 16, 1024);
            const uint64_t bdesc = B_MN ? umma_smem_desc(sb + kk * (UMMA_K * 128), BK * 128, 1024)
                                        : umma_smem_desc(sb + kk * (UMMA_K * 2), 16, 1024);
            umma_f16<1>(d_tmem, adesc, bdesc, idesc, (kb | kk) != 0 ? 1u : 0u);
          }
          umma_commit<1>(&empty_bar[stage]);  // frees the smem slot when these MMAs retire
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
        umma_commit<1>(&tmem_full[acc]);  // accumulator complete -> epilogue
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1;
      }
    }
  } else {
    // ------------------------------------------------------------- epilogue (warps 2..5)
    const int q = warp & 3;                 // TMEM lane quarter this warp may read
    const int row = q * 32 + lane;          // row of the 128-row tile held by this thread
    const int epi_tid = threadIdx.x - 64;   // 0..127
    constexpr int CW = OUT_F32 ? 32 : 64;   // columns per 128-byte staging row
    constexpr int NCHUNK = BN / CW;
    int acc = 0;
    uint32_t acc_phase = 0;
    uint32_t buf_idx = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      int m_blk, n_blk;
      tile_coords(tile, p.tiles_m, p.tiles_n, m_blk, n_blk);
      const int m0 = m_blk * BM, n0 = n_blk * BN;
      mbar_wait(&tmem_full[acc], acc_phase);
      tc_fence_after();
      const uint32_t t_acc = tmem_base + acc * BN + (static_cast<uint32_t>(q * 32) << 16);
#pragma unroll 1
      for (int c = 0; c < NCHUNK; ++c) {
        const bool live = (n0 + c * CW) < p.N;  // uniform across the CTA
        uint32_t v[32];
        uint32_t w[32];
        if (live) {
          tmem_ld_32x32(t_acc + c * CW, v);
          if constexpr (!OUT_F32) tmem_ld_32x32(t_acc + c * CW + 32, w);
          tmem_ld_wait();
        }
        if (c == NCHUNK - 1) {
          tc_fence_before();
          mbar_arrive(&tmem_empty[acc]);  // accumulator drained: MMA may overwrite it
        }
        if (!live) continue;
        uint8_t* buf = epi_buf + (buf_idx & 1) * EPI_BUF_BYTES;
        if (epi_tid == 0) tma_wait_group_read<1>();  // the store that last used this buffer has read it
        named_bar_sync(1, EPI_THREADS);
        uint8_t* rowp = buf + row * 128;
        const int sw = row & 7;
        if constexpr (OUT_F32) {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            uint4 val = make_uint4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
            *reinterpret_cast<uint4*>(rowp + ((j ^ sw) << 4)) = val;
          }
        } else {
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            uint4 val;
            val.x = pack_bf16x2(__uint_as_float(v[8 * j + 0]), __uint_as_float(v[8 * j + 1]));
            val.y = pack_bf16x2(__uint_as_float(v[8 * j + 2]), __uint_as_float(v[8 * j + 3]));
            val.z = pack_bf16x2(__uint_as_float(v[8 * j + 4]), __uint_as_float(v[8 * j + 5]));
            val.w = pack_bf16x2(__uint_as_float(v[8 * j + 6]), __uint_as_float(v[8 * j + 7]));
            *reinterpret_cast<uint4*>(rowp + ((j ^ sw) << 4)) = val;
          }
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            uint4 val;
            val.x = pack_bf16x2(__uint_as_float(w[8 * j + 0]), __uint_as_float(w[8 * j + 1]));
            val.y = pack_bf16x2(__uint_as_float(w[8 * j + 2]), __uint_as_float(w[8 * j + 3]));
            val.z = pack_bf16x2(__uint_as_float(w[8 * j + 4]), __uint_as_float(w[8 * j + 5]));
            val.w = pack_bf16x2(__uint_as_float(w[8 * j + 6]), __uint_as_float(w[8 * j + 7]));
            *reinterpret_cast<uint4*>(rowp + (((j + 4) ^ sw) << 4)) = val;
          }
        }
        fence_proxy_async_smem();
        named_bar_sync(1, EPI_THREADS);
        if (epi_tid == 0) {
          if (p.reduce_add)
            tma_reduce_add_2d(&tmC, buf, n0 + c * CW, m0);
          else
            tma_store_2d(&tmC, buf, n0 + c * CW, m0);
          tma_commit_group();
        }
        ++buf_idx;
      }
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1;
    }
    if (epi_tid == 0) tma_wait_group<0>();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<1>(tmem_base, 2 * BN);
  }
}

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
  if (N <= 128) return gemm::dispatch<128, 6>(A, B, C, M, N, K, lda, ldb, ldc, a_mn_major, b_mn_major, out_mode, max_ctas, stream);
  return gemm::dispatch<256, 4>(A, B, C, M, N, K, lda, ldb, ldc, a_mn_major, b_mn_major, out_mode, max_ctas, stream);
}

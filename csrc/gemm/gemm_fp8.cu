// FP8 (e4m3) GEMM with row / column scales on tcgen05 (kind::f8f6f4):
//     C[m][n] = sa[m] * sb[n] * sum_k A[m][k] * B[n][k]        A [M][K] e4m3, B [N][K] e4m3, C bf16, fp32 accumulation
// i.e. activations quantised per token and weights per output feature (W8A8).  The scales factor out of the sum, so the
// tensor core accumulates raw fp8 products in TMEM and the epilogue applies sa[m] * sb[n] once per element -- no per-k-block
// promotion through registers.  Twice the MMA rate of bf16 (K = 32 per instruction at the same issue cadence) and half
// the operand bytes: one 128-byte swizzle row holds 128 k-elements, so a pipeline stage covers BK = 128.
// Structure is the single-CTA persistent kernel of gemm_core.cuh (TMA producer warp, one MMA-issuing lane of a convergent
// warp, four epilogue warps, double-buffered TMEM accumulators, swizzled staging + TMA store).
// The reference has no fp8 GEMM (its low-precision inference path dequantises int8 weights to fp16 before cuBLAS:
// csrc/transformer/inference/csrc/dequantize.cu + pt_binding.cpp:1560-1650); this is the Blackwell-native counterpart.
#include "gemm_core.cuh"

namespace dsb {
namespace gemm8 {

constexpr int BM = 128;
constexpr int BN = 256;
constexpr int BKB = 128;  // bytes (= fp8 elements) of K per stage row
constexpr int STAGES = 4;
constexpr int A_BYTES = BM * BKB;
constexpr int B_BYTES = BN * BKB;
constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
constexpr int EPI_BUF = 128 * 128;
constexpr int BAR_OFFSET = STAGES * STAGE_BYTES + 2 * EPI_BUF;
constexpr int SMEM_TOTAL = BAR_OFFSET + 256 + 1024;
constexpr int NUM_THREADS = 192;
constexpr int EPI_THREADS = 128;

struct Params {
  int M, N, K;
  int tiles_m, tiles_n;
  const float* sa;  // [M] or nullptr (= 1)
  const float* sb;  // [N] or nullptr (= 1)
};

__device__ __forceinline__ void wait_lean(uint64_t* bar, uint32_t parity) { mbar_wait_lean(bar, parity); }

__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_fp8_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                const __grid_constant__ CUtensorMap tmC, const Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* epi_buf = smem + STAGES * STAGE_BYTES;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + BAR_OFFSET);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full = empty_bar + STAGES;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int num_tiles = p.tiles_m * p.tiles_n;
  const int num_kb = (p.K + BKB - 1) / BKB;

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
        gemm::tile_coords(tile, p.tiles_m, p.tiles_n, m_blk, n_blk);
        for (int kb = 0; kb < num_kb; ++kb) {
          wait_lean(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * STAGE_BYTES;
          mbar_expect_tx(&full_bar[stage], STAGE_BYTES);
          tma_load_2d(&tmA, &full_bar[stage], sa, kb * BKB, m_blk * BM);
          tma_load_2d(&tmB, &full_bar[stage], sa + A_BYTES, kb * BKB, n_blk * BN);
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------- MMA issuer (convergent warp, one elected lane)
    constexpr uint32_t idesc = umma_idesc(BM, BN, /*e4m3*/ 0, 0, 0, 0);
    constexpr uint32_t HI = (1024u >> 4) | (1u << 14) | (2u << 29);
    constexpr uint32_t KM = (16u >> 4) << 16;
    const uint32_t tm = __shfl_sync(0xffffffffu, tmem_base, 0);
    const bool leader = elect_one();
    const uint32_t base = smem_u32(smem);
    int stage = 0;
    uint32_t phase = 0;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      wait_lean(&tmem_empty[acc], acc_phase ^ 1);
      tc_fence_after();
      for (int kb = 0; kb < num_kb; ++kb) {
        wait_lean(&full_bar[stage], phase);
        tc_fence_after();
        if (leader) {
          const uint32_t a_lo = KM | ((base + stage * STAGE_BYTES) >> 4);
          const uint32_t b_lo = KM | ((base + stage * STAGE_BYTES + A_BYTES) >> 4);
#pragma unroll
          for (int kk = 0; kk < BKB / 32; ++kk)  // 32 fp8 elements (= 32 bytes) of K per instruction
            umma_f8(tm + acc * BN, (static_cast<uint64_t>(HI) << 32) | (a_lo + kk * 2), (static_cast<uint64_t>(HI) << 32) | (b_lo + kk * 2),
                    idesc, (kb | kk) != 0 ? 1u : 0u);
          umma_commit<1>(&empty_bar[stage]);
        }
        if (++stage == STAGES) {
          stage = 0;
          phase ^= 1;
        }
      }
      if (leader) umma_commit<1>(&tmem_full[acc]);
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1;
    }
  } else {
    // ------------------------------------------------------------- epilogue: TMEM -> * sa[m] * sb[n] -> bf16 -> TMA store
    const int q = warp & 3;
    const int row = q * 32 + lane;
    const int epi_tid = threadIdx.x - 64;
    constexpr int CW = 64;
    int acc = 0;
    uint32_t acc_phase = 0, buf_idx = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      int m_blk, n_blk;
      gemm::tile_coords(tile, p.tiles_m, p.tiles_n, m_blk, n_blk);
      const int m0 = m_blk * BM, n0 = n_blk * BN;
      const float sa = (p.sa != nullptr && m0 + row < p.M) ? p.sa[m0 + row] : 1.f;
      wait_lean(&tmem_full[acc], acc_phase);
      tc_fence_after();
      const uint32_t t_acc = tmem_base + acc * BN + (static_cast<uint32_t>(q * 32) << 16);
#pragma unroll 1
      for (int c = 0; c < BN / CW; ++c) {
        const bool live = (n0 + c * CW) < p.N;
        uint32_t v[32], w[32];
        if (live) {
          tmem_ld_32x32(t_acc + c * CW, v);
          tmem_ld_32x32(t_acc + c * CW + 32, w);
          tmem_ld_wait();
        }
        if (c == BN / CW - 1) {
          tc_fence_before();
          mbar_arrive(&tmem_empty[acc]);
        }
        if (!live) continue;
        uint8_t* buf = epi_buf + (buf_idx & 1) * EPI_BUF;
        if (epi_tid == 0) tma_wait_group_read<1>();
        named_bar_sync(1, EPI_THREADS);
        uint8_t* rowp = buf + row * 128;
        const int sw = row & 7;
        const int col0 = n0 + c * CW;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const uint32_t* src = j < 4 ? &v[8 * j] : &w[8 * (j - 4)];
          float o[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const int col = col0 + j * 8 + i;
            const float sb = (p.sb != nullptr && col < p.N) ? __ldg(p.sb + col) : 1.f;
            o[i] = __uint_as_float(src[i]) * sa * sb;
          }
          uint4 val;
          val.x = pack_bf16x2(o[0], o[1]);
          val.y = pack_bf16x2(o[2], o[3]);
          val.z = pack_bf16x2(o[4], o[5]);
          val.w = pack_bf16x2(o[6], o[7]);
          *reinterpret_cast<uint4*>(rowp + ((j ^ sw) << 4)) = val;
        }
        fence_proxy_async_smem();
        named_bar_sync(1, EPI_THREADS);
        if (epi_tid == 0) {
          tma_store_2d(&tmC, buf, col0, m0);
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

}  // namespace gemm8
}  // namespace dsb

// C[M,N] (bf16) = diag(sa) * (A * B^T) * diag(sb) with A [M][K], B [N][K] in e4m3 (one byte per element, K contiguous).
// sa [M] / sb [N] fp32 scales, either may be null (= 1).  lda / ldb in elements (= bytes), multiples of 16; ldc multiple of 8.
DSB_EXPORT int dsb200_gemm_fp8(const void* A, const void* B, void* C, const float* sa, const float* sb, int M, int N, int K, int lda,
                               int ldb, int ldc, int max_ctas, cudaStream_t stream) {
  using namespace dsb;
  DSB_REQUIRE(M > 0 && N > 0 && K > 0, "gemm_fp8: empty problem M=%d N=%d K=%d", M, N, K);
  DSB_REQUIRE((reinterpret_cast<uintptr_t>(A) & 15) == 0 && (reinterpret_cast<uintptr_t>(B) & 15) == 0 &&
                  (reinterpret_cast<uintptr_t>(C) & 15) == 0,
              "gemm_fp8: operands must be 16-byte aligned");
  DSB_REQUIRE(lda % 16 == 0 && ldb % 16 == 0 && ldc % 8 == 0, "gemm_fp8: lda/ldb must be multiples of 16, ldc of 8");
  CUtensorMap tmA, tmB, tmC;
  int rc = make_tensor_map_2d(&tmA, TMA_U8, A, M, K, lda, gemm8::BM, gemm8::BKB, TMA_SW_128);
  if (rc) return rc;
  rc = make_tensor_map_2d(&tmB, TMA_U8, B, N, K, ldb, gemm8::BN, gemm8::BKB, TMA_SW_128);
  if (rc) return rc;
  rc = make_tensor_map_2d(&tmC, TMA_BF16, C, M, N, ldc, gemm8::BM, 64, TMA_SW_128);
  if (rc) return rc;
  static bool attr_set = false;
  if (!attr_set) {
    DSB_CHECK_CUDA(cudaFuncSetAttribute(gemm8::gemm_fp8_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, gemm8::SMEM_TOTAL));
    attr_set = true;
  }
  gemm8::Params p;
  p.M = M;
  p.N = N;
  p.K = K;
  p.tiles_m = (M + gemm8::BM - 1) / gemm8::BM;
  p.tiles_n = (N + gemm8::BN - 1) / gemm8::BN;
  p.sa = sa;
  p.sb = sb;
  int grid = p.tiles_m * p.tiles_n;
  int cap = num_sms();
  if (max_ctas > 0 && max_ctas < cap) cap = max_ctas;
  if (grid > cap) grid = cap;
  gemm8::gemm_fp8_kernel<<<grid, gemm8::NUM_THREADS, gemm8::SMEM_TOTAL, stream>>>(tmA, tmB, tmC, p);
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

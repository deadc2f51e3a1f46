// Block-scaled FP8 GEMM on tcgen05 (kind::f8f6f4) with 1 x 128 scale blocks on BOTH operands:
//     C[m][n] = sum_kb  sa[m][kb] * sb[n][kb] * sum_{k in block kb} A[m][k] * B[n][k]
// A [M][K] e4m3 + sa [M][K/128] fp32, B [N][K] e4m3 + sb [N][K/128] fp32, C bf16.  This is exactly the ZeRO++ wire format of
// csrc/quant/quantizer.cu (fp8_quant_kernel: one fp32 scale per 128 consecutive elements of a row-major tensor whose rows are
// multiples of 128), so a weight shard that arrived as a qwZ payload is a valid B operand as it is -- it is never
// dequantised to HBM -- and activations are quantised by the same kernel.
//
// Scales vary along K, so they do not factor out of the sum: one pipeline stage (BK = 128 = one scale block) is multiplied
// into a scratch TMEM accumulator, and the four epilogue warps promote it -- acc += tmp * sa[row][kb] * sb[col][kb] -- into
// fp32 registers (thread == row, 128 columns) while the tensor core works on the next stage in the other TMEM buffer.
// Warp roles as in gemm_fp8.cu: TMA producer, one MMA-issuing lane, four promotion / epilogue warps; the sb values of a stage
// are staged through shared memory (thread t fetches column t's scale).
// Reference counterpart: none (its quantised inference path dequantises int8 weights to fp16 in HBM before cuBLAS,
// csrc/transformer/inference/csrc/dequantize.cu); role: SURVEY K2 / K5.
#include "gemm_core.cuh"

namespace dsb {
namespace gemm8b {

constexpr int BM = 128;
constexpr int BN = 128;
constexpr int BKB = 128;  // fp8 elements (= bytes) of K per stage = one scale block
constexpr int STAGES = 5;
constexpr int A_BYTES = BM * BKB;
constexpr int B_BYTES = BN * BKB;
constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
constexpr int EPI_BUF = 128 * 128;  // 128 rows x 64 bf16
constexpr int SB_OFFSET = STAGES * STAGE_BYTES + 2 * EPI_BUF;  // float sbuf[2][BN]
constexpr int BAR_OFFSET = SB_OFFSET + 2 * BN * 4;
constexpr int SMEM_TOTAL = BAR_OFFSET + 256 + 1024;
constexpr int NUM_THREADS = 192;
constexpr int EPI_THREADS = 128;

struct Params {
  int M, N, K, KB;  // KB = K / 128
  int tiles_m, tiles_n;
  const float* sa;  // [M][KB]
  const float* sb;  // [N][KB]
};

__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_fp8_block_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                      const __grid_constant__ CUtensorMap tmC, const Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* epi_buf = smem + STAGES * STAGE_BYTES;
  float* sbuf = reinterpret_cast<float*>(smem + SB_OFFSET);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + BAR_OFFSET);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full = empty_bar + STAGES;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int num_tiles = p.tiles_m * p.tiles_n;
  const int num_kb = p.KB;

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
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        int m_blk, n_blk;
        gemm::tile_coords(tile, p.tiles_m, p.tiles_n, m_blk, n_blk);
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait_lean(&empty_bar[stage], phase ^ 1);
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
    // one scratch accumulator per scale block, alternating between the two TMEM buffers
    constexpr uint32_t idesc = umma_idesc(BM, BN, /*e4m3*/ 0, 0, 0, 0);
    constexpr uint32_t HI = (1024u >> 4) | (1u << 14) | (2u << 29);
    constexpr uint32_t KM = (16u >> 4) << 16;
    const uint32_t tm = __shfl_sync(0xffffffffu, tmem_base, 0);
    const bool leader = elect_one();
    const uint32_t base = smem_u32(smem);
    int stage = 0;
    uint32_t phase = 0;
    int buf = 0;
    uint32_t buf_phase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait_lean(&tmem_empty[buf], buf_phase ^ 1);
        mbar_wait_lean(&full_bar[stage], phase);
        tc_fence_after();
        if (leader) {
          const uint32_t a_lo = KM | ((base + stage * STAGE_BYTES) >> 4);
          const uint32_t b_lo = KM | ((base + stage * STAGE_BYTES + A_BYTES) >> 4);
#pragma unroll
          for (int kk = 0; kk < BKB / 32; ++kk)
            umma_f8(tm + buf * BN, (static_cast<uint64_t>(HI) << 32) | (a_lo + kk * 2), (static_cast<uint64_t>(HI) << 32) | (b_lo + kk * 2),
                    idesc, kk != 0 ? 1u : 0u);
          umma_commit<1>(&empty_bar[stage]);
          umma_commit<1>(&tmem_full[buf]);
        }
        if (++stage == STAGES) {
          stage = 0;
          phase ^= 1;
        }
        buf ^= 1;
        if (buf == 0) buf_phase ^= 1;
      }
    }
  } else {
    const int q = warp & 3;
    const int row = q * 32 + lane;
    const int epi_tid = threadIdx.x - 64;  // 0..127 == row
    int buf = 0;
    uint32_t buf_phase = 0, out_idx = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      int m_blk, n_blk;
      gemm::tile_coords(tile, p.tiles_m, p.tiles_n, m_blk, n_blk);
      const int m0 = m_blk * BM, n0 = n_blk * BN;
      const bool row_ok = m0 + row < p.M;
      const bool col_ok = n0 + epi_tid < p.N;
      const float* sa_row = p.sa + static_cast<size_t>(row_ok ? m0 + row : 0) * p.KB;
      const float* sb_col = p.sb + static_cast<size_t>(col_ok ? n0 + epi_tid : 0) * p.KB;
      float acc[BN];
#pragma unroll
      for (int i = 0; i < BN; ++i) acc[i] = 0.f;
      for (int kb = 0; kb < num_kb; ++kb) {
        float* sbk = sbuf + (kb & 1) * BN;
        sbk[epi_tid] = col_ok ? __ldg(sb_col + kb) : 0.f;
        const float sa = row_ok ? __ldg(sa_row + kb) : 0.f;
        named_bar_sync(1, EPI_THREADS);  // this stage's column scales are in shared memory (and the buffer of kb - 2 is retired)
        mbar_wait_lean(&tmem_full[buf], buf_phase);
        tc_fence_after();
        const uint32_t t_acc = tmem_base + buf * BN + (static_cast<uint32_t>(q * 32) << 16);
#pragma unroll
        for (int c = 0; c < BN / 32; ++c) {
          uint32_t v[32];
          tmem_ld_32x32(t_acc + c * 32, v);
          tmem_ld_wait();
          if (c == BN / 32 - 1) {
            tc_fence_before();
            mbar_arrive(&tmem_empty[buf]);  // scratch accumulator drained: the tensor core may overwrite it
          }
#pragma unroll
          for (int i = 0; i < 32; ++i) acc[c * 32 + i] = fmaf(__uint_as_float(v[i]) * sa, sbk[c * 32 + i], acc[c * 32 + i]);
        }
        buf ^= 1;
        if (buf == 0) buf_phase ^= 1;
      }
      // ---- tile done: bf16 -> swizzled staging -> TMA store, 64 columns at a time
#pragma unroll
      for (int c = 0; c < BN / 64; ++c) {
        if (n0 + c * 64 >= p.N) continue;  // uniform across the CTA
        uint8_t* obuf = epi_buf + (out_idx & 1) * EPI_BUF;
        if (epi_tid == 0) tma_wait_group_read<1>();
        named_bar_sync(2, EPI_THREADS);
        uint8_t* rowp = obuf + row * 128;
        const int sw = row & 7;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float* o = &acc[c * 64 + j * 8];
          uint4 val;
          val.x = pack_bf16x2(o[0], o[1]);
          val.y = pack_bf16x2(o[2], o[3]);
          val.z = pack_bf16x2(o[4], o[5]);
          val.w = pack_bf16x2(o[6], o[7]);
          *reinterpret_cast<uint4*>(rowp + ((j ^ sw) << 4)) = val;
        }
        fence_proxy_async_smem();
        named_bar_sync(2, EPI_THREADS);
        if (epi_tid == 0) {
          tma_store_2d(&tmC, obuf, n0 + c * 64, m0);
          tma_commit_group();
        }
        ++out_idx;
      }
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

}  // namespace gemm8b
}  // namespace dsb

// C[M,N] (bf16) = sum over 128-wide K blocks of sa[m][kb] * sb[n][kb] * (A_kb * B_kb^T); A [M][K], B [N][K] e4m3 (K contiguous,
// K % 128 == 0), sa [M][K/128], sb [N][K/128] fp32 (the layout dsb200_fp8_block_quantize writes for row-major operands).
DSB_EXPORT int dsb200_gemm_fp8_block(const void* A, const void* B, void* C, const float* sa, const float* sb, int M, int N, int K,
                                     int lda, int ldb, int ldc, int max_ctas, cudaStream_t stream) {
  using namespace dsb;
  DSB_REQUIRE(M > 0 && N > 0 && K > 0 && K % 128 == 0, "gemm_fp8_block: K (%d) must be a positive multiple of 128", K);
  DSB_REQUIRE(sa != nullptr && sb != nullptr, "gemm_fp8_block: both scale tensors are required");
  DSB_REQUIRE((reinterpret_cast<uintptr_t>(A) & 15) == 0 && (reinterpret_cast<uintptr_t>(B) & 15) == 0 &&
                  (reinterpret_cast<uintptr_t>(C) & 15) == 0,
              "gemm_fp8_block: operands must be 16-byte aligned");
  DSB_REQUIRE(lda % 16 == 0 && ldb % 16 == 0 && ldc % 8 == 0, "gemm_fp8_block: lda/ldb must be multiples of 16, ldc of 8");
  CUtensorMap tmA, tmB, tmC;
  int rc = make_tensor_map_2d(&tmA, TMA_U8, A, M, K, lda, gemm8b::BM, gemm8b::BKB, TMA_SW_128);
  if (rc) return rc;
  rc = make_tensor_map_2d(&tmB, TMA_U8, B, N, K, ldb, gemm8b::BN, gemm8b::BKB, TMA_SW_128);
  if (rc) return rc;
  rc = make_tensor_map_2d(&tmC, TMA_BF16, C, M, N, ldc, gemm8b::BM, 64, TMA_SW_128);
  if (rc) return rc;
  static bool attr_set = false;
  if (!attr_set) {
    DSB_CHECK_CUDA(cudaFuncSetAttribute(gemm8b::gemm_fp8_block_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, gemm8b::SMEM_TOTAL));
    attr_set = true;
  }
  gemm8b::Params p;
  p.M = M; p.N = N; p.K = K; p.KB = K / 128;
  p.tiles_m = (M + gemm8b::BM - 1) / gemm8b::BM;
  p.tiles_n = (N + gemm8b::BN - 1) / gemm8b::BN;
  p.sa = sa;
  p.sb = sb;
  int grid = p.tiles_m * p.tiles_n;
  int cap = num_sms();
  if (max_ctas > 0 && max_ctas < cap) cap = max_ctas;
  if (grid > cap) grid = cap;
  gemm8b::gemm_fp8_block_kernel<<<grid, gemm8b::NUM_THREADS, gemm8b::SMEM_TOTAL, stream>>>(tmA, tmB, tmC, p);
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

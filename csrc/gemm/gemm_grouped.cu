// Grouped (per-expert) bf16 GEMM for MoE on tcgen05:  C[rows of expert e] = A[rows of expert e] * W_e^T  for every expert
// in ONE persistent launch, with the row ranges read on the device (no host synchronisation on the token counts).
//
//   A  [R][K]     token rows already sorted by expert (output of moe_scatter)
//   W  [E*N][K]   the experts' weight matrices stacked along rows (each [N][K], K contiguous)
//   C  [R][N]
//   cum[E]        int64 inclusive cumulative row counts ("total rows before expert e+1")
//
// Tiles never straddle two experts: expert e owns ceil(rows_e/128) m-tiles that start at its first row.  The last
// m-tile of an expert may cover rows of the next expert (or rows past R, which TMA zero-fills) -- those accumulator rows
// are computed and thrown away: full tiles leave through the swizzled-smem TMA store, the partial tile is written
// from registers with a row predicate.  Tile order is (expert, n-block, m-tile) with m fastest, so the CTAs of a wave share
// one B tile in L2 while the weights of an expert stream from HBM exactly once.
// Pipeline, roles and barriers are those of gemm_core.cuh (TMA producer warp, single MMA thread, 4 epilogue warps,
// double-buffered TMEM accumulators).
// Replaces the reference's CUTLASS 2.x grouped GEMM (deepspeed/inference/v2/kernels/cutlass_ops/moe_gemm/moe_gemm.cu:29-175,
// kernel body in the closed `dskernels` library) that is driven by the same `total_rows_before_expert` vector.
#include "gemm_core.cuh"

namespace dsb {
namespace gemm {

constexpr int GROUPED_MAX_E = 128;
constexpr int GROUPED_TABLE_BYTES = (2 * GROUPED_MAX_E + 4) * 4;

struct GroupedParams {
  int R, N, K, E;
  int tiles_n;
  const long long* cum;
  __nv_bfloat16* C;
  int ldc;
};

// smem table built once per CTA: mt_prefix[e] = number of m-tiles of experts < e ; row0[e] = first row of expert e
struct GroupedTable {
  int* mt_prefix;  // [E + 1]
  int* row0;       // [E + 1]  (row0[E] = R_used)
};

struct TileWalk {
  int e;  // running expert index: tile ids only grow, so the search never goes back
  __device__ __forceinline__ void locate(int tile, const GroupedTable& t, int tiles_n, int E, int& m0, int& n_blk,
                                         int& valid_rows, int& expert) {
    while (e + 1 < E && tile >= t.mt_prefix[e + 1] * tiles_n) ++e;
    const int mt = t.mt_prefix[e + 1] - t.mt_prefix[e];
    const int local = tile - t.mt_prefix[e] * tiles_n;
    n_blk = local / mt;
    const int m = local - n_blk * mt;
    m0 = t.row0[e] + m * BM;
    valid_rows = min(BM, t.row0[e + 1] - m0);
    expert = e;
  }
};

template <int BN, int STAGES>
__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_grouped_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    const __grid_constant__ CUtensorMap tmC, const GroupedParams p) {
  using S = Smem<BN, STAGES>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* epi_buf = smem + STAGES * S::STAGE_BYTES;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + S::BAR_OFFSET);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full = empty_bar + STAGES;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tmem_empty + 2);
  GroupedTable tab;
  tab.mt_prefix = reinterpret_cast<int*>(smem + S::BAR_OFFSET + 256);
  tab.row0 = tab.mt_prefix + GROUPED_MAX_E + 2;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
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
  if (warp == 2 && lane == 0) {
    // E <= 128 entries: a serial prefix by one thread costs less than a microsecond and runs once per CTA
    int mt = 0;
    long long prev = 0;
    for (int e = 0; e < p.E; ++e) {
      long long c = p.cum[e];
      if (c > p.R) c = p.R;
      if (c < prev) c = prev;
      tab.mt_prefix[e] = mt;
      tab.row0[e] = static_cast<int>(prev);
      mt += static_cast<int>((c - prev + BM - 1) / BM);
      prev = c;
    }
    tab.mt_prefix[p.E] = mt;
    tab.row0[p.E] = static_cast<int>(prev);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  const int num_tiles = tab.mt_prefix[p.E] * p.tiles_n;

  if (warp == 0) {
    // ------------------------------------------------------------- TMA producer
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      TileWalk walk{0};
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        int m0, n_blk, valid, e;
        walk.locate(tile, tab, p.tiles_n, p.E, m0, n_blk, valid, e);
        const int nrow = e * p.N + n_blk * BN;  // row of the stacked weight matrix
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * S::STAGE_BYTES;
          uint8_t* sb = sa + S::A_BYTES;
          mbar_expect_tx(&full_bar[stage], S::STAGE_BYTES);
          tma_load_2d(&tmA, &full_bar[stage], sa, kb * BK, m0);
          tma_load_2d(&tmB, &full_bar[stage], sb, kb * BK, nrow);
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
      constexpr uint32_t idesc = umma_idesc(BM, BN, 1, 1, 0, 0);
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
            const uint64_t adesc = umma_smem_desc(sa + kk * (UMMA_K * 2), 16, 1024);
            const uint64_t bdesc = umma_smem_desc(sb + kk * (UMMA_K * 2), 16, 1024);
            umma_f16<1>(d_tmem, adesc, bdesc, idesc, (kb | kk) != 0 ? 1u : 0u);
          }
          umma_commit<1>(&empty_bar[stage]);
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
        umma_commit<1>(&tmem_full[acc]);
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1;
      }
    }
  } else {
    // ------------------------------------------------------------- epilogue (warps 2..5)
    const int q = warp & 3;
    const int row = q * 32 + lane;
    const int epi_tid = threadIdx.x - 64;
    constexpr int CW = 64;
    constexpr int NCHUNK = BN / CW;
    int acc = 0;
    uint32_t acc_phase = 0;
    uint32_t buf_idx = 0;
    TileWalk walk{0};
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      int m0, n_blk, valid, e;
      walk.locate(tile, tab, p.tiles_n, p.E, m0, n_blk, valid, e);
      const int n0 = n_blk * BN;
      const bool full = valid == BM;  // uniform across the CTA
      mbar_wait(&tmem_full[acc], acc_phase);
      tc_fence_after();
      const uint32_t t_acc = tmem_base + acc * BN + (static_cast<uint32_t>(q * 32) << 16);
#pragma unroll 1
      for (int c = 0; c < NCHUNK; ++c) {
        const bool live = (n0 + c * CW) < p.N;
        uint32_t v[32];
        uint32_t w[32];
        if (live) {
          tmem_ld_32x32(t_acc + c * CW, v);
          tmem_ld_32x32(t_acc + c * CW + 32, w);
          tmem_ld_wait();
        }
        if (c == NCHUNK - 1) {
          tc_fence_before();
          mbar_arrive(&tmem_empty[acc]);
        }
        if (!live) continue;
        uint4 out[8];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          out[j].x = pack_bf16x2(__uint_as_float(v[8 * j + 0]), __uint_as_float(v[8 * j + 1]));
          out[j].y = pack_bf16x2(__uint_as_float(v[8 * j + 2]), __uint_as_float(v[8 * j + 3]));
          out[j].z = pack_bf16x2(__uint_as_float(v[8 * j + 4]), __uint_as_float(v[8 * j + 5]));
          out[j].w = pack_bf16x2(__uint_as_float(v[8 * j + 6]), __uint_as_float(v[8 * j + 7]));
          out[j + 4].x = pack_bf16x2(__uint_as_float(w[8 * j + 0]), __uint_as_float(w[8 * j + 1]));
          out[j + 4].y = pack_bf16x2(__uint_as_float(w[8 * j + 2]), __uint_as_float(w[8 * j + 3]));
          out[j + 4].z = pack_bf16x2(__uint_as_float(w[8 * j + 4]), __uint_as_float(w[8 * j + 5]));
          out[j + 4].w = pack_bf16x2(__uint_as_float(w[8 * j + 6]), __uint_as_float(w[8 * j + 7]));
        }
        if (full) {
          uint8_t* buf = epi_buf + (buf_idx & 1) * EPI_BUF_BYTES;
          if (epi_tid == 0) tma_wait_group_read<1>();
          named_bar_sync(1, EPI_THREADS);
          uint8_t* rowp = buf + row * 128;
          const int sw = row & 7;
#pragma unroll
          for (int j = 0; j < 8; ++j) *reinterpret_cast<uint4*>(rowp + ((j ^ sw) << 4)) = out[j];
          fence_proxy_async_smem();
          named_bar_sync(1, EPI_THREADS);
          if (epi_tid == 0) {
            tma_store_2d(&tmC, buf, n0 + c * CW, m0);
            tma_commit_group();
          }
          ++buf_idx;
        } else if (row < valid) {
          // last m-tile of an expert: rows past `valid` belong to the next expert -- each thread writes its own row
          __nv_bfloat16* dst = p.C + static_cast<long long>(m0 + row) * p.ldc + n0 + c * CW;
          const int cols_left = p.N - (n0 + c * CW);
#pragma unroll
          for (int j = 0; j < 8; ++j)
            if (j * 8 + 8 <= cols_left) *reinterpret_cast<uint4*>(dst + j * 8) = out[j];
        }
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

template <int BN, int STAGES>
static int launch_grouped(const void* A, const void* W, void* C, const long long* cum, int R, int N, int K, int E, int lda,
                          int ldw, int ldc, cudaStream_t stream) {
  using S = Smem<BN, STAGES>;
  constexpr int TOTAL = S::TOTAL + GROUPED_TABLE_BYTES;
  static_assert(TOTAL <= 232448, "grouped GEMM shared memory exceeds 227 KB");
  CUtensorMap tmA, tmB, tmC;
  int rc = make_tensor_map_2d(&tmA, TMA_BF16, A, R, K, lda, BM, BK, TMA_SW_128);
  if (rc) return rc;
  rc = make_tensor_map_2d(&tmB, TMA_BF16, W, static_cast<uint64_t>(E) * N, K, ldw, BN, BK, TMA_SW_128);
  if (rc) return rc;
  rc = make_tensor_map_2d(&tmC, TMA_BF16, C, R, N, ldc, BM, 64, TMA_SW_128);
  if (rc) return rc;
  auto kern = gemm_grouped_kernel<BN, STAGES>;
  static bool attr_set = false;
  if (!attr_set) {
    DSB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, TOTAL));
    attr_set = true;
  }
  GroupedParams p;
  p.R = R;
  p.N = N;
  p.K = K;
  p.E = E;
  p.tiles_n = (N + BN - 1) / BN;
  p.cum = cum;
  p.C = static_cast<__nv_bfloat16*>(C);
  p.ldc = ldc;
  // upper bound of the tile count (the exact one is known only on the device): every expert adds at most one partial m-tile
  long long max_tiles = (static_cast<long long>(R) / BM + E) * p.tiles_n;
  int grid = num_sms();
  if (max_tiles < grid) grid = static_cast<int>(max_tiles);
  kern<<<grid, NUM_THREADS, TOTAL, stream>>>(tmA, tmB, tmC, p);
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

}  // namespace gemm
}  // namespace dsb

// C[r, :] = A[r, :] * W[expert(r)]^T for rows sorted by expert; `cum` (device, int64[E]) holds inclusive cumulative row counts.
// Rows at or beyond cum[E-1] are left untouched.  N must be a multiple of 8; A, W, C 16-byte aligned with 16-byte row strides.
DSB_EXPORT int dsb200_moe_gemm_bf16(const void* A, const void* W, void* C, const long long* cum, int R, int N, int K, int E,
                                    int lda, int ldw, int ldc, cudaStream_t stream) {
  using namespace dsb;
  DSB_REQUIRE(R > 0 && N > 0 && K > 0 && E > 0, "moe_gemm: empty problem R=%d N=%d K=%d E=%d", R, N, K, E);
  DSB_REQUIRE(E <= gemm::GROUPED_MAX_E, "moe_gemm: at most %d experts per launch (got %d)", gemm::GROUPED_MAX_E, E);
  DSB_REQUIRE(N % 8 == 0, "moe_gemm: N must be a multiple of 8 (got %d)", N);
  DSB_REQUIRE((reinterpret_cast<uintptr_t>(A) & 15) == 0 && (reinterpret_cast<uintptr_t>(W) & 15) == 0 &&
                  (reinterpret_cast<uintptr_t>(C) & 15) == 0,
              "moe_gemm: operands must be 16-byte aligned");
  DSB_REQUIRE(lda % 8 == 0 && ldw % 8 == 0 && ldc % 8 == 0, "moe_gemm: row strides must be multiples of 8 elements");
  // wide tiles once the average expert has a few m-tiles of work; narrow tiles (more CTAs streaming weights) for decode
  if (static_cast<long long>(R) / E >= 256 && N >= 256)
    return gemm::launch_grouped<256, 4>(A, W, C, cum, R, N, K, E, lda, ldw, ldc, stream);
  return gemm::launch_grouped<128, 6>(A, W, C, cum, R, N, K, E, lda, ldw, ldc, stream);
}

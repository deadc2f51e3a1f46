// GEMM -> all-reduce in ONE kernel (tensor-parallel row-parallel linear: o_proj / down_proj):
//     out = sum over ranks r of  X_r W_r^T   (+ residual),   X_r [M][K_r], W_r [N][K_r] bf16, out [M][N] bf16 on every rank.
// Each rank runs the persistent tcgen05 GEMM of gemm_core.cuh.  Its epilogue TMA-stores every 128 x 256 partial tile into the
// rank's OWN symmetric buffer and then publishes flag[tile] = epoch with a system-scope release.  Four extra "reducer" warps
// of the same CTA walk the same tile sequence: they wait until flag[tile] has reached `epoch` on EVERY rank (all ranks use
// the same tile order, so the peers' tiles arrive at about the same time), pull the W partial tiles with peer loads, add
// them in fp32 (+ residual) and write the bf16 result locally.  The reduction of tile t therefore runs under the MMAs of
// tiles t+1, t+2, ...: no separate all-reduce launch and no device-wide barriers around it.
//
// Re-use across calls: partials are double-buffered by epoch parity and flags carry the monotonically increasing epoch.
// A rank can enter call n+2 (which overwrites the buffer of call n) only after reducing call n+1, which needed every peer's
// call-n+1 partials, i.e. every peer had finished call n entirely (kernels of one stream run in order).
// Replaces the reference's cuBLAS GEMM + torch.distributed all_reduce pair of its tensor-parallel inference path
// (module_inject/layers.py LinearAllreduce.forward; csrc/transformer/inference `vector_matmul` + `dist.all_reduce`).
#include "gemm_core.cuh"

namespace dsb {
namespace gemm_ar {

constexpr int BM = 128;
constexpr int BN = 256;
constexpr int BK = 64;
constexpr int STAGES = 4;
constexpr int A_BYTES = BM * BK * 2;
constexpr int B_BYTES = BN * BK * 2;
constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
constexpr int EPI_BUF = 128 * 128;
constexpr int BAR_OFFSET = STAGES * STAGE_BYTES + 2 * EPI_BUF;
constexpr int SMEM_TOTAL = BAR_OFFSET + 256 + 1024;
constexpr int EPI_THREADS = 128;
constexpr int RED_THREADS = 128;
constexpr int NUM_THREADS = 64 + EPI_THREADS + RED_THREADS;
constexpr int MAX_RANKS = 16;

struct Params {
  int M, N, K;
  int tiles_m, tiles_n;
  void* peer[MAX_RANKS];    // base of every rank's symmetric buffer as mapped in this process (own rank included)
  unsigned long long part_off;  // byte offset of the [M][N] bf16 partial buffer of this call's parity
  unsigned long long flag_off;  // byte offset of the uint32 flags[tiles]
  uint32_t epoch;
  // graph-capturable form: the epoch is read from device memory (bumped by a one-thread kernel in front of this one) and picks
  // the partial area by its parity: part_off (even) / part_off_alt (odd).  nullptr: `epoch` / `part_off` as given by the host.
  const uint32_t* epoch_ptr;
  unsigned long long part_off_alt;
  int rank, world;
  __nv_bfloat16* out;
  int ldo;
  const __nv_bfloat16* residual;  // optional, added once (same on every rank); indexed like `out`
  int ldr;
  int row_lo, row_hi;  // rows this rank reduces and writes: [0, M) = all-reduce, its own slice = reduce-scatter; out row = r - row_lo
};

__device__ __forceinline__ void wait_lean(uint64_t* bar, uint32_t parity) { mbar_wait_lean(bar, parity); }
__device__ __forceinline__ uint32_t cur_epoch(const Params& p) { return p.epoch_ptr ? *p.epoch_ptr : p.epoch; }
__device__ __forceinline__ unsigned long long cur_part_off(const Params& p) {
  return (p.epoch_ptr && (*p.epoch_ptr & 1u)) ? p.part_off_alt : p.part_off;
}
__global__ void bump_epoch_kernel(uint32_t* ctr) { *ctr += 1; }
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint4 ld_peer_v4(const void* p) {
  uint4 r;
  asm volatile("ld.relaxed.sys.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
  return r;
}
__device__ __forceinline__ void acc_bf16x8(float (&a)[8], const uint4& v) {
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    a[2 * i] += __uint_as_float(w[i] << 16);
    a[2 * i + 1] += __uint_as_float(w[i] & 0xffff0000u);
  }
}

// Sum of one tile over the ranks, compile-time world W, U vectors per thread in one batch, so W*U 16-byte loads are in flight
// per thread before the first add, 16 KB per SM (a peer load is a ~2-3 us NVLink round trip; 128 threads x 1 load could not cover it).
template <int W, int U>
__device__ __forceinline__ void reduce_tile(const Params& p, int m0, int n0, int rt) {
  constexpr int VEC_PER_ROW = BN / 8;
  constexpr int NW = W;
  constexpr int world = W;
  const unsigned long long part_off = cur_part_off(p);
#pragma unroll 1
  for (int v0 = rt; v0 < BM * VEC_PER_ROW; v0 += RED_THREADS * U) {
    uint4 part[U][NW];
    uint32_t off[U];  // in 16-byte vectors (M*N/8 < 2^32)
    bool ok[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int v = v0 + u * RED_THREADS;
      const int r = v / VEC_PER_ROW, cv = v - r * VEC_PER_ROW;
      const int gr = m0 + r, gc = n0 + cv * 8;
      ok[u] = gr < p.row_hi && gr >= p.row_lo && gc < p.N;
      off[u] = ok[u] ? static_cast<uint32_t>((static_cast<size_t>(gr) * p.N + gc) >> 3) : 0u;
    }
#pragma unroll
    for (int k = 0; k < NW; ++k) {
      if (k < world) {
        const int src = (k + p.rank) % world;
        const uint8_t* base = static_cast<const uint8_t*>(p.peer[src]) + part_off;
#pragma unroll
        for (int u = 0; u < U; ++u)
          if (ok[u]) part[u][k] = ld_peer_v4(base + static_cast<size_t>(off[u]) * 16);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (!ok[u]) continue;
      float a[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int k = 0; k < NW; ++k)
        if (k < world) acc_bf16x8(a, part[u][k]);
      const int v = v0 + u * RED_THREADS;
      const int r = v / VEC_PER_ROW, cv = v - r * VEC_PER_ROW;
      const int gr = m0 + r, gc = n0 + cv * 8;
      const size_t lr = static_cast<size_t>(gr - p.row_lo);
      if (p.residual != nullptr) acc_bf16x8(a, *reinterpret_cast<const uint4*>(p.residual + lr * p.ldr + gc));
      uint4 o;
      o.x = pack_bf16x2(a[0], a[1]);
      o.y = pack_bf16x2(a[2], a[3]);
      o.z = pack_bf16x2(a[4], a[5]);
      o.w = pack_bf16x2(a[6], a[7]);
      *reinterpret_cast<uint4*>(p.out + lr * p.ldo + gc) = o;
    }
  }
}

// Any world size: one vector at a time, peers summed in arrival order of the loads the compiler can keep in flight.
__device__ __noinline__ void reduce_tile_any(const Params& p, int m0, int n0, int rt) {
  const unsigned long long part_off = cur_part_off(p);
  for (int v = rt; v < BM * (BN / 8); v += RED_THREADS) {
    const int r = v / (BN / 8), cv = v - r * (BN / 8);
    const int gr = m0 + r, gc = n0 + cv * 8;
    if (gr >= p.row_hi || gr < p.row_lo || gc >= p.N) continue;
    const size_t off = part_off + (static_cast<size_t>(gr) * p.N + gc) * 2;
    const size_t lr = static_cast<size_t>(gr - p.row_lo);
    float a[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int k = 0; k < p.world; ++k) acc_bf16x8(a, ld_peer_v4(static_cast<const uint8_t*>(p.peer[(k + p.rank) % p.world]) + off));
    if (p.residual != nullptr) acc_bf16x8(a, *reinterpret_cast<const uint4*>(p.residual + lr * p.ldr + gc));
    uint4 o;
    o.x = pack_bf16x2(a[0], a[1]);
    o.y = pack_bf16x2(a[2], a[3]);
    o.z = pack_bf16x2(a[4], a[5]);
    o.w = pack_bf16x2(a[6], a[7]);
    *reinterpret_cast<uint4*>(p.out + lr * p.ldo + gc) = o;
  }
}

__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_ar_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
               const __grid_constant__ CUtensorMap tmP, const __grid_constant__ CUtensorMap tmP1, const __grid_constant__ Params p) {
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
  const int num_kb = (p.K + BK - 1) / BK;
  const uint32_t epoch = cur_epoch(p);
  const CUtensorMap* tmPart = (p.epoch_ptr && (epoch & 1u)) ? &tmP1 : &tmP;  // partial area of this call's parity

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    tma_prefetch_desc(&tmP);
    tma_prefetch_desc(&tmP1);
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
          tma_load_2d(&tmA, &full_bar[stage], sa, kb * BK, m_blk * BM);
          tma_load_2d(&tmB, &full_bar[stage], sa + A_BYTES, kb * BK, n_blk * BN);
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------- MMA issuer (convergent warp, one elected lane)
    constexpr uint32_t idesc = umma_idesc(BM, BN, 1, 1, 0, 0);
    constexpr uint32_t HI = (1024u >> 4) | (1u << 14) | (2u << 29);
    constexpr uint32_t KM = (16u >> 4) << 16;
    const uint32_t tm = __shfl_sync(0xffffffffu, tmem_base, 0);
    const bool leader = elect_one();
    const uint32_t base = smem_u32(smem);
    int stage = 0, acc = 0;
    uint32_t phase = 0, acc_phase = 0;
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
          for (int kk = 0; kk < BK / 16; ++kk)
            umma_f16<1>(tm + acc * BN, (static_cast<uint64_t>(HI) << 32) | (a_lo + kk * 2), (static_cast<uint64_t>(HI) << 32) | (b_lo + kk * 2),
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
  } else if (warp < 6) {
    // ------------------------------------------------------------- epilogue: partial tile -> own symmetric buffer, then flag
    const int q = warp & 3;
    const int row = q * 32 + lane;
    const int epi_tid = threadIdx.x - 64;
    constexpr int CW = 64;
    uint32_t* my_flags = reinterpret_cast<uint32_t*>(static_cast<uint8_t*>(p.peer[p.rank]) + p.flag_off);
    int acc = 0;
    uint32_t acc_phase = 0, buf_idx = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      int m_blk, n_blk;
      gemm::tile_coords(tile, p.tiles_m, p.tiles_n, m_blk, n_blk);
      const int m0 = m_blk * BM, n0 = n_blk * BN;
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
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const uint32_t* s = j < 4 ? &v[8 * j] : &w[8 * (j - 4)];
          uint4 val;
          val.x = pack_bf16x2(__uint_as_float(s[0]), __uint_as_float(s[1]));
          val.y = pack_bf16x2(__uint_as_float(s[2]), __uint_as_float(s[3]));
          val.z = pack_bf16x2(__uint_as_float(s[4]), __uint_as_float(s[5]));
          val.w = pack_bf16x2(__uint_as_float(s[6]), __uint_as_float(s[7]));
          *reinterpret_cast<uint4*>(rowp + ((j ^ sw) << 4)) = val;
        }
        fence_proxy_async_smem();
        named_bar_sync(1, EPI_THREADS);
        if (epi_tid == 0) {
          tma_store_2d(tmPart, buf, n0 + c * CW, m0);
          tma_commit_group();
        }
        ++buf_idx;
      }
      if (epi_tid == 0) {
        tma_wait_group<0>();      // every byte of this tile's partial has been written
        __threadfence_system();
        st_release_sys(my_flags + tile, epoch);
      }
      acc ^= 1;
      if (acc == 0) acc_phase ^= 1;
    }
  } else {
    // ------------------------------------------------------------- reducers: out tile = sum of every rank's partial tile
    const int rt = threadIdx.x - (64 + EPI_THREADS);
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      int m_blk, n_blk;
      gemm::tile_coords(tile, p.tiles_m, p.tiles_n, m_blk, n_blk);
      const int m0 = m_blk * BM, n0 = n_blk * BN;
      if (m0 >= p.row_hi || m0 + BM <= p.row_lo) continue;  // reduce-scatter: another rank reduces this tile (uniform per CTA)
      if (rt < p.world) {
        const uint32_t* f = reinterpret_cast<const uint32_t*>(static_cast<const uint8_t*>(p.peer[rt]) + p.flag_off) + tile;
        uint32_t spins = 0;
        while (static_cast<int32_t>(ld_acquire_sys(f) - epoch) < 0) {
          if (++spins > DSB_SPIN_LIMIT) {
            printf("dsb200: GEMM-allreduce rank %d timed out waiting for tile %d of rank %d (epoch %u)\n", p.rank, tile, rt, epoch);
            __trap();
          }
          __nanosleep(32);
        }
      }
      named_bar_sync(2, RED_THREADS);
      switch (p.world) {
        case 1: reduce_tile<1, 4>(p, m0, n0, rt); break;
        case 2: reduce_tile<2, 4>(p, m0, n0, rt); break;
        case 4: reduce_tile<4, 2>(p, m0, n0, rt); break;
        case 8: reduce_tile<8, 1>(p, m0, n0, rt); break;
        default: reduce_tile_any(p, m0, n0, rt); break;
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<1>(tmem_base, 2 * BN);
  }
}

}  // namespace gemm_ar
}  // namespace dsb

// out[M,N] (bf16, local) = sum_r A_r B_r^T (+ residual).  A [M][K] / B [N][K] bf16 are this rank's operands; peer_ptrs[world] are the
// bases of every rank's symmetric buffer (own included) as mapped here.  The caller reserves inside the symmetric buffer
//   * two partial areas of M*N*2 bytes (16-byte aligned): pass part_off = area[epoch & 1]
//   * flags: ceil(M/128)*ceil(N/256) uint32, zero-initialised once (flag_off, 4-byte aligned)
// and passes the same, strictly increasing `epoch` (>= 1) on every rank for each call.  N must be a multiple of 8.
static int launch_gemm_ar(const void* A, const void* B, void* const* peer_ptrs, long long part_off, long long flag_off, unsigned int epoch,
                          unsigned int* epoch_ctr, long long part_off_alt, void* out, const void* residual, int M, int N, int K, int lda, int ldb, int ldo, int ldr, int rank, int world,
                          int row_lo, int row_hi, int max_ctas, cudaStream_t stream) {
  using namespace dsb;
  namespace G = gemm_ar;
  DSB_REQUIRE(row_lo >= 0 && row_lo <= row_hi && row_hi <= M, "gemm_ar: rows [%d, %d) outside [0, %d)", row_lo, row_hi, M);
  DSB_REQUIRE(M > 0 && N > 0 && K > 0, "gemm_ar: empty problem M=%d N=%d K=%d", M, N, K);
  DSB_REQUIRE(world >= 1 && world <= G::MAX_RANKS && rank >= 0 && rank < world, "gemm_ar: rank %d / world %d", rank, world);
  DSB_REQUIRE(N % 8 == 0 && lda % 8 == 0 && ldb % 8 == 0 && ldo % 8 == 0, "gemm_ar: N, lda, ldb, ldo must be multiples of 8");
  DSB_REQUIRE(part_off % 16 == 0 && part_off_alt % 16 == 0 && flag_off % 4 == 0 && (epoch >= 1 || epoch_ctr != nullptr),
              "gemm_ar: part_off must be 16-byte aligned, flag_off 4-byte, epoch >= 1");
  DSB_REQUIRE(residual == nullptr || ldr % 8 == 0, "gemm_ar: ldr must be a multiple of 8");
  CUtensorMap tmA, tmB, tmP, tmP1;
  int rc = make_tensor_map_2d(&tmA, TMA_BF16, A, M, K, lda, G::BM, G::BK, TMA_SW_128);
  if (rc) return rc;
  rc = make_tensor_map_2d(&tmB, TMA_BF16, B, N, K, ldb, G::BN, G::BK, TMA_SW_128);
  if (rc) return rc;
  // the partial tiles are stored into THIS rank's own buffer (the tensor map covers local memory only)
  void* my_part = static_cast<uint8_t*>(peer_ptrs[rank]) + part_off;
  rc = make_tensor_map_2d(&tmP, TMA_BF16, my_part, M, N, N, G::BM, 64, TMA_SW_128);
  if (rc) return rc;
  void* my_part1 = static_cast<uint8_t*>(peer_ptrs[rank]) + (epoch_ctr ? part_off_alt : part_off);
  rc = make_tensor_map_2d(&tmP1, TMA_BF16, my_part1, M, N, N, G::BM, 64, TMA_SW_128);
  if (rc) return rc;
  static bool attr_set = false;
  if (!attr_set) {
    DSB_CHECK_CUDA(cudaFuncSetAttribute(G::gemm_ar_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, G::SMEM_TOTAL));
    attr_set = true;
  }
  G::Params p;
  p.M = M;
  p.N = N;
  p.K = K;
  p.tiles_m = (M + G::BM - 1) / G::BM;
  p.tiles_n = (N + G::BN - 1) / G::BN;
  for (int i = 0; i < G::MAX_RANKS; ++i) p.peer[i] = i < world ? peer_ptrs[i] : nullptr;
  p.part_off = static_cast<unsigned long long>(part_off);
  p.flag_off = static_cast<unsigned long long>(flag_off);
  p.epoch = epoch;
  p.epoch_ptr = epoch_ctr;
  p.part_off_alt = static_cast<unsigned long long>(part_off_alt);
  p.rank = rank;
  p.world = world;
  p.out = static_cast<__nv_bfloat16*>(out);
  p.ldo = ldo;
  p.residual = static_cast<const __nv_bfloat16*>(residual);
  p.ldr = ldr;
  p.row_lo = row_lo;
  p.row_hi = row_hi;
  int grid = p.tiles_m * p.tiles_n;
  int cap = num_sms();
  if (max_ctas > 0 && max_ctas < cap) cap = max_ctas;
  if (grid > cap) grid = cap;
  if (epoch_ctr) G::bump_epoch_kernel<<<1, 1, 0, stream>>>(epoch_ctr);
  G::gemm_ar_kernel<<<grid, G::NUM_THREADS, G::SMEM_TOTAL, stream>>>(tmA, tmB, tmP, tmP1, p);
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

DSB_EXPORT int dsb200_gemm_ar_bf16(const void* A, const void* B, void* const* peer_ptrs, long long part_off, long long flag_off,
                                   unsigned int epoch, void* out, const void* residual, int M, int N, int K, int lda, int ldb, int ldo,
                                   int ldr, int rank, int world, int max_ctas, cudaStream_t stream) {
  return launch_gemm_ar(A, B, peer_ptrs, part_off, flag_off, epoch, nullptr, part_off, out, residual, M, N, K, lda, ldb, ldo, ldr, rank, world,
                        0, M, max_ctas, stream);
}

// GEMM -> reduce-scatter over rows (sequence-parallel tensor parallelism): same kernel, but this rank reduces and writes only rows
// [row_lo, row_hi) of the sum -- out and residual are [row_hi - row_lo][N].  Every rank still publishes its whole partial (the
// owners of the other rows need it); the row ranges of the ranks must cover [0, M) and every rank must own at least one row.
DSB_EXPORT int dsb200_gemm_rs_bf16(const void* A, const void* B, void* const* peer_ptrs, long long part_off, long long flag_off,
                                   unsigned int epoch, void* out, const void* residual, int M, int N, int K, int lda, int ldb, int ldo,
                                   int ldr, int rank, int world, int row_lo, int row_hi, int max_ctas, cudaStream_t stream) {
  DSB_REQUIRE(row_hi > row_lo, "gemm_rs: rank %d owns no rows", rank);
  return launch_gemm_ar(A, B, peer_ptrs, part_off, flag_off, epoch, nullptr, part_off, out, residual, M, N, K, lda, ldb, ldo, ldr, rank, world,
                        row_lo, row_hi, max_ctas, stream);
}

// Graph-capturable forms: the epoch lives in device memory (`epoch_ctr`, one uint32 of LOCAL memory per rank, all ranks start it at
// the same value); every call bumps it with a one-thread kernel in front of the GEMM, and the GEMM picks the partial area
// part_off0 / part_off1 by the bumped value's parity.  A captured call therefore replays correctly any number of times.
// row_lo = 0, row_hi = M: all-reduce; a sub-range: reduce-scatter over rows.
DSB_EXPORT int dsb200_gemm_ar_bf16_dev(const void* A, const void* B, void* const* peer_ptrs, long long part_off0, long long part_off1,
                                       long long flag_off, unsigned int* epoch_ctr, void* out, const void* residual, int M, int N, int K,
                                       int lda, int ldb, int ldo, int ldr, int rank, int world, int row_lo, int row_hi, int max_ctas,
                                       cudaStream_t stream) {
  DSB_REQUIRE(epoch_ctr != nullptr && row_hi > row_lo, "gemm_ar_dev: epoch counter missing or empty row range");
  return launch_gemm_ar(A, B, peer_ptrs, part_off0, flag_off, 0, epoch_ctr, part_off1, out, residual, M, N, K, lda, ldb, ldo, ldr, rank, world,
                        row_lo, row_hi, max_ctas, stream);
}

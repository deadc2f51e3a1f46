// CTA-pair (cta_group::2) variant of the tcgen05 bf16 GEMM.
//
// Two CTAs on the two SMs of a TPC form a cluster and compute one 256 x 256 output tile together:
//   * each CTA stages ITS 128 rows of A and ITS 128 rows (N-columns) of B per 64-wide k-block: 32 KB per stage instead
//     of 48 KB for the single-CTA 128 x 256 tile, while the tensor cores of both SMs see the full 256-wide B -- the B
//     halves are read across the pair by the MMA unit.  Shared-memory traffic per FLOP halves, 6 stages fit.
//   * only the leader CTA (cluster rank 0) issues `tcgen05.mma.cta_group::2` (M = 256, N = 256, K = 16); the
//     accumulator of each CTA's 128 rows lives in that CTA's own TMEM (double buffered, 2 x 256 columns).
//   * barrier protocol
//       full[s]       leader only, count 1: leader's producer posts expect_tx for BOTH CTAs' bytes; every TMA load of
//                     either CTA signals the leader's barrier (cp.async.bulk.tensor ... cta_group::2)
//       empty[s]      one per CTA, count 1: tcgen05.commit multicast to both CTAs when the MMAs reading slot s retire
//       tmem_full[a]  one per CTA, count 1: commit multicast when an accumulator is complete
//       tmem_empty[a] leader only, count 2: one remote arrive per CTA after its epilogue drained accumulator a
//   * each CTA's epilogue converts and TMA-stores (or reduce-adds) its own 128 x 256 half of the tile.
// Everything else (persistent tile loop with grouped raster, 128B-swizzled operand tiles, operand layouts) is as in
// gemm_core.cuh.
#pragma once
#include "gemm_core.cuh"

namespace dsb {
namespace gemm {

constexpr int BN2 = 256;      // N of the pair tile
constexpr int BN2_CTA = 128;  // rows of B staged by each CTA
constexpr int BM2 = 256;      // M of the pair tile (128 per CTA)

template <int STAGES>
struct Smem2 {
  static constexpr int A_BYTES = BM * BK * 2;
  static constexpr int B_BYTES = BN2_CTA * BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int BAR_OFFSET = STAGES * STAGE_BYTES + 2 * EPI_BUF_BYTES;
  static constexpr int TOTAL = BAR_OFFSET + 256 + 1024;
};

__device__ __forceinline__ uint32_t mapa_u32(uint32_t smem_addr, uint32_t cta_rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_addr), "r"(cta_rank));
  return r;
}

// AG = true adds the all-gather of the sharded B operand (see AgArgs in gemm_core.cuh): two extra warps per CTA pull
// 256-row chunks of B from the owning peers over NVLink in the order the (n-major) tile loop consumes them and publish
// a completion counter per chunk; the TMA producers wait for the counter of their n-block.
template <int STAGES, bool A_MN, bool B_MN, bool OUT_F32, bool AG>
__device__ __forceinline__ void gemm2_body(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmC,
                                           const Params& p, const AgArgs* ag) {
  using S = Smem2<STAGES>;
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
  const uint32_t rank = cluster_ctarank();
  const int cluster_id = blockIdx.x >> 1;
  const int num_clusters = gridDim.x >> 1;
  const int num_tiles = p.tiles_m * p.tiles_n;  // 256 x 256 tiles
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
      mbar_init(&tmem_empty[i], 2);
    }
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc<2>(tmem_ptr, 2 * BN2);
    tmem_relinquish<2>();
  }
  tc_fence_before();
  cluster_sync_all();  // barriers of BOTH CTAs initialised and visible before any remote arrive / TMA completion
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  if (warp == 0) {
    // ------------------------------------------------------------- TMA producer (both CTAs, each its own halves)
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
        int m_blk, n_blk;
        if constexpr (AG) { n_blk = tile / p.tiles_m; m_blk = tile - n_blk * p.tiles_m; }
        else tile_coords(tile, p.tiles_m, p.tiles_n, m_blk, n_blk);
        const int m0 = m_blk * BM2 + static_cast<int>(rank) * BM;
        const int nb0 = n_blk * BN2 + static_cast<int>(rank) * BN2_CTA;
        if constexpr (AG) {
          const int target = static_cast<int>(gridDim.x);  // every CTA's copy warps landed their share of this chunk
          uint32_t spins = 0;
          int seen;
          do {
            asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(seen) : "l"(ag->flags + n_blk) : "memory");
            if (seen < target && ++spins > DSB_SPIN_LIMIT) {
              printf("dsb200: AG-GEMM(2cta) wait for n-block %d timed out (%d/%d)\n", n_blk, seen, target);
              __trap();
            }
          } while (seen < target);
          asm volatile("fence.proxy.async;" ::: "memory");
        }
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * S::STAGE_BYTES;
          uint8_t* sb = sa + S::A_BYTES;
          if (rank == 0) mbar_expect_tx(&full_bar[stage], 2 * S::STAGE_BYTES);
          const uint32_t bar = mapa_u32(smem_u32(&full_bar[stage]), 0);  // the leader's barrier
          const int k0 = kb * BK;
          if constexpr (!A_MN) {
            tma_load_2d_bar_addr(&tmA, bar, sa, k0, m0);
          } else {
#pragma unroll
            for (int c = 0; c < BM / 64; ++c) tma_load_2d_bar_addr(&tmA, bar, sa + c * (BK * 128), m0 + c * 64, k0);
          }
          if constexpr (!B_MN) {
            tma_load_2d_bar_addr(&tmB, bar, sb, k0, nb0);
          } else {
#pragma unroll
            for (int c = 0; c < BN2_CTA / 64; ++c) tma_load_2d_bar_addr(&tmB, bar, sb + c * (BK * 128), nb0 + c * 64, k0);
          }
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------- MMA issuer (leader CTA only)
    if (lane == 0 && rank == 0) {
      constexpr uint32_t idesc = umma_idesc(BM2, BN2, /*bf16*/ 1, 1, A_MN ? 1 : 0, B_MN ? 1 : 0);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
        mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN2;
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
            umma_f16<2>(d_tmem, adesc, bdesc, idesc, (kb | kk) != 0 ? 1u : 0u);
          }
          umma_commit_mcast2(&empty_bar[stage], 0b11);  // slot s is free in BOTH CTAs once these MMAs retire
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
        umma_commit_mcast2(&tmem_full[acc], 0b11);
        acc ^= 1;
        if (acc == 0) acc_phase ^= 1;
      }
    }
  } else if (AG && warp >= 6) {
    // ------------------------------------------------------------- all-gather copy warps (AG only): peers -> b_local
    if constexpr (AG) {
      const int ct = threadIdx.x - 192;  // 0..63
      // Copy work = tiles_n x gridDim.x chunks, handed out in consumption order through ONE global claim counter
      // (flags[tiles_n]): whichever CTAs are resident copy everything, so the kernel makes progress even when other
      // kernels hold some SMs and part of this grid has not started yet (no CTA waits for a specific other CTA).
      __shared__ int s_claim;
      const long long row_elems = p.K;
      const int chunks_per_nb = static_cast<int>(gridDim.x);
      const int total_chunks = p.tiles_n * chunks_per_nb;
      int* claim = ag->flags + p.tiles_n;
      while (true) {
        if (ct == 0) s_claim = atomicAdd(claim, 1);
        named_bar_sync(2, 64);
        const int idx = s_claim;
        named_bar_sync(2, 64);  // everyone has read s_claim before the next claim overwrites it
        if (idx >= total_chunks) break;
        const int nb = idx / chunks_per_nb, chunk = idx - nb * chunks_per_nb;
        const long long rows = min(BN2, p.N - nb * BN2);
        const long long chunk_elems = rows * row_elems;
        const long long chunk_off = static_cast<long long>(nb) * BN2 * row_elems;  // offset inside B
        const long long vecs = chunk_elems / 8;
        const long long per_chunk = (vecs + chunks_per_nb - 1) / chunks_per_nb;
        const long long v_lo = per_chunk * chunk;
        const long long v_hi = min(vecs, v_lo + per_chunk);
        constexpr int UNR = 8;
        for (long long v = v_lo + ct; v < v_hi; v += 64 * UNR) {
          uint4 buf[UNR];
#pragma unroll
          for (int u = 0; u < UNR; ++u) {
            const long long vv = v + static_cast<long long>(u) * 64;
            if (vv < v_hi) {
              const long long e = ag->w_off + chunk_off + vv * 8;  // flat element index in the unit
              const int peer = static_cast<int>(e / ag->shard_numel);
              const long long le = e - static_cast<long long>(peer) * ag->shard_numel;
              const void* src = static_cast<const __nv_bfloat16*>(ag->peer_shard[peer]) + le;
              asm volatile("ld.relaxed.sys.global.v4.u32 {%0, %1, %2, %3}, [%4];"
                           : "=r"(buf[u].x), "=r"(buf[u].y), "=r"(buf[u].z), "=r"(buf[u].w)
                           : "l"(src)
                           : "memory");
            }
          }
#pragma unroll
          for (int u = 0; u < UNR; ++u) {
            const long long vv = v + static_cast<long long>(u) * 64;
            if (vv < v_hi) *reinterpret_cast<uint4*>(ag->b_local + chunk_off + vv * 8) = buf[u];
          }
        }
        __threadfence();
        named_bar_sync(2, 64);
        if (ct == 0) asm volatile("red.release.gpu.global.add.s32 [%0], 1;" ::"l"(ag->flags + nb) : "memory");
      }
    }
  } else {
    // ------------------------------------------------------------- epilogue (warps 2..5 of both CTAs)
    const int q = warp & 3;
    const int row = q * 32 + lane;
    const int epi_tid = threadIdx.x - 64;
    constexpr int CW = OUT_F32 ? 32 : 64;
    constexpr int NCHUNK = BN2 / CW;
    int acc = 0;
    uint32_t acc_phase = 0;
    uint32_t buf_idx = 0;
    for (int tile = cluster_id; tile < num_tiles; tile += num_clusters) {
      int m_blk, n_blk;
      if constexpr (AG) { n_blk = tile / p.tiles_m; m_blk = tile - n_blk * p.tiles_m; }
      else tile_coords(tile, p.tiles_m, p.tiles_n, m_blk, n_blk);
      const int m0 = m_blk * BM2 + static_cast<int>(rank) * BM, n0 = n_blk * BN2;
      mbar_wait(&tmem_full[acc], acc_phase);
      tc_fence_after();
      const uint32_t t_acc = tmem_base + acc * BN2 + (static_cast<uint32_t>(q * 32) << 16);
      const bool rows_live = m0 < p.M;  // the second CTA's half may lie entirely below the matrix
#pragma unroll 1
      for (int c = 0; c < NCHUNK; ++c) {
        const bool live = rows_live && (n0 + c * CW) < p.N;  // uniform across the CTA
        uint32_t v[32];
        uint32_t w[32];
        if (live) {
          tmem_ld_32x32(t_acc + c * CW, v);
          if constexpr (!OUT_F32) tmem_ld_32x32(t_acc + c * CW + 32, w);
          tmem_ld_wait();
        }
        if (c == NCHUNK - 1) {
          // accumulator drained by this CTA: one arrive per CTA on the leader's barrier
          tc_fence_before();
          named_bar_sync(3, EPI_THREADS);
          if (epi_tid == 0) mbar_arrive_cluster(&tmem_empty[acc], 0);
        }
        if (!live) continue;
        uint8_t* buf = epi_buf + (buf_idx & 1) * EPI_BUF_BYTES;
        if (epi_tid == 0) tma_wait_group_read<1>();
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

  // no CTA of the pair may exit (or free TMEM) while its partner can still read its shared memory / signal its barriers
  tc_fence_before();
  cluster_sync_all();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<2>(tmem_base, 2 * BN2);
  }
}

template <int STAGES, bool A_MN, bool B_MN, bool OUT_F32>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(NUM_THREADS, 1)
gemm2_bf16_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                  const __grid_constant__ CUtensorMap tmC, const Params p) {
  gemm2_body<STAGES, A_MN, B_MN, OUT_F32, false>(tmA, tmB, tmC, p, nullptr);
}

template <int STAGES>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(NUM_THREADS_AG, 1)
gemm2_ag_bf16_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                     const __grid_constant__ CUtensorMap tmC, const Params p, const __grid_constant__ AgArgs ag) {
  gemm2_body<STAGES, false, false, false, true>(tmA, tmB, tmC, p, &ag);
}

}  // namespace gemm
}  // namespace dsb

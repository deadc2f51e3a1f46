// Shared body of the tcgen05 bf16 GEMM kernels (plain and all-gather-fused).  See gemm_bf16.cu for the design notes.
#pragma once
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

// All-gather fused into the GEMM (AG = true): the B operand (a weight matrix [N][K], K contiguous) is sharded over
// `world` ranks as one flat buffer -- rank r holds flat elements [r*shard, (r+1)*shard) of the enclosing ZeRO unit.
// Two extra warps per CTA pull the rows of B from the owning peers over NVLink into `b_local` (chunk = one BN-row
// block, in the order the tiles consume them) and publish per-chunk completion counters; the TMA producer waits on
// the counter of its n-block before loading B tiles.  The transfer therefore overlaps the math tile by tile.
constexpr int AG_MAX_RANKS = 16;
struct AgArgs {
  const void* peer_shard[AG_MAX_RANKS];  // base of every rank's parameter shard region (mapped peer pointers)
  long long w_off;                        // flat element offset of B[0][0] inside the unit
  long long shard_numel;                  // elements per rank shard of the unit
  __nv_bfloat16* b_local;                 // gathered destination == the tensor the TMA descriptor of B points to
  int* flags;                             // [tiles_n] zero-initialised completion counters
  int world;
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

template <int BN, int STAGES, bool A_MN, bool B_MN, bool OUT_F32, bool AG>
__device__ __forceinline__ void gemm_body(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmC,
                                          const Params& p, const AgArgs* ag) {
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
        if constexpr (AG) { n_blk = tile / p.tiles_m; m_blk = tile - n_blk * p.tiles_m; }
        else tile_coords(tile, p.tiles_m, p.tiles_n, m_blk, n_blk);
        const int m0 = m_blk * BM, n0 = n_blk * BN;
        if constexpr (AG) {
          // wait until every CTA's copy warps have landed their share of this n-block of B in local memory
          const int target = static_cast<int>(gridDim.x);
          uint32_t spins = 0;
          int seen;
          do {
            asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(seen) : "l"(ag->flags + n_blk) : "memory");
            if (seen < target && ++spins > DSB_SPIN_LIMIT) {
              printf("dsb200: AG-GEMM wait for n-block %d timed out (%d/%d)\n", n_blk, seen, target);
              __trap();
            }
          } while (seen < target);
          asm volatile("fence.proxy.async;" ::: "memory");  // peer data was written with generic stores, TMA reads it
        }
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
        const long long rows = min(BN, p.N - nb * BN);
        const long long chunk_elems = rows * row_elems;
        const long long chunk_off = static_cast<long long>(nb) * BN * row_elems;  // offset inside B
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
      if constexpr (AG) { n_blk = tile / p.tiles_m; m_blk = tile - n_blk * p.tiles_m; }
      else tile_coords(tile, p.tiles_m, p.tiles_n, m_blk, n_blk);
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
__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                 const __grid_constant__ CUtensorMap tmC, const Params p) {
  gemm_body<BN, STAGES, A_MN, B_MN, OUT_F32, false>(tmA, tmB, tmC, p, nullptr);
}

constexpr int NUM_THREADS_AG = NUM_THREADS + 64;
template <int BN, int STAGES>
__global__ void __launch_bounds__(NUM_THREADS_AG, 1)
gemm_ag_bf16_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    const __grid_constant__ CUtensorMap tmC, const Params p, const __grid_constant__ AgArgs ag) {
  gemm_body<BN, STAGES, false, false, false, true>(tmA, tmB, tmC, p, &ag);
}

}  // namespace gemm
}  // namespace dsb

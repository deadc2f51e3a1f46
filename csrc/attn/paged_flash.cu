// Paged flash attention for the FastGen ragged engine on tcgen05 tensor cores (sm_100a): chunked prefill and decode in one
// kernel, KV read page by page out of the blocked cache with TMA.
//
// Unit of work = atom x kv head.  An atom is a run of up to 128/group consecutive new tokens of ONE sequence (group = q heads
// per kv head); its `group` q heads are packed into the M dimension of the MMA tile, row r = token_in_atom * group + head_in_group,
// so a decode step of a GQA 8:1 model still feeds 8 rows per token and every KV byte is read once per kv head.  Atoms are
// found on the device from the ragged batch metadata (no host round trip): token t opens an atom when it is the first token of
// the batch, the first token of its sequence in this batch, or its position is a multiple of 128/group.
//
//   warp 0   TMA producer : Q tile once (3-D map over [token][head][D], box [128/group tokens][group heads][64]), then K and V
//                           pages (4-D map over [block*2+kv][slot][kv head][D]; block ids from the block table) through a ring of
//                           KV_STAGES K tiles and KV_STAGES V tiles; a K tile is released as soon as its QK^T has retired and a
//                           V tile when its PV has, so two to three pages of loads are in flight per SM (decode is a pure
//                           KV-bandwidth problem: with one page in flight the kernel sat at 35 % of HBM)
//   warp 1   MMA issuer   : S = Q K^T (128 x 128 x D, SS) into TMEM, PV = P V (TS: P read from TMEM, V MN-major from smem) into TMEM
//   warps 2-5 softmax     : thread == row: S row from TMEM, per-row causal limit (position + 1), online softmax in exp2 domain,
//                           P (bf16, two per column) stored back over the first 64 columns of its S buffer, O kept in registers
// Long contexts at small batch do not fill 148 SMs with one CTA per (atom, kv head): with n_splits > 1 every CTA handles a
// slice of the pages and writes an unnormalised partial (O, running max, running sum) to scratch; a second kernel merges.
// Reference counterpart: inference/v2/kernels/ragged_ops/blocked_flash/blocked_flash.cpp:21-101 (closed flash_attn_by_atoms)
// and atom_builder/atom_builder.cpp:10-53.
#include "../common/host.h"
#include "../common/ptx.cuh"

namespace dsb {
namespace paged {

constexpr int BM = 128;
constexpr int BN = 128;
constexpr int NUM_THREADS = 320;  // TMA warp, MMA warp, 8 softmax warps
constexpr int SM_THREADS = 256;
constexpr int CHUNK_BYTES = 128 * 128;  // one [128 rows x 64 bf16] swizzled chunk

template <int D>
struct Smem {
  static constexpr int NCH = D / 64;
  static constexpr int TILE = NCH * CHUNK_BYTES;
  static constexpr int STAGES = D == 128 ? 3 : 6;  // K and V tiles each; Q + 2 * STAGES tiles = 224 KB (d = 128) / 208 KB (d = 64)
  static constexpr int Q_OFF = 0;
  static constexpr int K_OFF = Q_OFF + TILE;
  static constexpr int V_OFF = K_OFF + STAGES * TILE;
  static constexpr int BAR_OFF = V_OFF + STAGES * TILE;
  static constexpr int X_OFF = BAR_OFF + 384;  // 1 KB: page maxima exchanged between the two column halves of a row
  static constexpr int TOTAL = X_OFF + 1024 + 1024;
};

__device__ __forceinline__ float ex2_fast(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

struct Params {
  const int* tok_seq;      // [T] sequence slot of every token of the ragged batch
  const int* tok_pos;      // [T] position of the token inside its sequence
  const int* block_table;  // [seqs][max_blocks]
  const int* atoms;        // [n_atoms] first token of every atom (device-built)
  const int* n_atoms;      // device scalar
  __nv_bfloat16* out;      // [T][hq*D]
  float* part_o;           // [T][hq][n_splits][D]   (n_splits > 1)
  float* part_ml;          // [T][hq][n_splits][2]
  int T, hq, hkv, group, nt_tile, block_size, max_blocks, n_splits;
  float scale_log2;
};

// ---------------------------------------------------------------- atom discovery (one thread per token)
__global__ void find_atoms_kernel(const int* __restrict__ tok_seq, const int* __restrict__ tok_pos, int T, int nt_tile,
                                  int* __restrict__ atoms, int* __restrict__ n_atoms, int max_atoms) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= T) return;
  const bool opens = t == 0 || tok_seq[t] != tok_seq[t - 1] || tok_pos[t] != tok_pos[t - 1] + 1 || (tok_pos[t] % nt_tile) == 0;
  if (opens) {
    const int idx = atomicAdd(n_atoms, 1);
    if (idx < max_atoms) atoms[idx] = t;  // a caller that under-sized max_atoms loses atoms (reported by the host wrapper) but never writes out of bounds
  }
}

template <int D>
__global__ void __launch_bounds__(NUM_THREADS, 1)
paged_flash_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmKV, const Params p) {
  using SMEM = Smem<D>;
  constexpr int NCH = SMEM::NCH;
  if (static_cast<int>(blockIdx.x) >= *p.n_atoms) return;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + SMEM::BAR_OFF);
  constexpr int NS = SMEM::STAGES;
  uint64_t* q_full = bars + 0;
  uint64_t* k_full = bars + 1;             // [NS]
  uint64_t* v_full = k_full + NS;          // [NS]
  uint64_t* k_empty = v_full + NS;         // [NS]
  uint64_t* v_empty = k_empty + NS;        // [NS]
  uint64_t* s_full = v_empty + NS;         // [2]
  uint64_t* p_full = s_full + 2;           // [2]
  uint64_t* o_full = p_full + 2;           // [2]
  uint64_t* o_free = o_full + 2;           // [2]
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(o_free + 2);
  int* s_nt = reinterpret_cast<int*>(o_free + 3);  // tokens in this atom

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int t0 = p.atoms[blockIdx.x];
  const int kvh = blockIdx.y, split = blockIdx.z;
  const int seq = p.tok_seq[t0], pos0 = p.tok_pos[t0];

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmQ);
    tma_prefetch_desc(&tmKV);
    mbar_init(q_full, 1);
    for (int i = 0; i < NS; ++i) {
      mbar_init(&k_full[i], 1);
      mbar_init(&v_full[i], 1);
      mbar_init(&k_empty[i], 1);
      mbar_init(&v_empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&s_full[i], 1);
      mbar_init(&p_full[i], SM_THREADS);
      mbar_init(&o_full[i], 1);
      mbar_init(&o_free[i], 1 * SM_THREADS);
    }
    fence_barrier_init();
    // length of the atom: consecutive tokens of the same sequence, up to the next multiple of nt_tile
    int nt = 1;
    while (nt < p.nt_tile && t0 + nt < p.T && p.tok_seq[t0 + nt] == seq && p.tok_pos[t0 + nt] == pos0 + nt && ((pos0 + nt) % p.nt_tile) != 0) ++nt;
    *s_nt = nt;
  }
  if (warp == 1) {
    tmem_alloc<1>(tmem_ptr, 512);
    tmem_relinquish<1>();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  const uint32_t TM_S = tmem_base;
  const uint32_t TM_PV = tmem_base + 256;
  const int nt = *s_nt;
  // pages visible to the last token of the atom, and this CTA's slice of them
  const int n_pages = (pos0 + nt + BN - 1) / BN;
  const int per_split = (n_pages + p.n_splits - 1) / p.n_splits;
  const int pg0 = split * per_split;
  const int pg1 = min(n_pages, pg0 + per_split);
  const int n_blocks = max(0, pg1 - pg0);
  const bool partial = p.n_splits > 1;

  if (n_blocks > 0) {
    if (warp == 0) {
      if (lane == 0) {
        mbar_expect_tx(q_full, SMEM::TILE);
#pragma unroll
        for (int c = 0; c < NCH; ++c) tma_load_3d(&tmQ, q_full, smem + SMEM::Q_OFF + c * CHUNK_BYTES, c * 64, kvh * p.group, t0);
        const int* bt = p.block_table + static_cast<size_t>(seq) * p.max_blocks;
        const int sub = BN / p.block_size;  // cache blocks per 128-key page
        const int last_blk = (pos0 + nt - 1) / p.block_size;
        int st = 0, ph = 0;
        for (int j = 0; j < n_blocks; ++j) {
#pragma unroll
          for (int is_v = 0; is_v < 2; ++is_v) {
            uint64_t* full = is_v ? &v_full[st] : &k_full[st];
            mbar_wait(is_v ? &v_empty[st] : &k_empty[st], ph ^ 1);
            mbar_expect_tx(full, SMEM::TILE);
            uint8_t* tile = smem + (is_v ? SMEM::V_OFF : SMEM::K_OFF) + st * SMEM::TILE;
            for (int s = 0; s < sub; ++s) {
              // blocks past the sequence's extent are never unmasked, but their bytes must be finite: re-read a valid block
              const int logical = min((pg0 + j) * sub + s, last_blk);
              const int blk = bt[logical];
              const int row_off = s * p.block_size * 128;
#pragma unroll
              for (int c = 0; c < NCH; ++c) tma_load_4d(&tmKV, full, tile + c * CHUNK_BYTES + row_off, c * 64, kvh, 0, blk * 2 + is_v);
            }
          }
          if (++st == NS) { st = 0; ph ^= 1; }
        }
      }
    } else if (warp == 1) {
      if (lane == 0) {
        constexpr uint32_t idesc_qk = umma_idesc(BM, BN, 1, 1, 0, 0);
        constexpr uint32_t idesc_pv = umma_idesc(BM, D, 1, 1, 0, 1);
        const uint32_t q_s = smem_u32(smem + SMEM::Q_OFF);
        // S / P / partial-O buffers alternate with the page parity; K and V tiles cycle through their NS-deep rings.
        // The tensor pipe executes this thread's MMAs in issue order, so QK(j+1) cannot overwrite S buffer (j+1)&1 before
        // PV(j-1), which reads P out of the same columns, has retired.
        auto issue_qk = [&](int j, int st, int ph) {
          const int b = j & 1;
          mbar_wait(&k_full[st], ph);
          tc_fence_after();
          const uint32_t k_s = smem_u32(smem + SMEM::K_OFF + st * SMEM::TILE);
#pragma unroll
          for (int kk = 0; kk < D / 16; ++kk) {
            const uint32_t off = (kk >> 2) * CHUNK_BYTES + (kk & 3) * 32;
            umma_f16<1>(TM_S + b * 128, umma_smem_desc(q_s + off, 16, 1024), umma_smem_desc(k_s + off, 16, 1024), idesc_qk,
                        kk != 0 ? 1u : 0u);
          }
          umma_commit<1>(&s_full[b]);
          umma_commit<1>(&k_empty[st]);
        };
        auto issue_pv = [&](int j, int st, int ph) {
          const int b = j & 1;
          mbar_wait(&v_full[st], ph);
          mbar_wait(&p_full[b], (j >> 1) & 1);
          if (j >= 2) mbar_wait(&o_free[b], ((j >> 1) - 1) & 1);
          tc_fence_after();
          const uint32_t v_s = smem_u32(smem + SMEM::V_OFF + st * SMEM::TILE);
#pragma unroll
          for (int kk = 0; kk < BN / 16; ++kk)
            // P of keys [16 kk, 16 kk + 16): the softmax half that owns them packed it into the first 32 columns of its 64
            umma_f16_ts(TM_PV + b * 128, TM_S + b * 128 + (kk >> 2) * 64 + (kk & 3) * 8,
                        umma_smem_desc(v_s + kk * 2048, CHUNK_BYTES, 1024), idesc_pv, kk != 0 ? 1u : 0u);
          umma_commit<1>(&o_full[b]);
          umma_commit<1>(&v_empty[st]);
        };
        mbar_wait(q_full, 0);
        issue_qk(0, 0, 0);
        int st = 0, ph = 0;  // ring position of page j
        for (int j = 0; j < n_blocks; ++j) {
          int st1 = st + 1, ph1 = ph;
          if (st1 == NS) { st1 = 0; ph1 ^= 1; }
          if (j + 1 < n_blocks) issue_qk(j + 1, st1, ph1);
          issue_pv(j, st, ph);
          st = st1; ph = ph1;
        }
      }
    } else {
      // 8 softmax warps: warp quarter q4 (the TMEM lanes a warp may touch) x column half.  Thread == (row, half): it owns keys
      // [64 half, 64 half + 64) of the row's S and columns [D/2 half, D/2 half + D/2) of its O.  One warp per scheduler doing
      // a whole row was latency-bound on its own instruction stream (~1500 dependent-ish instructions per page); two warps per
      // scheduler with half the work each overlap each other's latencies.  The halves agree on the page maximum through shared
      // memory (rounded up to bf16: any common reference >= the true maximum is exact for the result); sums merge at the end.
      const int q4 = warp & 3, half = (warp - 2) >> 2;
      const int row = q4 * 32 + lane;
      const uint32_t lane_off = static_cast<uint32_t>(q4 * 32) << 16;
      const int ti = row / p.group, g = row - ti * p.group;
      const bool live = ti < nt;
      // keys [0, limit) are visible.  Dead rows (past the atom) see everything: their operands are finite (TMA zero fill /
      // other tokens' rows, zero-initialised cache) and nothing of theirs is stored
      const int limit = live ? pos0 + ti + 1 : 0x3fffffff;
      constexpr int DH = D / 2;
      float o_acc[DH];
#pragma unroll
      for (int i = 0; i < DH; ++i) o_acc[i] = 0.f;
      float m_run = -INFINITY, l_run = 0.f;
      __nv_bfloat16* xmax = reinterpret_cast<__nv_bfloat16*>(smem + SMEM::X_OFF);  // [page parity][half][row]
      const uint32_t pair_bar = 1 + q4;

      // O (in the scale of the previous page's maximum) * alpha + partial PV of page j
      auto accumulate_pv = [&](int j, float alpha) {
        const int b = j & 1;
        mbar_wait(&o_full[b], (j >> 1) & 1);
        tc_fence_after();
#pragma unroll
        for (int c = 0; c < DH / 32; ++c) {
          uint32_t v[32];
          tmem_ld_32x32(TM_PV + b * 128 + half * DH + c * 32 + lane_off, v);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) o_acc[c * 32 + i] = fmaf(o_acc[c * 32 + i], alpha, __uint_as_float(v[i]));
        }
        tc_fence_before();
        mbar_arrive(&o_free[b]);
      };

      float alpha_prev = 1.f;
      for (int j = 0; j < n_blocks; ++j) {
        const int b = j & 1;
        mbar_wait(&s_full[b], (j >> 1) & 1);
        tc_fence_after();
        const int visible = limit - ((pg0 + j) * BN + half * 64);  // of this thread's 64 keys, [0, visible) count
        const bool all_visible = __all_sync(0xffffffffu, visible >= 64);  // every page but the diagonal one(s)
        const uint32_t s_addr = TM_S + b * 128 + half * 64 + lane_off;
        // pass 1: maximum of the thread's 64 scores (four independent chains)
        float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          uint32_t sv[32];
          tmem_ld_32x32(s_addr + c * 32, sv);
          tmem_ld_wait();
          if (all_visible) {
#pragma unroll
            for (int i = 0; i < 32; ++i) mx[i & 3] = fmaxf(mx[i & 3], __uint_as_float(sv[i]));
          } else {
#pragma unroll
            for (int i = 0; i < 32; ++i) mx[i & 3] = fmaxf(mx[i & 3], c * 32 + i >= visible ? -INFINITY : __uint_as_float(sv[i]));
          }
        }
        const __nv_bfloat16 mine = __float2bfloat16_ru(fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3])));
        xmax[(b * 2 + half) * 128 + row] = mine;
        asm volatile("bar.sync %0, 64;" ::"r"(pair_bar) : "memory");
        const float m_blk = fmaxf(__bfloat162float(mine), __bfloat162float(xmax[(b * 2 + (half ^ 1)) * 128 + row]));
        // a page with no visible key for this row leaves the running max alone (and -inf - -inf is never formed)
        const float m_new = fmaxf(m_run, m_blk * p.scale_log2);
        const float m_use = m_new == -INFINITY ? 0.f : m_new;
        const float alpha = m_run == -INFINITY ? 0.f : exp2f(m_run - m_use);
        float ls[4] = {0.f, 0.f, 0.f, 0.f};
        // pass 2: P chunk c (32 keys -> 16 packed columns) goes to columns [16 c, 16 c + 16) of the thread's own half of the
        // S buffer: always columns it has already consumed (chunk 0) or holds in registers (chunk c)
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          uint32_t sv[32];
          tmem_ld_32x32(s_addr + c * 32, sv);
          tmem_ld_wait();
          uint32_t pk[16];
          if (all_visible) {
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              const float e0 = ex2_fast(fmaf(__uint_as_float(sv[2 * i]), p.scale_log2, -m_use));
              const float e1 = ex2_fast(fmaf(__uint_as_float(sv[2 * i + 1]), p.scale_log2, -m_use));
              ls[i & 1] += e0;
              ls[2 + (i & 1)] += e1;
              pk[i] = pack_bf16x2(e0, e1);
            }
          } else {
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              const float v0 = fmaf(__uint_as_float(sv[2 * i]), p.scale_log2, -m_use);
              const float v1 = fmaf(__uint_as_float(sv[2 * i + 1]), p.scale_log2, -m_use);
              const float e0 = (c * 32 + 2 * i >= visible) ? 0.f : ex2_fast(v0);
              const float e1 = (c * 32 + 2 * i + 1 >= visible) ? 0.f : ex2_fast(v1);
              ls[i & 1] += e0;
              ls[2 + (i & 1)] += e1;
              pk[i] = pack_bf16x2(e0, e1);
            }
          }
          tmem_st_32x16(s_addr + c * 16, pk);
        }
        l_run = l_run * alpha + ((ls[0] + ls[1]) + (ls[2] + ls[3]));
        m_run = m_new;
        tmem_st_wait();
        tc_fence_before();
        mbar_arrive(&p_full[b]);
        if (j > 0) accumulate_pv(j - 1, alpha_prev);
        alpha_prev = alpha;
      }
      accumulate_pv(n_blocks - 1, alpha_prev);

      // the row's sum = both halves' sums (same sequence of maxima): exchange through the (now idle) max buffer
      float* xl = reinterpret_cast<float*>(smem + SMEM::X_OFF);  // [half][row]
      asm volatile("bar.sync %0, 64;" ::"r"(pair_bar) : "memory");
      xl[half * 128 + row] = l_run;
      asm volatile("bar.sync %0, 64;" ::"r"(pair_bar) : "memory");
      l_run += xl[(half ^ 1) * 128 + row];

      if (live) {
        const int t = t0 + ti, head = kvh * p.group + g;
        if (!partial) {
          const float inv_l = 1.f / l_run;
          __nv_bfloat16* dst = p.out + (static_cast<size_t>(t) * p.hq + head) * D + half * DH;
#pragma unroll
          for (int u = 0; u < DH / 8; ++u) {
            uint4 val;
            const float* o = &o_acc[u * 8];
            val.x = pack_bf16x2(o[0] * inv_l, o[1] * inv_l);
            val.y = pack_bf16x2(o[2] * inv_l, o[3] * inv_l);
            val.z = pack_bf16x2(o[4] * inv_l, o[5] * inv_l);
            val.w = pack_bf16x2(o[6] * inv_l, o[7] * inv_l);
            *reinterpret_cast<uint4*>(dst + u * 8) = val;
          }
        } else {
          const size_t slot = (static_cast<size_t>(t) * p.hq + head) * p.n_splits + split;
          float* dst = p.part_o + slot * D + half * DH;
#pragma unroll
          for (int u = 0; u < DH / 4; ++u)
            *reinterpret_cast<float4*>(dst + u * 4) = make_float4(o_acc[u * 4], o_acc[u * 4 + 1], o_acc[u * 4 + 2], o_acc[u * 4 + 3]);
          if (half == 0) {
            p.part_ml[slot * 2] = m_run;
            p.part_ml[slot * 2 + 1] = l_run;
          }
        }
      }
    }
  } else if (partial && warp >= 2 && warp < 6) {
    // this split has no pages (short context): publish an empty partial for every live row
    const int row = (warp & 3) * 32 + lane;
    const int ti = row / p.group, g = row - ti * p.group;
    if (ti < nt) {
      const size_t slot = (static_cast<size_t>(t0 + ti) * p.hq + kvh * p.group + g) * p.n_splits + split;
      p.part_ml[slot * 2] = -INFINITY;
      p.part_ml[slot * 2 + 1] = 0.f;
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<1>(tmem_base, 512);
  }
}

// out[t][h][:] = sum_s 2^(m_s - M) O_s / sum_s 2^(m_s - M) l_s ; one warp per (token, head)
template <int D>
__global__ void combine_splits_kernel(const float* __restrict__ part_o, const float* __restrict__ part_ml, __nv_bfloat16* __restrict__ out,
                                      int rows, int n_splits) {
  const int wid = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (wid >= rows) return;
  constexpr int EPL = D / 32;
  float M = -INFINITY;
  for (int s = 0; s < n_splits; ++s) M = fmaxf(M, part_ml[(static_cast<size_t>(wid) * n_splits + s) * 2]);
  float acc[EPL];
#pragma unroll
  for (int e = 0; e < EPL; ++e) acc[e] = 0.f;
  float L = 0.f;
  for (int s = 0; s < n_splits; ++s) {
    const size_t slot = static_cast<size_t>(wid) * n_splits + s;
    const float m = part_ml[slot * 2];
    if (m == -INFINITY) continue;
    const float w = exp2f(m - M);
    L = fmaf(w, part_ml[slot * 2 + 1], L);
#pragma unroll
    for (int e = 0; e < EPL; ++e) acc[e] = fmaf(w, part_o[slot * D + lane * EPL + e], acc[e]);
  }
  const float inv = 1.f / L;
#pragma unroll
  for (int e = 0; e < EPL; ++e) out[static_cast<size_t>(wid) * D + lane * EPL + e] = __float2bfloat16_rn(acc[e] * inv);
}

template <int D>
static int launch(const void* qkv, const void* kv_cache, void* out, const int* tok_seq, const int* tok_pos, const int* block_table,
                  int T, int hq, int hkv, int block_size, int max_blocks, int num_blocks, float scale, int max_atoms, int n_splits,
                  void* scratch, cudaStream_t stream) {
  using SMEM = Smem<D>;
  const int group = hq / hkv, nt_tile = BM / group;
  const int heads = hq + 2 * hkv;
  CUtensorMap tmQ, tmKV;
  {
    const uint64_t dims[3] = {static_cast<uint64_t>(D), static_cast<uint64_t>(heads), static_cast<uint64_t>(T)};
    const uint64_t strides[2] = {static_cast<uint64_t>(D) * 2, static_cast<uint64_t>(heads) * D * 2};
    const uint32_t box[3] = {64, static_cast<uint32_t>(group), static_cast<uint32_t>(nt_tile)};
    int rc = make_tensor_map(&tmQ, TMA_BF16, 3, qkv, dims, strides, box, TMA_SW_128);
    if (rc) return rc;
  }
  {
    const uint64_t dims[4] = {static_cast<uint64_t>(D), static_cast<uint64_t>(hkv), static_cast<uint64_t>(block_size),
                              static_cast<uint64_t>(num_blocks) * 2};
    const uint64_t strides[3] = {static_cast<uint64_t>(D) * 2, static_cast<uint64_t>(hkv) * D * 2,
                                 static_cast<uint64_t>(block_size) * hkv * D * 2};
    const uint32_t box[4] = {64, 1, static_cast<uint32_t>(block_size), 1};
    int rc = make_tensor_map(&tmKV, TMA_BF16, 4, kv_cache, dims, strides, box, TMA_SW_128);
    if (rc) return rc;
  }
  // scratch: [max_atoms + 1 ints (atoms, count)] then, for split runs, the partials
  int* atoms = static_cast<int*>(scratch);
  int* n_atoms = atoms + max_atoms;
  const size_t head_ints = (static_cast<size_t>(max_atoms) + 1 + 3) / 4 * 4;
  float* part_o = reinterpret_cast<float*>(atoms + head_ints);
  float* part_ml = part_o + static_cast<size_t>(T) * hq * n_splits * D;
  DSB_CHECK_CUDA(cudaMemsetAsync(n_atoms, 0, sizeof(int), stream));
  find_atoms_kernel<<<(T + 255) / 256, 256, 0, stream>>>(tok_seq, tok_pos, T, nt_tile, atoms, n_atoms, max_atoms);
  auto kern = paged_flash_kernel<D>;
  static bool attr_set = false;
  if (!attr_set) {
    DSB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM::TOTAL));
    attr_set = true;
  }
  Params p;
  p.tok_seq = tok_seq; p.tok_pos = tok_pos; p.block_table = block_table; p.atoms = atoms; p.n_atoms = n_atoms;
  p.out = static_cast<__nv_bfloat16*>(out); p.part_o = part_o; p.part_ml = part_ml;
  p.T = T; p.hq = hq; p.hkv = hkv; p.group = group; p.nt_tile = nt_tile; p.block_size = block_size; p.max_blocks = max_blocks;
  p.n_splits = n_splits; p.scale_log2 = scale * 1.4426950408889634f;
  dim3 grid(max_atoms, hkv, n_splits);
  kern<<<grid, NUM_THREADS, SMEM::TOTAL, stream>>>(tmQ, tmKV, p);
  DSB_CHECK_CUDA(cudaGetLastError());
  if (n_splits > 1) {
    const int rows = T * hq;
    combine_splits_kernel<D><<<(rows * 32 + 255) / 256, 256, 0, stream>>>(part_o, part_ml, static_cast<__nv_bfloat16*>(out), rows, n_splits);
    DSB_CHECK_CUDA(cudaGetLastError());
  }
  return DSB_OK;
}

}  // namespace paged
}  // namespace dsb

// Bytes of scratch dsb200_paged_flash_attention needs: atom list + count, plus the split-KV partials when n_splits > 1.
DSB_EXPORT long long dsb200_paged_flash_scratch_bytes(int n_tokens, int hq, int head_dim, int max_atoms, int n_splits) {
  const long long head_ints = (static_cast<long long>(max_atoms) + 1 + 3) / 4 * 4;
  long long bytes = head_ints * 4;
  if (n_splits > 1) bytes += static_cast<long long>(n_tokens) * hq * n_splits * (head_dim + 2) * 4;
  return bytes;
}

// out[t][hq*D] = attention of ragged token t over keys [0, tok_pos[t]] of its sequence, KV read from the blocked cache
// [num_blocks][2][block_size][hkv][D] (bf16) through the block table.  max_atoms >= number of atoms of the batch
// (<= n_tokens; ceil(n_tokens / (128/group)) + number of sequences is enough).  n_splits >= 1 splits the KV range of every atom.
DSB_EXPORT int dsb200_paged_flash_attention(const void* qkv, const void* kv_cache, void* out, const int* tok_seq, const int* tok_pos,
                                            const int* block_table, int n_tokens, int hq, int hkv, int head_dim, int block_size,
                                            int max_blocks_per_seq, int num_blocks, float scale, int max_atoms, int n_splits,
                                            void* scratch, cudaStream_t stream) {
  using namespace dsb;
  if (n_tokens == 0) return DSB_OK;
  DSB_REQUIRE(head_dim == 64 || head_dim == 128, "paged_flash: head_dim %d unsupported (64 or 128)", head_dim);
  DSB_REQUIRE(hq % hkv == 0 && 128 % (hq / hkv) == 0, "paged_flash: q/kv head ratio %d must divide 128", hq / hkv);
  DSB_REQUIRE(block_size >= 8 && block_size <= 128 && 128 % block_size == 0, "paged_flash: KV block size %d must divide 128", block_size);
  DSB_REQUIRE(max_atoms >= 1 && n_splits >= 1 && n_splits <= 64 && scratch != nullptr, "paged_flash: bad max_atoms / n_splits / scratch");
  if (head_dim == 128)
    return paged::launch<128>(qkv, kv_cache, out, tok_seq, tok_pos, block_table, n_tokens, hq, hkv, block_size, max_blocks_per_seq,
                              num_blocks, scale, max_atoms, n_splits, scratch, stream);
  return paged::launch<64>(qkv, kv_cache, out, tok_seq, tok_pos, block_table, n_tokens, hq, hkv, block_size, max_blocks_per_seq,
                           num_blocks, scale, max_atoms, n_splits, scratch, stream);
}

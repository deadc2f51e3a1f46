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
//   warps 2-17 softmax    : thread == (row, quarter of the columns): S from TMEM, per-row causal limit (position + 1), online
//                           softmax in exp2 domain, P (bf16, two per column) stored back over the thread's own S columns, its
//                           quarter of O kept in registers
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
constexpr int NUM_THREADS = 576;  // TMA warp, MMA warp, 16 softmax warps
constexpr int SM_THREADS = 512;
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
  static constexpr int X_OFF = BAR_OFF + 400;  // 1 KB: page maxima exchanged between the four column parts of a row
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
  int window;          // local attention: keys [pos - window + 1, pos] are visible; 0 = the whole history
  const float* alibi;  // optional [hq] ALiBi slopes: score += slope[head] * key position (BLOOM / MPT / falcon-rw); nullptr = none
  long long* trace;  // diagnostics only (DSB200_PF_TRACE=1): clock64 stamps of the first CTA with >= 8 pages, [page][8 events]
};
// events: 0 K landed (MMA thread)  1 P ready seen by the MMA thread  2 S ready seen by softmax  3 P published  4 partial O read
#define DSB_PF_TRACE(ev, j) \
  if (tracing && (j) < 64) p.trace[8 + (j) * 8 + (ev)] = clock64();

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
  uint64_t* o_full = p_full + 2;           // [0] only: phase j completes when PV(j) has retired
  uint64_t* o_free = o_full + 2;           // [0]: "Q is in TMEM" (softmax warps -> MMA thread); [1] unused
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
      mbar_init(&o_free[i], 128 * (D / 32));
    }
    fence_barrier_init();
  }
  if (warp == 2) {
    // length of the atom: consecutive tokens of the same sequence, up to the next multiple of nt_tile.  Lane i tests token
    // t0 + i (one round of loads for 32 candidates instead of a serial chain of dependent global loads)
    int nt = p.nt_tile;
    for (int base = 0; base < p.nt_tile; base += 32) {
      const int i = base + lane;
      const bool ok = i == 0 || (i < p.nt_tile && t0 + i < p.T && p.tok_seq[t0 + i] == seq && p.tok_pos[t0 + i] == pos0 + i &&
                                 ((pos0 + i) % p.nt_tile) != 0);
      const uint32_t bad = __ballot_sync(0xffffffffu, !ok);
      if (bad) {
        nt = base + __ffs(bad) - 1;
        break;
      }
    }
    if (lane == 0) *s_nt = nt;
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
  const uint32_t TM_O = tmem_base + 256;  // O accumulates here across the pages of the CTA
  const uint32_t TM_Q = tmem_base + 384;  // Q as the A operand of QK^T: two bf16 per column, D/2 columns
  const int nt = *s_nt;
  // pages visible to the last token of the atom, and this CTA's slice of them
  const int n_pages = (pos0 + nt + BN - 1) / BN;
  const int per_split = (n_pages + p.n_splits - 1) / p.n_splits;
  const int pg0 = split * per_split;
  const int pg1 = min(n_pages, pg0 + per_split);
  const int n_blocks = max(0, pg1 - pg0);
  const bool partial = p.n_splits > 1;
  bool tracing = false;
  if (p.trace != nullptr) {  // one CTA with enough pages claims the trace buffer
    int* claim = reinterpret_cast<int*>(bars + 48);
    if (threadIdx.x == 0) *claim = (n_blocks >= 8 && atomicCAS(reinterpret_cast<int*>(p.trace), 0, 1) == 0) ? 1 : 0;
    __syncthreads();
    tracing = *claim != 0;
    if (tracing && threadIdx.x == 0) p.trace[1] = clock64();
  }

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
      // The whole warp stays converged and one elected lane issues: inside an `if (lane == 0)` region every descriptor and
      // TMEM address is a per-thread value that the compiler moves to the uniform registers with an elect / broadcast loop
      // per instruction, ~81 cycles per tcgen05.mma against a 64-cycle floor for 128 x 128 x 16 (tools/microbench/mma_rate.cu).
      // Descriptors are built once; a k-step or a ring stage is a 32-bit add to the low word.
      const bool leader = elect_one();
      constexpr uint32_t idesc_qk = umma_idesc(BM, BN, 1, 1, 0, 0);
      constexpr uint32_t idesc_pv = umma_idesc(BM, D, 1, 1, 0, 1);
      const uint32_t tm_s = __shfl_sync(0xffffffffu, TM_S, 0), tm_o = tm_s + 256, tm_q = tm_s + 384;
      const uint64_t k_desc0 = umma_smem_desc(smem_u32(smem + SMEM::K_OFF), 16, 1024);
      const uint64_t v_desc0 = umma_smem_desc(smem_u32(smem + SMEM::V_OFF), CHUNK_BYTES, 1024);
      // S / P buffers alternate with the page parity; K and V tiles cycle through their NS-deep rings.  The tensor pipe
      // executes this warp's MMAs in issue order, so QK(j+1) cannot overwrite S buffer (j+1)&1 before PV(j-1), which reads P
      // out of the same columns, has retired.
      auto issue_qk = [&](int j, int st, int ph) {
        const int b = j & 1;
        mbar_wait_lean(&k_full[st], ph);
        if (leader) { DSB_PF_TRACE(0, j) }
        tc_fence_after();
        if (leader) {
          const uint64_t kd = k_desc0 + static_cast<uint32_t>((st * SMEM::TILE) >> 4);
#pragma unroll
          for (int kk = 0; kk < D / 16; ++kk)
            // A = Q from TMEM (stationary for the CTA): only the K tile crosses the shared-memory port
            umma_f16_ts(tm_s + b * 128, tm_q + kk * 8, kd + static_cast<uint32_t>(((kk >> 2) * CHUNK_BYTES + (kk & 3) * 32) >> 4), idesc_qk,
                        kk != 0 ? 1u : 0u);
          umma_commit<1>(&s_full[b]);
          umma_commit<1>(&k_empty[st]);
        }
      };
      auto issue_pv = [&](int j, int st, int ph) {
        const int b = j & 1;
        mbar_wait_lean(&v_full[st], ph);
        mbar_wait_lean(&p_full[b], (j >> 1) & 1);
        if (leader) { DSB_PF_TRACE(1, j) }
        tc_fence_after();
        if (leader) {
          const uint64_t vd = v_desc0 + static_cast<uint32_t>((st * SMEM::TILE) >> 4);
#pragma unroll
          for (int kk = 0; kk < BN / 16; ++kk)
            // P of keys [16 kk, 16 kk + 16): the softmax part that owns them packed it into the first 16 columns of its 32
            umma_f16_ts(tm_o, tm_s + b * 128 + (kk >> 1) * 32 + (kk & 1) * 8, vd + static_cast<uint32_t>((kk * 2048) >> 4), idesc_pv,
                        (j | kk) != 0 ? 1u : 0u);
          umma_commit<1>(&o_full[0]);  // phase j completes when PV(j) has retired
          umma_commit<1>(&v_empty[st]);
        }
      };
      mbar_wait_lean(&o_free[0], 0);  // Q copied into TMEM
      tc_fence_after();
      issue_qk(0, 0, 0);
      int st = 0, ph = 0;  // ring position of page j
      for (int j = 0; j < n_blocks; ++j) {
        int st1 = st + 1, ph1 = ph;
        if (st1 == NS) { st1 = 0; ph1 ^= 1; }
        if (j + 1 < n_blocks) issue_qk(j + 1, st1, ph1);
        issue_pv(j, st, ph);
        st = st1; ph = ph1;
      }
    } else {
      // 16 softmax warps: warp quarter q4 (the TMEM lanes a warp may touch) x column quarter `part`.  Thread == (row, part): it
      // owns keys [32 part, 32 part + 32) of the row's S and columns [D/4 part, D/4 part + D/4) of its O.  Four warps per
      // scheduler, each with a quarter of a row, hide one another's TMEM / MUFU / dependent-issue latencies (one warp per
      // scheduler owning whole rows ran at 0.3 IPC and ~2 us a page).  The four parts agree on the page maximum through shared
      // memory (rounded up to bf16: any common reference >= the true maximum is exact for the result); sums merge at the end.
      const int q4 = warp & 3, part = (warp - 2) >> 2;
      const int row = q4 * 32 + lane;
      const uint32_t lane_off = static_cast<uint32_t>(q4 * 32) << 16;
      const int ti = row / p.group, g = row - ti * p.group;
      const bool live = ti < nt;
      // keys [0, limit) are visible.  Dead rows (past the atom) see everything: their operands are finite (TMA zero fill /
      // other tokens' rows, zero-initialised cache) and nothing of theirs is stored
      const int limit = live ? pos0 + ti + 1 : 0x3fffffff;
      const int first = (live && p.window > 0) ? max(0, pos0 + ti + 1 - p.window) : 0;  // first visible key (local attention)
      constexpr int DQ = D / 4;
      // m_run is the reference maximum P and O are expressed in.  It follows the true running maximum lazily: only when that
      // has grown by more than 2^8 is O (in TMEM, accumulated by the tensor core) read, rescaled and written back.  Reading
      // the partial O of every page back into registers cost as much TMEM bandwidth as reading S (64 B/clk/SM, the limiter).
      float m_run = -INFINITY, l_run = 0.f;
      // ALiBi: the biased score is formed in the exp2 domain right after the mask (s * scale_log2 + slope_log2 * key), after which
      // the running maximum / exponent arithmetic below uses a scale of 1
      const bool has_alibi = p.alibi != nullptr;
      const float slope_l2 = (has_alibi && live) ? p.alibi[kvh * p.group + g] * 1.4426950408889634f : 0.f;
      const float sl2 = has_alibi ? 1.f : p.scale_log2;
      __nv_bfloat16* xmax = reinterpret_cast<__nv_bfloat16*>(smem + SMEM::X_OFF);  // [part][row]
      const uint32_t bar_a = 1 + q4, bar_b = 5 + q4;  // the 4 warps (128 threads) that share the rows of this quarter
      const uint32_t o_addr = TM_O + part * DQ + lane_off;
      const bool warp_live = q4 * 32 < nt * p.group;  // rows are packed from 0: a warp past them has nothing to compute

      // Q: swizzled smem tile (TMA) -> TMEM.  Parts 0 .. D/32-1 of every row quarter move 32 elements (16 packed columns) each
      if (part < D / 32) {
        mbar_wait(q_full, 0);
        const uint8_t* qrow = smem + SMEM::Q_OFF + (part >> 1) * CHUNK_BYTES + row * 128;
        uint32_t qv[16];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const uint4 x = *reinterpret_cast<const uint4*>(qrow + ((((part & 1) * 4 + u) ^ (row & 7)) << 4));
          qv[u * 4] = x.x; qv[u * 4 + 1] = x.y; qv[u * 4 + 2] = x.z; qv[u * 4 + 3] = x.w;
        }
        tmem_st_32x16(TM_Q + part * 16 + lane_off, qv);
        tmem_st_wait();
        tc_fence_before();
        mbar_arrive(&o_free[0]);
      }
      if (!warp_live) {
        // keep the P barrier's arrival count: one arrival per page, paced by the completion of that page's phase
        for (int j = 0; j < n_blocks; ++j) {
          mbar_arrive(&p_full[j & 1]);
          mbar_wait(&p_full[j & 1], (j >> 1) & 1);
        }
      } else {
      for (int j = 0; j < n_blocks; ++j) {
        const int b = j & 1;
        mbar_wait(&s_full[b], (j >> 1) & 1);
        if (warp == 4 && lane == 0) { DSB_PF_TRACE(2, j) }
        tc_fence_after();
        const int visible = limit - ((pg0 + j) * BN + part * 32);  // of this thread's 32 keys, [0, visible) count
        const uint32_t s_addr = TM_S + b * 128 + part * 32 + lane_off;
        uint32_t sv[32];  // the thread's scores stay in registers for both passes
        tmem_ld_32x32(s_addr, sv);
        tmem_ld_wait();
        const int hidden = first - ((pg0 + j) * BN + part * 32);  // keys [0, hidden) of the thread's 32 precede the local window
        if (!__all_sync(0xffffffffu, visible >= 32 && hidden <= 0)) {  // only the diagonal page(s) / the window's first page(s)
#pragma unroll
          for (int i = 0; i < 32; ++i)
            if (i >= visible || i < hidden) sv[i] = 0xff800000u;  // -inf: exp2 gives exactly 0 below
        }
        if (has_alibi) {
          const float key0 = static_cast<float>((pg0 + j) * BN + part * 32);
#pragma unroll
          for (int i = 0; i < 32; ++i) sv[i] = __float_as_uint(fmaf(__uint_as_float(sv[i]), p.scale_log2, slope_l2 * (key0 + i)));
        }
        float mx[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
        for (int i = 0; i < 32; ++i) mx[i & 3] = fmaxf(mx[i & 3], __uint_as_float(sv[i]));
        const __nv_bfloat16 mine = __float2bfloat16_ru(fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3])));
        if (j > 0) asm volatile("bar.sync %0, 128;" ::"r"(bar_b) : "memory");  // everyone has read the previous page's maxima
        xmax[part * 128 + row] = mine;
        asm volatile("bar.sync %0, 128;" ::"r"(bar_a) : "memory");
        const float m_blk = fmaxf(fmaxf(__bfloat162float(xmax[row]), __bfloat162float(xmax[128 + row])),
                                  fmaxf(__bfloat162float(xmax[256 + row]), __bfloat162float(xmax[384 + row])));
        // a page with no visible key for this row leaves the reference alone (and -inf - -inf is never formed)
        const float m_cand = fmaxf(m_run, m_blk * sl2);
        const bool fresh = m_run == -INFINITY;  // nothing accumulated for this row yet: O row and l are exactly 0
        const float m_new = (fresh || m_cand - m_run > 8.f) ? m_cand : m_run;
        const bool rescale = !fresh && m_new != m_run && j > 0;
        if (__any_sync(0xffffffffu, rescale)) {  // identical decision in the 4 warps that share these rows
          const float alpha = rescale ? exp2f(m_run - m_new) : 1.f;
          mbar_wait(&o_full[0], (j - 1) & 1);  // PV(j-1) has retired; PV(j) cannot start before this thread publishes P(j)
          tc_fence_after();
          uint32_t v[DQ];
          if constexpr (DQ == 32) tmem_ld_32x32(o_addr, v);
          else tmem_ld_32x16(o_addr, v);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < DQ; ++i) v[i] = __float_as_uint(__uint_as_float(v[i]) * alpha);
          if constexpr (DQ == 32) tmem_st_32x32(o_addr, v);
          else tmem_st_32x16(o_addr, v);
          l_run *= alpha;
        }
        m_run = m_new;
        const float m_use = m_new == -INFINITY ? 0.f : m_new;
        float ls[2];
        // P (32 keys -> 16 packed columns) goes to the first 16 columns of the thread's own quarter of the S buffer
        // The MUFU unit (16 ex2 / clk / SM: 1024 cycles for a 128 x 128 page) is the limiter of this loop: 3 pairs of every 8
        // take the polynomial exp2 on the FMA pipe, in packed fp32x2 arithmetic
        uint32_t pk[16];
        const uint64_t sc2 = pack2(sl2, sl2), negm2 = pack2(-m_use, -m_use);
        uint64_t l2a = 0ull, l2b = 0ull;
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const uint64_t x2 = fma2(pack2(__uint_as_float(sv[2 * i]), __uint_as_float(sv[2 * i + 1])), sc2, negm2);
          const uint64_t e2 = ((i & 7) < 3) ? exp2_poly2(x2) : exp2_mufu2(x2);
          if (i & 1) l2b = add2(l2b, e2); else l2a = add2(l2a, e2);
          float e0, e1;
          unpack2(e2, e0, e1);
          pk[i] = pack_bf16x2(e0, e1);
        }
        tmem_st_32x16(s_addr, pk);
        unpack2(add2(l2a, l2b), ls[0], ls[1]);
        l_run += ls[0] + ls[1];
        tmem_st_wait();
        tc_fence_before();
        mbar_arrive(&p_full[b]);
        if (warp == 4 && lane == 0) { DSB_PF_TRACE(3, j) }
      }
      mbar_wait(&o_full[0], (n_blocks - 1) & 1);
      tc_fence_after();
      uint32_t o_bits[DQ];
      if constexpr (DQ == 32) tmem_ld_32x32(o_addr, o_bits);
      else tmem_ld_32x16(o_addr, o_bits);
      tmem_ld_wait();
      float o_acc[DQ];
#pragma unroll
      for (int i = 0; i < DQ; ++i) o_acc[i] = __uint_as_float(o_bits[i]);

      // the row's sum = the four parts' sums (same sequence of maxima).  The last partial product has been read, so every MMA
      // and every TMA load of this CTA has retired: the first K tile is free to carry the exchange
      float* xl = reinterpret_cast<float*>(smem + SMEM::K_OFF);  // [part][row]
      xl[part * 128 + row] = l_run;
      asm volatile("bar.sync %0, 128;" ::"r"(bar_a) : "memory");
      l_run = (xl[row] + xl[128 + row]) + (xl[256 + row] + xl[384 + row]);

      if (live) {
        const int t = t0 + ti, head = kvh * p.group + g;
        if (!partial) {
          const float inv_l = 1.f / l_run;
          __nv_bfloat16* dst = p.out + (static_cast<size_t>(t) * p.hq + head) * D + part * DQ;
#pragma unroll
          for (int u = 0; u < DQ / 8; ++u) {
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
          float* dst = p.part_o + slot * D + part * DQ;
#pragma unroll
          for (int u = 0; u < DQ / 4; ++u)
            *reinterpret_cast<float4*>(dst + u * 4) = make_float4(o_acc[u * 4], o_acc[u * 4 + 1], o_acc[u * 4 + 2], o_acc[u * 4 + 3]);
          if (part == 0) {
            p.part_ml[slot * 2] = m_run;
            p.part_ml[slot * 2 + 1] = l_run;
          }
        }
      }
      }  // warp_live
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

// ================================================================================================ decode (small atoms)
// Transposed formulation for atoms of at most 16 rows (decode steps: 1 token x `group` heads).  In the kernel above a decode
// atom fills 8 of the 128 MMA rows, yet the softmax warps of that lane quarter still execute a full 32 x 128 tile per page --
// all on one scheduler -- and the tensor core a full 128 x 128 x 128 product: 1.7 us a page, 54 % of HBM.  Here the keys take
// the M dimension:
//     S^T [128 keys x 16] = K_page (smem, K-major A)  x  Q^T (16 q rows of the atom, smem, K-major B)        8 MMAs of 128 x 16 x 16
//     O^T [128 d    x 16] += V_page^T (smem, MN-major A)  x  P^T (smem, K-major B [16 rows][128 keys])       8 MMAs of 128 x 16 x 16
// thread == (key, 4 of the 16 columns): every lane live, the work spread over all four schedulers.  Per-column
// maxima are reduced across the 128 keys (shuffles + one shared-memory exchange); the sums stay per-thread partials until
// the end.  O^T accumulates in TMEM and is rescaled lazily (reference maximum moves only by > 2^8).  head_dim 128 only.
constexpr int NQ = 16;              // q rows (columns of S^T / O^T) per atom
constexpr int DEC_THREADS = 576;    // TMA warp, MMA warp, 16 softmax warps
struct DecSmem {
  static constexpr int TILE = 2 * CHUNK_BYTES;       // one K or V page: [128 keys][128 d] as two 64-column chunks
  static constexpr int STAGES = 3;
  static constexpr int Q_BYTES = 2 * NQ * 128;       // [2 chunks][16 rows][128 B]
  static constexpr int P_BYTES = 2 * NQ * 128;       // [2 key chunks][16 rows][128 B]
  static constexpr int K_OFF = 0;
  static constexpr int V_OFF = K_OFF + STAGES * TILE;
  static constexpr int Q_OFF = V_OFF + STAGES * TILE;
  static constexpr int P_OFF = Q_OFF + Q_BYTES;      // two buffers
  static constexpr int BAR_OFF = P_OFF + 2 * P_BYTES;
  static constexpr int X_OFF = BAR_OFF + 256;        // [2 parities][4 warps][16] page maxima, then the final sums
  static constexpr int TOTAL = X_OFF + 1024 + 1024;
};

__global__ void __launch_bounds__(DEC_THREADS, 1)
paged_decode_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmKV, const Params p) {
  using SMEM = DecSmem;
  constexpr int D = 128, NS = SMEM::STAGES;
  // One CTA per token: it runs only when its token opens an atom (first of the batch / of its sequence in the batch / of a
  // 16/group-aligned run).  Decode batches are a few hundred tokens at most, and this drops the atom-list kernel and its
  // counter memset from the launch chain (they cost more than a tenth of a 64-sequence decode step).
  const int t0 = blockIdx.x;
  const int seq = p.tok_seq[t0], pos0 = p.tok_pos[t0];
  if (t0 > 0 && p.tok_seq[t0 - 1] == seq && p.tok_pos[t0 - 1] + 1 == pos0 && (pos0 % p.nt_tile) != 0) return;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + SMEM::BAR_OFF);
  uint64_t* q_full = bars + 0;
  uint64_t* k_full = bars + 1;       // [NS]
  uint64_t* v_full = k_full + NS;    // [NS]
  uint64_t* k_empty = v_full + NS;   // [NS]
  uint64_t* v_empty = k_empty + NS;  // [NS]
  uint64_t* s_full = v_empty + NS;   // [2]
  uint64_t* p_full = s_full + 2;     // [2] softmax -> MMA: P^T(j) is in shared memory
  uint64_t* p_free = p_full + 2;     // [2] MMA -> softmax: PV(j) has read P^T buffer j & 1
  uint64_t* o_full = p_free + 2;     // phase j completes when PV(j) has retired
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(o_full + 1);
  int* s_nt = reinterpret_cast<int*>(o_full + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int kvh = blockIdx.y, split = blockIdx.z;

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
      mbar_init(&p_full[i], DEC_THREADS - 64);
      mbar_init(&p_free[i], 1);
    }
    mbar_init(o_full, 1);
    fence_barrier_init();
  }
  if (warp == 2) {  // atom length, as in the kernel above (nt_tile = 16 / group here)
    int nt = p.nt_tile;
    const int i = lane;
    const bool ok = i == 0 || (i < p.nt_tile && t0 + i < p.T && p.tok_seq[t0 + i] == seq && p.tok_pos[t0 + i] == pos0 + i &&
                               ((pos0 + i) % p.nt_tile) != 0);
    const uint32_t bad = __ballot_sync(0xffffffffu, !ok);
    if (bad) nt = __ffs(bad) - 1;
    if (lane == 0) *s_nt = nt;
  }
  if (warp == 1) {
    tmem_alloc<1>(tmem_ptr, 64);
    tmem_relinquish<1>();
  }
  // P^T buffers start zeroed: rows of q columns past the atom are never written
  for (int i = threadIdx.x; i < 2 * SMEM::P_BYTES / 16; i += DEC_THREADS)
    reinterpret_cast<uint4*>(smem + SMEM::P_OFF)[i] = make_uint4(0, 0, 0, 0);
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  const int nt = *s_nt;
  const int nq = nt * p.group;  // live columns
  const int n_pages = (pos0 + nt + BN - 1) / BN;
  const int per_split = (n_pages + p.n_splits - 1) / p.n_splits;
  const int pg0 = split * per_split;
  const int pg1 = min(n_pages, pg0 + per_split);
  const int n_blocks = max(0, pg1 - pg0);
  const bool partial = p.n_splits > 1;
  bool tracing = false;
  if (p.trace != nullptr) {
    int* claim = reinterpret_cast<int*>(bars + 30);
    if (threadIdx.x == 0) *claim = (n_blocks >= 8 && atomicCAS(reinterpret_cast<int*>(p.trace), 0, 1) == 0) ? 1 : 0;
    __syncthreads();
    tracing = *claim != 0;
    if (tracing && threadIdx.x == 0) p.trace[1] = clock64();
  }

  if (n_blocks > 0) {
    if (warp == 0) {
      if (lane == 0) {
        mbar_expect_tx(q_full, SMEM::Q_BYTES);
#pragma unroll
        for (int c = 0; c < 2; ++c) tma_load_3d(&tmQ, q_full, smem + SMEM::Q_OFF + c * NQ * 128, c * 64, kvh * p.group, t0);
        const int* bt = p.block_table + static_cast<size_t>(seq) * p.max_blocks;
        const int sub = BN / p.block_size;
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
              const int logical = min((pg0 + j) * sub + s, last_blk);
              const int blk = bt[logical];
              const int row_off = s * p.block_size * 128;
#pragma unroll
              for (int c = 0; c < 2; ++c) tma_load_4d(&tmKV, full, tile + c * CHUNK_BYTES + row_off, c * 64, kvh, 0, blk * 2 + is_v);
            }
          }
          if (++st == NS) { st = 0; ph ^= 1; }
        }
      }
    } else if (warp == 1) {
      const bool leader = elect_one();
      constexpr uint32_t idesc_qk = umma_idesc(128, NQ, 1, 1, 0, 0);  // A = K page (K-major), B = Q rows (K-major)
      constexpr uint32_t idesc_pv = umma_idesc(128, NQ, 1, 1, 1, 0);  // A = V page^T (MN-major), B = P^T (K-major)
      const uint32_t tm = __shfl_sync(0xffffffffu, tmem_base, 0);
      const uint32_t tm_o = tm + 32;
      const uint64_t k_desc0 = umma_smem_desc(smem_u32(smem + SMEM::K_OFF), 16, 1024);
      const uint64_t v_desc0 = umma_smem_desc(smem_u32(smem + SMEM::V_OFF), CHUNK_BYTES, 1024);
      const uint64_t q_desc0 = umma_smem_desc(smem_u32(smem + SMEM::Q_OFF), 16, 1024);
      const uint64_t p_desc0 = umma_smem_desc(smem_u32(smem + SMEM::P_OFF), 16, 1024);
      auto issue_qk = [&](int j, int st, int ph) {
        const int b = j & 1;
        mbar_wait_lean(&k_full[st], ph);
        tc_fence_after();
        if (leader) {
          DSB_PF_TRACE(0, j)
          const uint64_t kd = k_desc0 + static_cast<uint32_t>((st * SMEM::TILE) >> 4);
#pragma unroll
          for (int kk = 0; kk < D / 16; ++kk)
            umma_f16<1>(tm + b * NQ, kd + static_cast<uint32_t>(((kk >> 2) * CHUNK_BYTES + (kk & 3) * 32) >> 4),
                        q_desc0 + static_cast<uint32_t>(((kk >> 2) * NQ * 128 + (kk & 3) * 32) >> 4), idesc_qk, kk != 0 ? 1u : 0u);
          umma_commit<1>(&s_full[b]);
          umma_commit<1>(&k_empty[st]);
        }
      };
      auto issue_pv = [&](int j, int st, int ph) {
        const int b = j & 1;
        mbar_wait_lean(&v_full[st], ph);
        mbar_wait_lean(&p_full[b], (j >> 1) & 1);
        tc_fence_after();
        if (leader) {
          DSB_PF_TRACE(1, j)
          const uint64_t vd = v_desc0 + static_cast<uint32_t>((st * SMEM::TILE) >> 4);
          const uint64_t pd = p_desc0 + static_cast<uint32_t>((b * SMEM::P_BYTES) >> 4);
#pragma unroll
          for (int kk = 0; kk < BN / 16; ++kk)
            umma_f16<1>(tm_o, vd + static_cast<uint32_t>((kk * 2048) >> 4),
                        pd + static_cast<uint32_t>(((kk >> 2) * NQ * 128 + (kk & 3) * 32) >> 4), idesc_pv, (j | kk) != 0 ? 1u : 0u);
          umma_commit<1>(o_full);
          umma_commit<1>(&v_empty[st]);
          umma_commit<1>(&p_free[b]);
        }
      };
      mbar_wait_lean(q_full, 0);
      tc_fence_after();
      issue_qk(0, 0, 0);
      int st = 0, ph = 0;
      for (int j = 0; j < n_blocks; ++j) {
        int st1 = st + 1, ph1 = ph;
        if (st1 == NS) { st1 = 0; ph1 ^= 1; }
        if (j + 1 < n_blocks) issue_qk(j + 1, st1, ph1);
        issue_pv(j, st, ph);
        st = st1; ph = ph1;
      }
    } else {
      // softmax: thread == (key of the page = TMEM lane, 4 of the 16 q columns).  16 warps: lane quarter w x column part; the
      // parts past the atom's live columns only keep the barrier counts.  One warp per scheduler owning all 16 columns ran
      // ~600 latency-bound instructions a page.
      const int w = warp & 3, part = (warp - 2) >> 2;
      constexpr int CPT = 4;
      const int c0 = part * CPT;
      const int row = w * 32 + lane;
      const uint32_t lane_off = static_cast<uint32_t>(w * 32) << 16;
      if (c0 >= nq) {
        for (int j = 0; j < n_blocks; ++j) {
          mbar_arrive(&p_full[j & 1]);
          mbar_wait(&p_full[j & 1], (j >> 1) & 1);
        }
      } else {
      float* xmax = reinterpret_cast<float*>(smem + SMEM::X_OFF);  // [parity][warp quarter][16]
      const uint32_t bar_id = 1 + part;  // the 4 warps (128 key-lanes) that share these columns
      float m_run[CPT], l_part[CPT];  // reference maximum per column (uniform over the part), this key-lane's share of the sums
      int limit[CPT], first[CPT];     // keys [first, limit) are visible to column c
      float slope_l2[CPT];            // ALiBi slope of the column's head, exp2 domain (0 without ALiBi)
      const bool has_alibi = p.alibi != nullptr;
      const float sl2 = has_alibi ? 1.f : p.scale_log2;
#pragma unroll
      for (int i = 0; i < CPT; ++i) {
        m_run[i] = -INFINITY;
        l_part[i] = 0.f;
        limit[i] = c0 + i < nq ? pos0 + (c0 + i) / p.group + 1 : 0x3fffffff;
        first[i] = (c0 + i < nq && p.window > 0) ? max(0, limit[i] - p.window) : 0;
        slope_l2[i] = (has_alibi && c0 + i < nq) ? p.alibi[kvh * p.group + (c0 + i) % p.group] * 1.4426950408889634f : 0.f;
      }
      const uint32_t s_addr0 = tmem_base + c0 + lane_off, o_addr = tmem_base + 32 + c0 + lane_off;
      for (int j = 0; j < n_blocks; ++j) {
        const int b = j & 1;
        mbar_wait(&s_full[b], (j >> 1) & 1);
        if (warp == 4 && lane == 0) { DSB_PF_TRACE(2, j) }
        tc_fence_after();
        uint32_t sv[CPT];
        tmem_ld_32x4(s_addr0 + b * NQ, sv);
        tmem_ld_wait();
        const int key = (pg0 + j) * BN + row;
        float s[CPT], wm[CPT];
        // column maxima over the 128 keys: one REDUX per column inside the warp (floats mapped monotonically to unsigned),
        // then the 4 lane-quarter warps meet in shared memory
#pragma unroll
        for (int i = 0; i < CPT; ++i) {
          s[i] = (key >= limit[i] || key < first[i]) ? -INFINITY : __uint_as_float(sv[i]);
          if (has_alibi) s[i] = fmaf(s[i], p.scale_log2, slope_l2[i] * static_cast<float>(key));
          const uint32_t bits = __float_as_uint(s[i]);
          const uint32_t r = __reduce_max_sync(0xffffffffu, bits ^ ((bits >> 31) ? 0xffffffffu : 0x80000000u));
          wm[i] = __uint_as_float(r ^ ((r >> 31) ? 0x80000000u : 0xffffffffu));
        }
        if (lane == 0) *reinterpret_cast<float4*>(xmax + (b * 4 + w) * NQ + c0) = make_float4(wm[0], wm[1], wm[2], wm[3]);
        asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
        float comb[CPT];
        {
          const float4 a0 = *reinterpret_cast<const float4*>(xmax + (b * 4 + 0) * NQ + c0);
          const float4 a1 = *reinterpret_cast<const float4*>(xmax + (b * 4 + 1) * NQ + c0);
          const float4 a2 = *reinterpret_cast<const float4*>(xmax + (b * 4 + 2) * NQ + c0);
          const float4 a3 = *reinterpret_cast<const float4*>(xmax + (b * 4 + 3) * NQ + c0);
          comb[0] = fmaxf(fmaxf(a0.x, a1.x), fmaxf(a2.x, a3.x));
          comb[1] = fmaxf(fmaxf(a0.y, a1.y), fmaxf(a2.y, a3.y));
          comb[2] = fmaxf(fmaxf(a0.z, a1.z), fmaxf(a2.z, a3.z));
          comb[3] = fmaxf(fmaxf(a0.w, a1.w), fmaxf(a2.w, a3.w));
        }
        bool any_rescale = false;
        float alpha[CPT], m_use[CPT];
#pragma unroll
        for (int i = 0; i < CPT; ++i) {
          const float m_cand = fmaxf(m_run[i], comb[i] * sl2);
          const bool fresh = m_run[i] == -INFINITY;
          const float m_new = (fresh || m_cand - m_run[i] > 8.f) ? m_cand : m_run[i];
          const bool rescale = !fresh && m_new != m_run[i];
          alpha[i] = rescale ? ex2_fast(m_run[i] - m_new) : 1.f;
          any_rescale |= rescale;
          m_run[i] = m_new;
          m_use[i] = m_new == -INFINITY ? 0.f : m_new;
        }
        if (any_rescale && j > 0) {  // uniform over the part: every thread of it sees the same maxima
          mbar_wait(o_full, (j - 1) & 1);  // PV(j-1) has retired; PV(j) cannot start before this thread publishes P(j)
          tc_fence_after();
          uint32_t ov[CPT];
          tmem_ld_32x4(o_addr, ov);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < CPT; ++i) {
            ov[i] = __float_as_uint(__uint_as_float(ov[i]) * alpha[i]);
            l_part[i] *= alpha[i];
          }
          tmem_st_32x4(o_addr, ov);
          tmem_st_wait();
          tc_fence_before();
        }
        if (j >= 2) mbar_wait(&p_free[b], ((j >> 1) - 1) & 1);  // PV(j-2) has read this P^T buffer
        // P^T [16 rows][128 keys] bf16, K-major, 128-byte swizzle rows of 64 keys: this thread owns column `row` of its q rows
        uint8_t* pbuf = smem + SMEM::P_OFF + b * SMEM::P_BYTES + (row >> 6) * (NQ * 128);
        const int k64 = row & 63;
#pragma unroll
        for (int i = 0; i < CPT; ++i) {
          const int c = c0 + i;
          if (c < nq) {
            const float e = ex2_fast(fmaf(s[i], sl2, -m_use[i]));
            l_part[i] += e;
            *reinterpret_cast<__nv_bfloat16*>(pbuf + c * 128 + ((((k64 >> 3) ^ (c & 7)) << 4) | ((k64 & 7) << 1))) = __float2bfloat16_rn(e);
          }
        }
        fence_proxy_async_smem();
        mbar_arrive(&p_full[b]);
        if (warp == 4 && lane == 0) { DSB_PF_TRACE(3, j) }
      }
      // epilogue: O^T [d = row][this part's columns]; column sums over the 128 key-lanes
      mbar_wait(o_full, (n_blocks - 1) & 1);
      tc_fence_after();
      uint32_t ov[CPT];
      tmem_ld_32x4(o_addr, ov);
      tmem_ld_wait();
      float* xl = reinterpret_cast<float*>(smem + SMEM::X_OFF) + 2 * 4 * NQ;  // [warp quarter][16], behind the maxima
      float tsum[CPT];
#pragma unroll
      for (int i = 0; i < CPT; ++i) tsum[i] = warp_sum(l_part[i]);
      if (lane == 0) *reinterpret_cast<float4*>(xl + w * NQ + c0) = make_float4(tsum[0], tsum[1], tsum[2], tsum[3]);
      asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
#pragma unroll
      for (int i = 0; i < CPT; ++i) {
        const int c = c0 + i;
        const float l = (xl[c] + xl[NQ + c]) + (xl[2 * NQ + c] + xl[3 * NQ + c]);
        if (c < nq) {
          const int t = t0 + c / p.group, head = kvh * p.group + c % p.group;
          if (!partial) {
            p.out[(static_cast<size_t>(t) * p.hq + head) * D + row] = __float2bfloat16_rn(__uint_as_float(ov[i]) / l);
          } else {
            const size_t slot = (static_cast<size_t>(t) * p.hq + head) * p.n_splits + split;
            p.part_o[slot * D + row] = __uint_as_float(ov[i]);
            if (row == 0) {
              p.part_ml[slot * 2] = m_run[i];
              p.part_ml[slot * 2 + 1] = l;
            }
          }
        }
      }
      }  // live part
    }
  } else if (partial && warp == 2) {
    const int c = lane;  // this split has no pages: an empty partial per live column
    if (c < nq) {
      const size_t slot = (static_cast<size_t>(t0 + c / p.group) * p.hq + kvh * p.group + c % p.group) * p.n_splits + split;
      p.part_ml[slot * 2] = -INFINITY;
      p.part_ml[slot * 2 + 1] = 0.f;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<1>(tmem_base, 64);
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
                  void* scratch, const float* alibi, int window, cudaStream_t stream) {
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
  p.n_splits = n_splits; p.scale_log2 = scale * 1.4426950408889634f; p.alibi = alibi; p.window = window;
  static const bool want_trace = [] { const char* e = getenv("DSB200_PF_TRACE"); return e && atoi(e) != 0; }();
  static long long* trace_buf = nullptr;
  p.trace = nullptr;
  if (want_trace) {
    if (!trace_buf) DSB_CHECK_CUDA(cudaMalloc(&trace_buf, (8 + 64 * 8) * sizeof(long long)));
    DSB_CHECK_CUDA(cudaMemsetAsync(trace_buf, 0, (8 + 64 * 8) * sizeof(long long), stream));
    p.trace = trace_buf;
  }
  dim3 grid(max_atoms, hkv, n_splits);
  kern<<<grid, NUM_THREADS, SMEM::TOTAL, stream>>>(tmQ, tmKV, p);
  DSB_CHECK_CUDA(cudaGetLastError());
  if (want_trace) {  // diagnostics: synchronous dump of the claimed CTA's timeline (cycles relative to its start)
    long long h[8 + 64 * 8];
    DSB_CHECK_CUDA(cudaStreamSynchronize(stream));
    DSB_CHECK_CUDA(cudaMemcpy(h, trace_buf, sizeof(h), cudaMemcpyDeviceToHost));
    printf("paged_flash trace (T=%d splits=%d): page  K_landed  P_seen_by_mma  S_seen  P_published  O_read\n", T, n_splits);
    for (int j = 0; j < 64 && h[8 + j * 8 + 2]; ++j)
      printf("  %2d %8lld %8lld %8lld %8lld %8lld\n", j, h[8 + j * 8] - h[1], h[8 + j * 8 + 1] - h[1], h[8 + j * 8 + 2] - h[1],
             h[8 + j * 8 + 3] - h[1], h[8 + j * 8 + 4] - h[1]);
    fflush(stdout);
  }
  if (n_splits > 1) {
    const int rows = T * hq;
    combine_splits_kernel<D><<<(rows * 32 + 255) / 256, 256, 0, stream>>>(part_o, part_ml, static_cast<__nv_bfloat16*>(out), rows, n_splits);
    DSB_CHECK_CUDA(cudaGetLastError());
  }
  return DSB_OK;
}

// decode variant: atoms of at most 16 / group tokens, head_dim 128
static int launch_decode(const void* qkv, const void* kv_cache, void* out, const int* tok_seq, const int* tok_pos, const int* block_table,
                         int T, int hq, int hkv, int block_size, int max_blocks, int num_blocks, float scale, int max_atoms,
                         int n_splits, void* scratch, const float* alibi, int window, cudaStream_t stream) {
  constexpr int D = 128;
  const int group = hq / hkv, nt_tile = NQ / group;
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
  int* atoms = static_cast<int*>(scratch);  // (layout shared with the general kernel; the atom list itself is not used here)
  const size_t head_ints = (static_cast<size_t>(max_atoms) + 1 + 3) / 4 * 4;
  float* part_o = reinterpret_cast<float*>(atoms + head_ints);
  float* part_ml = part_o + static_cast<size_t>(T) * hq * n_splits * D;
  static bool attr_set = false;
  if (!attr_set) {
    DSB_CHECK_CUDA(cudaFuncSetAttribute(paged_decode_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, DecSmem::TOTAL));
    attr_set = true;
  }
  Params p;
  p.tok_seq = tok_seq; p.tok_pos = tok_pos; p.block_table = block_table; p.atoms = nullptr; p.n_atoms = nullptr;
  p.out = static_cast<__nv_bfloat16*>(out); p.part_o = part_o; p.part_ml = part_ml;
  p.T = T; p.hq = hq; p.hkv = hkv; p.group = group; p.nt_tile = nt_tile; p.block_size = block_size; p.max_blocks = max_blocks;
  p.n_splits = n_splits; p.scale_log2 = scale * 1.4426950408889634f; p.trace = nullptr; p.alibi = alibi; p.window = window;
  static const bool want_trace = [] { const char* e = getenv("DSB200_PF_TRACE"); return e && atoi(e) != 0; }();
  static long long* trace_buf = nullptr;
  if (want_trace) {
    if (!trace_buf) DSB_CHECK_CUDA(cudaMalloc(&trace_buf, (8 + 64 * 8) * sizeof(long long)));
    DSB_CHECK_CUDA(cudaMemsetAsync(trace_buf, 0, (8 + 64 * 8) * sizeof(long long), stream));
    p.trace = trace_buf;
  }
  dim3 grid(T, hkv, n_splits);
  paged_decode_kernel<<<grid, DEC_THREADS, DecSmem::TOTAL, stream>>>(tmQ, tmKV, p);
  DSB_CHECK_CUDA(cudaGetLastError());
  if (want_trace) {
    long long h[8 + 64 * 8];
    DSB_CHECK_CUDA(cudaStreamSynchronize(stream));
    DSB_CHECK_CUDA(cudaMemcpy(h, trace_buf, sizeof(h), cudaMemcpyDeviceToHost));
    printf("paged_decode trace (T=%d splits=%d): page  QK_issue  P_seen_by_mma  S_seen  max_exchanged  alpha_done  p_free_seen  stored  P_published\n", T, n_splits);
    for (int j = 0; j < 64 && h[8 + j * 8 + 2]; ++j)
      printf("  %2d %8lld %8lld %8lld %8lld %8lld %8lld %8lld %8lld\n", j, h[8 + j * 8] - h[1], h[8 + j * 8 + 1] - h[1], h[8 + j * 8 + 2] - h[1],
             h[8 + j * 8 + 4] - h[1], h[8 + j * 8 + 5] - h[1], h[8 + j * 8 + 6] - h[1], h[8 + j * 8 + 7] - h[1], h[8 + j * 8 + 3] - h[1]);
    fflush(stdout);
  }
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
// alibi_slopes: optional [hq] fp32 (device): score += slope[head] * key position; NULL = none.
// window > 0: local attention, a token sees its last `window` keys (itself included); 0 = the whole history.
DSB_EXPORT int dsb200_paged_flash_attention(const void* qkv, const void* kv_cache, void* out, const int* tok_seq, const int* tok_pos,
                                            const int* block_table, int n_tokens, int hq, int hkv, int head_dim, int block_size,
                                            int max_blocks_per_seq, int num_blocks, float scale, int max_atoms, int n_splits,
                                            void* scratch, const float* alibi_slopes, int window, cudaStream_t stream) {
  using namespace dsb;
  if (n_tokens == 0) return DSB_OK;
  DSB_REQUIRE(head_dim == 64 || head_dim == 128, "paged_flash: head_dim %d unsupported (64 or 128)", head_dim);
  DSB_REQUIRE(hq % hkv == 0 && 128 % (hq / hkv) == 0, "paged_flash: q/kv head ratio %d must divide 128", hq / hkv);
  DSB_REQUIRE(block_size >= 8 && block_size <= 128 && 128 % block_size == 0, "paged_flash: KV block size %d must divide 128", block_size);
  DSB_REQUIRE(max_atoms >= 1 && n_splits >= 1 && n_splits <= 64 && scratch != nullptr, "paged_flash: bad max_atoms / n_splits / scratch");
  if (head_dim == 128)
    return paged::launch<128>(qkv, kv_cache, out, tok_seq, tok_pos, block_table, n_tokens, hq, hkv, block_size, max_blocks_per_seq,
                              num_blocks, scale, max_atoms, n_splits, scratch, alibi_slopes, window, stream);
  return paged::launch<64>(qkv, kv_cache, out, tok_seq, tok_pos, block_table, n_tokens, hq, hkv, block_size, max_blocks_per_seq,
                           num_blocks, scale, max_atoms, n_splits, scratch, alibi_slopes, window, stream);
}

// Same contract as dsb200_paged_flash_attention for batches of small atoms (decode steps): atoms hold at most 16 / group tokens,
// keys take the M dimension of the MMAs (see paged_decode_kernel).  head_dim 128, q:kv head ratio dividing 16.
// max_atoms >= ceil(n_tokens / (16/group)) + number of sequences (<= n_tokens).
DSB_EXPORT int dsb200_paged_flash_decode(const void* qkv, const void* kv_cache, void* out, const int* tok_seq, const int* tok_pos,
                                         const int* block_table, int n_tokens, int hq, int hkv, int head_dim, int block_size,
                                         int max_blocks_per_seq, int num_blocks, float scale, int max_atoms, int n_splits,
                                         void* scratch, const float* alibi_slopes, int window, cudaStream_t stream) {
  using namespace dsb;
  if (n_tokens == 0) return DSB_OK;
  DSB_REQUIRE(head_dim == 128, "paged_flash_decode: head_dim %d unsupported (128)", head_dim);
  DSB_REQUIRE(hq % hkv == 0 && 16 % (hq / hkv) == 0, "paged_flash_decode: q/kv head ratio %d must divide 16", hq / hkv);
  DSB_REQUIRE(block_size >= 8 && block_size <= 128 && 128 % block_size == 0, "paged_flash_decode: KV block size %d must divide 128", block_size);
  DSB_REQUIRE(max_atoms >= 1 && n_splits >= 1 && n_splits <= 64 && scratch != nullptr, "paged_flash_decode: bad max_atoms / n_splits / scratch");
  return paged::launch_decode(qkv, kv_cache, out, tok_seq, tok_pos, block_table, n_tokens, hq, hkv, block_size, max_blocks_per_seq,
                              num_blocks, scale, max_atoms, n_splits, scratch, alibi_slopes, window, stream);
}

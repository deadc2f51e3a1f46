// Flash attention forward, second generation: two 128-row Q tiles per CTA in ping-pong, P and O resident in TMEM.
//
// The first-generation kernel (flash_fwd.cu) spends its time in the softmax warps: S is read from TMEM twice, P goes
// through shared memory, and O is folded in registers from a per-block PV accumulator (another full TMEM read per block).
// Here, per 128-key block and per Q tile:
//   QK^T  : SS MMA  Q_t (smem) x K (smem)            -> S_t  (TMEM, 128 fp32 columns)
//   softmax (one warpgroup per tile, thread == row)  : ONE tcgen05.ld pass over S_t into registers, exp2, P packed to bf16
//           and written back with tcgen05.st over the first 64 columns of S_t (P aliases S)
//   PV    : TS MMA  P_t (TMEM) x V (smem, MN-major)  -> O_t  (TMEM, accumulated across blocks by the tensor core)
// O is only touched by threads when the running row max grows by more than 2^8 (lazy rescale: until then P is computed
// against the stale max, which is exact after the final 1/l normalisation), and once at the end.
// The MMA thread alternates the two tiles (PV0, QK0', PV1, QK1'), so one tile's softmax runs under the other tile's MMAs.
// K and V travel through one ring of five 128-row slots (K0 V0 K1 V1 ...), each with its own full/empty barrier.
//   warp 0 TMA producer | warp 1 MMA issuer + TMEM owner | warps 2-3 idle (their registers go to the softmax warps via
//   setmaxnreg) | warps 4-7 softmax tile 0 | warps 8-11 softmax tile 1
// TMEM: S0 [0,128)  S1 [128,256)  O0 [256,256+D)  O1 [384,384+D).
#include "../common/host.h"
#include "../common/ptx.cuh"

namespace dsb {
namespace attn2 {

constexpr int BM = 128;
constexpr int BN = 128;
constexpr int NUM_THREADS = 384;  // warpgroup 0: TMA, MMA, 2 idle warps (registers donated) | warpgroups 1, 2: softmax of tile 0, 1
constexpr int WG_THREADS = 128;
constexpr int CHUNK_BYTES = 128 * 128;
constexpr int KV_SLOTS = 5;
constexpr float RESCALE_THRESHOLD = 8.f;  // log2 units

template <int N>
__device__ __forceinline__ void reg_dec() {
  asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(N));
}
template <int N>
__device__ __forceinline__ void reg_inc() {
  asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(N));
}


constexpr int POLY_OF_8 = 3;  // pairs out of every 8 that use exp2_poly2

template <int D>
struct Smem {
  static constexpr int NCH = D / 64;
  static constexpr int TILE = NCH * CHUNK_BYTES;
  static constexpr int Q_OFF = 0;                       // 2 tiles
  static constexpr int KV_OFF = 2 * TILE;               // KV_SLOTS tiles
  static constexpr int BAR_OFF = KV_OFF + KV_SLOTS * TILE;
  static constexpr int TOTAL = BAR_OFF + 256 + 1024;
};

struct Params {
  int B, S, hq, hkv;
  int causal;
  float scale_log2;
  float* lse;
  long long* trace;  // debug: clock64 stamps of CTA (0,0,0), [block j][tile][8 events]; nullptr in production
};
#define DSB_TRACE(ev, j, t)                                                                            \
  if (p.trace && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0) p.trace[((j) * 2 + (t)) * 8 + (ev)] = clock64();

template <int D>
__global__ void __launch_bounds__(NUM_THREADS, 1)
flash_fwd2_kernel(const __grid_constant__ CUtensorMap tmQKV, const __grid_constant__ CUtensorMap tmO, const Params p) {
  using SM = Smem<D>;
  constexpr int NCH = SM::NCH;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + SM::BAR_OFF);
  uint64_t* q_full = bars + 0;                  // [2]
  uint64_t* kv_full = bars + 2;                 // [KV_SLOTS]
  uint64_t* kv_empty = kv_full + KV_SLOTS;      // [KV_SLOTS]
  uint64_t* s_full = kv_empty + KV_SLOTS;       // [2] MMA -> softmax : S_t(j) complete
  uint64_t* p_full = s_full + 2;                // [2] softmax -> MMA : P_t(j) stored, O_t rescaled if needed
  uint64_t* o_done = p_full + 2;                // [2] MMA -> softmax : PV_t(j) complete
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(o_done + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_pairs = p.S / (2 * BM);
  const int qp = n_pairs - 1 - static_cast<int>(blockIdx.x);  // heavy (late) tiles first
  const int head = blockIdx.y, batch = blockIdx.z;
  const int kv_head = head / (p.hq / p.hkv);
  const int nb0 = p.causal ? (2 * qp + 1) : (p.S / BN);  // key blocks of tile 0
  const int nb1 = p.causal ? (2 * qp + 2) : (p.S / BN);  // key blocks of tile 1 (>= nb0)
  const int row0 = batch * p.S + qp * 2 * BM;
  const int q_col = head * D;
  const int k_col = (p.hq + kv_head) * D;
  const int v_col = (p.hq + p.hkv + kv_head) * D;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmQKV);
    tma_prefetch_desc(&tmO);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&q_full[i], 1);
      mbar_init(&s_full[i], 1);
      mbar_init(&p_full[i], WG_THREADS);
      mbar_init(&o_done[i], 1);
    }
    for (int i = 0; i < KV_SLOTS; ++i) {
      mbar_init(&kv_full[i], 1);
      mbar_init(&kv_empty[i], 1);
    }
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc<1>(tmem_ptr, 512);
    tmem_relinquish<1>();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  if (warp == 0) {
    // ------------------------------------------------------------- TMA producer
    reg_dec<56>();
    if (lane == 0) {
      for (int t = 0; t < 2; ++t) {
        mbar_expect_tx(&q_full[t], SM::TILE);
#pragma unroll
        for (int c = 0; c < NCH; ++c)
          tma_load_2d(&tmQKV, &q_full[t], smem + SM::Q_OFF + t * SM::TILE + c * CHUNK_BYTES, q_col + c * 64, row0 + t * BM);
      }
      int slot = 0;
      uint32_t phase = 0;
      for (int item = 0; item < 2 * nb1; ++item) {  // item 2j = K(j), 2j+1 = V(j)
        const int j = item >> 1;
        const int col = (item & 1) ? v_col : k_col;
        mbar_wait(&kv_empty[slot], phase ^ 1);
        mbar_expect_tx(&kv_full[slot], SM::TILE);
        uint8_t* dst = smem + SM::KV_OFF + slot * SM::TILE;
#pragma unroll
        for (int c = 0; c < NCH; ++c) tma_load_2d(&tmQKV, &kv_full[slot], dst + c * CHUNK_BYTES, col + c * 64, batch * p.S + j * BN);
        if (++slot == KV_SLOTS) {
          slot = 0;
          phase ^= 1;
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------- MMA issuer
    // The whole warp runs this loop convergently and one elected lane issues: every operand below derives from kernel
    // parameters, loop counters and one shuffled TMEM base, so descriptors stay cheap to form (a low-word add per MMA).
    reg_dec<56>();
    {
      constexpr uint32_t idesc_qk = umma_idesc(BM, BN, 1, 1, 0, 0);
      constexpr uint32_t idesc_pv = umma_idesc(BM, D, 1, 1, 0, 1);  // A = P (TMEM), B = V (MN-major)
      const uint32_t tm = __shfl_sync(0xffffffffu, tmem_base, 0);
      const bool leader = elect_one();
      // descriptor = {hi: SBO | version | swizzle (constant per operand kind), lo: (addr >> 4) | LBO << 16}
      constexpr uint32_t HI = (1024u >> 4) | (1u << 14) | (2u << 29);
      constexpr uint32_t LO_KMAJ = (16u >> 4) << 16;              // Q, K : K-major, LBO 16
      constexpr uint32_t LO_V = (uint32_t(CHUNK_BYTES) >> 4) << 16;  // V : MN-major, LBO = chunk stride
      const uint32_t smem_base = smem_u32(smem);
      auto desc = [](uint32_t lo) { return (static_cast<uint64_t>(HI) << 32) | lo; };
      int slot = 0;  // ring position of the next item consumed (K0 V0 K1 V1 ...)
      uint32_t phase = 0;
      auto advance = [&]() {
        if (++slot == KV_SLOTS) {
          slot = 0;
          phase ^= 1;
        }
      };
      auto issue_qk = [&](int t, uint32_t k_lo) {
        if (leader) {
          const uint32_t q_lo = LO_KMAJ | ((smem_base + SM::Q_OFF + t * SM::TILE) >> 4);
#pragma unroll
          for (int kk = 0; kk < D / 16; ++kk) {
            const uint32_t off = ((kk >> 2) * CHUNK_BYTES + (kk & 3) * 32) >> 4;
            umma_f16<1>(tm + t * 128, desc(q_lo + off), desc(k_lo + off), idesc_qk, kk != 0 ? 1u : 0u);
          }
          umma_commit<1>(&s_full[t]);
        }
      };
      auto issue_pv = [&](int t, int j, uint32_t v_lo) {
        mbar_wait_lean(&p_full[t], j & 1);
        tc_fence_after();
        if (leader) {
          DSB_TRACE(4, j, t)
#pragma unroll
          for (int kk = 0; kk < BN / 16; ++kk)
            umma_f16_ts(tm + 256 + t * 128, tm + t * 128 + kk * 8, desc(v_lo + ((kk * 2048) >> 4)), idesc_pv, (j | kk) != 0 ? 1u : 0u);
          umma_commit<1>(&o_done[t]);
          DSB_TRACE(5, j, t)
        }
      };
      // K(0)
      mbar_wait_lean(&kv_full[slot], phase);
      tc_fence_after();
      {
        const uint32_t k_lo = LO_KMAJ | ((smem_base + SM::KV_OFF + slot * SM::TILE) >> 4);
        mbar_wait_lean(&q_full[0], 0);
        issue_qk(0, k_lo);
        mbar_wait_lean(&q_full[1], 0);
        issue_qk(1, k_lo);
        if (leader) umma_commit<1>(&kv_empty[slot]);
        advance();
      }
      for (int j = 0; j < nb1; ++j) {
        // V(j)
        const int v_slot = slot;
        mbar_wait_lean(&kv_full[v_slot], phase);
        const uint32_t v_lo = LO_V | ((smem_base + SM::KV_OFF + v_slot * SM::TILE) >> 4);
        advance();
        // K(j+1)
        const bool more = j + 1 < nb1;
        const int k_slot = slot;
        uint32_t k_lo = 0;
        if (more) {
          mbar_wait_lean(&kv_full[k_slot], phase);
          k_lo = LO_KMAJ | ((smem_base + SM::KV_OFF + k_slot * SM::TILE) >> 4);
          advance();
        }
        tc_fence_after();
        if (j < nb0) issue_pv(0, j, v_lo);
        if (j + 1 < nb0) issue_qk(0, k_lo);
        issue_pv(1, j, v_lo);
        if (leader) umma_commit<1>(&kv_empty[v_slot]);
        if (more) {
          issue_qk(1, k_lo);
          if (leader) umma_commit<1>(&kv_empty[k_slot]);
        }
      }
    }
  } else if (warp < 4) {
    reg_dec<56>();  // idle warps of warpgroup 0: setmaxnreg is warpgroup-collective
  } else {
    // ------------------------------------------------------------- softmax warpgroups (thread == row)
    reg_inc<224>();
    const int t = (warp - 4) >> 2;  // tile
    const int q4 = warp & 3;        // TMEM lane quarter this warp may access
    const int row = q4 * 32 + lane;
    const uint32_t lane_off = static_cast<uint32_t>(q4 * 32) << 16;
    const uint32_t TM_S = tmem_base + t * 128 + lane_off;
    const uint32_t TM_O = tmem_base + 256 + t * 128 + lane_off;
    const int nb = t == 0 ? nb0 : nb1;
    const int diag_blk = p.causal ? (2 * qp + t) : -1;
    float m_run = -INFINITY, l_run = 0.f;

    for (int j = 0; j < nb; ++j) {
      mbar_wait(&s_full[t], j & 1);
      tc_fence_after();
      if (row == 0) { DSB_TRACE(0, j, t) }
      uint32_t s0[32], s1[32], s2[32], s3[32];
      tmem_ld_32x32(TM_S + 0, s0);
      tmem_ld_32x32(TM_S + 32, s1);
      tmem_ld_32x32(TM_S + 64, s2);
      tmem_ld_32x32(TM_S + 96, s3);
      tmem_ld_wait();
      if (row == 0) { DSB_TRACE(1, j, t) }
      if (j == diag_blk) {
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          if (i > row) s0[i] = 0xff800000u;  // -inf
          if (i + 32 > row) s1[i] = 0xff800000u;
          if (i + 64 > row) s2[i] = 0xff800000u;
          if (i + 96 > row) s3[i] = 0xff800000u;
        }
      }
      float mx0 = -INFINITY, mx1 = -INFINITY, mx2 = -INFINITY, mx3 = -INFINITY;
#pragma unroll
      for (int i = 0; i < 32; ++i) {
        mx0 = fmaxf(mx0, __uint_as_float(s0[i]));
        mx1 = fmaxf(mx1, __uint_as_float(s1[i]));
        mx2 = fmaxf(mx2, __uint_as_float(s2[i]));
        mx3 = fmaxf(mx3, __uint_as_float(s3[i]));
      }
      const float m_blk = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3)) * p.scale_log2;
      if (row == 0) { DSB_TRACE(2, j, t) }
      if (j == 0) {
        m_run = m_blk;
      } else {
        // lazy rescale: only when some row of this warp moved its max by more than the threshold (warp-uniform branch,
        // the TMEM accesses below are warp-collective)
        const bool grow = m_blk > m_run + RESCALE_THRESHOLD;
        if (__any_sync(0xffffffffu, grow)) {
          const float m_new = grow ? m_blk : m_run;
          const float alpha = ex2_approx(m_run - m_new);
          mbar_wait(&o_done[t], (j - 1) & 1);  // PV_t(j-1) has landed in O_t; nothing writes O_t until P_t(j) is published
          tc_fence_after();
#pragma unroll
          for (int c = 0; c < D / 32; ++c) {
            uint32_t o[32];
            tmem_ld_32x32(TM_O + c * 32, o);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
            tmem_st_32x32(TM_O + c * 32, o);
          }
          l_run *= alpha;
          m_run = m_new;
        }
      }
      // P = exp2(s * scale - m_run), packed two bf16 per 32-bit TMEM column over the first 64 columns of S_t.
      // Packed fp32x2 math (FFMA2 / FADD2) halves the issue slots; 3 of every 8 pairs take the polynomial exp2 on the FMA
      // pipe so the MUFU unit (16 ex2 / clk / SM, shared by both tiles' warps) is not the limiter.
      const uint64_t sc2 = pack2(p.scale_log2, p.scale_log2), negm2 = pack2(-m_run, -m_run);
      uint64_t l2a = 0ull, l2b = 0ull;  // two (0.f, 0.f) accumulators
#define DSB_SOFTMAX_HALF(SRC_A, SRC_B, COL)                                                                  \
  {                                                                                                          \
    uint32_t pk[32];                                                                                         \
    _Pragma("unroll") for (int i = 0; i < 32; ++i) {                                                         \
      const uint32_t* src = i < 16 ? SRC_A : SRC_B;                                                          \
      const int k = (i & 15) * 2;                                                                            \
      const uint64_t x2 = fma2(pack2(__uint_as_float(src[k]), __uint_as_float(src[k + 1])), sc2, negm2);     \
      const uint64_t e2 = ((i & 7) < POLY_OF_8) ? exp2_poly2(x2) : exp2_mufu2(x2);                            \
      if (i & 1) l2b = add2(l2b, e2); else l2a = add2(l2a, e2);                                              \
      float e0, e1;                                                                                          \
      unpack2(e2, e0, e1);                                                                                   \
      pk[i] = pack_bf16x2(e0, e1);                                                                           \
    }                                                                                                        \
    tmem_st_32x32(TM_S + COL, pk);                                                                           \
  }
      DSB_SOFTMAX_HALF(s0, s1, 0)
      DSB_SOFTMAX_HALF(s2, s3, 32)
#undef DSB_SOFTMAX_HALF
      float la, lb;
      unpack2(add2(l2a, l2b), la, lb);
      const float l_blk = la + lb;
      l_run += l_blk;
      tmem_st_wait();
      tc_fence_before();
      mbar_arrive(&p_full[t]);
      if (row == 0) { DSB_TRACE(3, j, t) }
    }

    // ---- epilogue: O_t / l -> bf16 -> swizzled staging in the Q_t tile (all QK MMAs of the tile have retired) -> TMA store
    mbar_wait(&o_done[t], (nb - 1) & 1);
    tc_fence_after();
    const float inv_l = 1.f / l_run;
    uint8_t* obuf = smem + SM::Q_OFF + t * SM::TILE;
    const int sw = row & 7;
#pragma unroll
    for (int c = 0; c < D / 32; ++c) {
      uint32_t o[32];
      tmem_ld_32x32(TM_O + c * 32, o);
      tmem_ld_wait();
      uint8_t* rowp = obuf + (c >> 1) * CHUNK_BYTES + row * 128;
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        uint4 val;
        val.x = pack_bf16x2(__uint_as_float(o[8 * u + 0]) * inv_l, __uint_as_float(o[8 * u + 1]) * inv_l);
        val.y = pack_bf16x2(__uint_as_float(o[8 * u + 2]) * inv_l, __uint_as_float(o[8 * u + 3]) * inv_l);
        val.z = pack_bf16x2(__uint_as_float(o[8 * u + 4]) * inv_l, __uint_as_float(o[8 * u + 5]) * inv_l);
        val.w = pack_bf16x2(__uint_as_float(o[8 * u + 6]) * inv_l, __uint_as_float(o[8 * u + 7]) * inv_l);
        const int unit = (c & 1) * 4 + u;
        *reinterpret_cast<uint4*>(rowp + ((unit ^ sw) << 4)) = val;
      }
    }
    if (p.lse)
      p.lse[(static_cast<size_t>(batch) * p.hq + head) * p.S + qp * 2 * BM + t * BM + row] = (m_run + log2f(l_run)) * 0.6931471805599453f;
    fence_proxy_async_smem();
    named_bar_sync(1 + t, WG_THREADS);
    if ((threadIdx.x & (WG_THREADS - 1)) == 0) {
#pragma unroll
      for (int c = 0; c < NCH; ++c) tma_store_2d(&tmO, obuf + c * CHUNK_BYTES, head * D + c * 64, row0 + t * BM);
      tma_commit_group();
      tma_wait_group<0>();
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<1>(tmem_base, 512);
  }
}

static long long* g_trace = nullptr;

template <int D>
int launch(const void* qkv, void* out, float* lse, int B, int S, int hq, int hkv, int causal, float scale, cudaStream_t stream) {
  using SM = Smem<D>;
  static_assert(SM::TOTAL <= 232448, "flash_fwd2 shared memory");
  CUtensorMap tmQKV, tmO;
  const int row_elems = (hq + 2 * hkv) * D;
  int rc = make_tensor_map_2d(&tmQKV, TMA_BF16, qkv, static_cast<uint64_t>(B) * S, row_elems, row_elems, 128, 64, TMA_SW_128);
  if (rc) return rc;
  rc = make_tensor_map_2d(&tmO, TMA_BF16, out, static_cast<uint64_t>(B) * S, hq * D, hq * D, 128, 64, TMA_SW_128);
  if (rc) return rc;
  auto kern = flash_fwd2_kernel<D>;
  static bool attr_set = false;
  if (!attr_set) {
    DSB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SM::TOTAL));
    attr_set = true;
  }
  Params p;
  p.B = B; p.S = S; p.hq = hq; p.hkv = hkv; p.causal = causal;
  p.scale_log2 = scale * 1.4426950408889634f;
  p.lse = lse;
  p.trace = g_trace;
  dim3 grid(S / (2 * BM), hq, B);
  kern<<<grid, NUM_THREADS, SM::TOTAL, stream>>>(tmQKV, tmO, p);
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

}  // namespace attn2

void flash_fwd2_set_trace(long long* buf) { attn2::g_trace = buf; }

int flash_fwd2_launch(const void* qkv, void* out, float* lse, int B, int S, int hq, int hkv, int head_dim, int causal, float scale,
                      cudaStream_t stream) {
  if (head_dim == 128) return attn2::launch<128>(qkv, out, lse, B, S, hq, hkv, causal, scale, stream);
  return attn2::launch<64>(qkv, out, lse, B, S, hq, hkv, causal, scale, stream);
}

}  // namespace dsb

// Flash attention backward for sm_100a on tcgen05 tensor cores (causal / full, MHA / GQA, head_dim 64/128).
//
// One CTA per (batch, kv-head, 128-key block); it keeps K and V resident in shared memory and dK, dV as fp32
// accumulators in TMEM, and walks over every (q-head of the group, q block) that attends to its keys.  Per step, five
// tcgen05 GEMMs run off the SAME swizzled smem tiles, only the UMMA descriptors (K-major vs MN-major) differ:
//     S  = Q K^T            A = Q  (K-major)   B = K  (K-major)      -> TMEM [0,128)
//     dP = dO V^T           A = dO (K-major)   B = V  (K-major)      -> TMEM [128,256)
//     P  = exp2(S*c - lse), dS = P * (dP - delta) * scale            (softmax warps, thread == q row, bf16 to smem)
//     dV += P^T dO          A = P  (MN-major)  B = dO (MN-major)     -> TMEM [256,256+D)
//     dK += dS^T Q          A = dS (MN-major)  B = Q  (MN-major)     -> TMEM [384,384+D)
//     dQ  = dS K            A = dS (K-major)   B = K  (MN-major)     -> TMEM [0,128) -> fp32 smem -> TMA reduce-add
// dQ tiles from different key blocks are summed in an fp32 scratch buffer by TMA reduce-add and converted to bf16 by
// a small follow-up kernel; delta = rowsum(dO * O) comes from a small pre-kernel.
// Reference counterpart: csrc/transformer/softmax_kernels.cu (softmax_backward_kernel*) + cuBLAS strided-batched GEMMs
// in ds_transformer_cuda.cpp (attention backward of the training kernel).
#include "../common/host.h"
#include "../common/ptx.cuh"

namespace dsb {
namespace attn {

// One softmax thread per q row (two threads per row were measured: no gain, see the dQ drain notes in the kernel).
constexpr int BW_THREADS = 192;
constexpr int BW_SM_THREADS = 128;
constexpr int BW_CHUNK = 128 * 128;  // [128 rows x 64 bf16] swizzled chunk (also [128 rows x 32 fp32])

__device__ __forceinline__ void bwd_wait(uint64_t* bar, uint32_t parity) { mbar_wait_lean(bar, parity); }
__device__ __forceinline__ float bwd_ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

__device__ __forceinline__ uint64_t bw_pack2(float lo, float hi) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void bw_unpack2(uint64_t v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ uint64_t bw_fma2(uint64_t a, uint64_t b, uint64_t c) {
  uint64_t d;
  asm("fma.rn.ftz.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__device__ __forceinline__ uint64_t bw_mul2(uint64_t a, uint64_t b) {
  uint64_t d;
  asm("mul.rn.ftz.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}

template <int D>
struct BwdSmem {
  static constexpr int NCH = D / 64;
  static constexpr int TILE = NCH * BW_CHUNK;
  static constexpr int K_OFF = 0;
  static constexpr int V_OFF = K_OFF + TILE;
  static constexpr int Q_OFF = V_OFF + TILE;             // 2 stages: Q of the next step is prefetched
  static constexpr int DO_OFF = Q_OFF + 2 * TILE;
  static constexpr int P_OFF = DO_OFF + TILE;           // 128 x 128 bf16 = 2 chunks
  static constexpr int DS_OFF = P_OFF + 2 * BW_CHUNK;   // 128 x 128 bf16 = 2 chunks
  static constexpr int BAR_OFF = DS_OFF + 2 * BW_CHUNK;
  static constexpr int TOTAL = BAR_OFF + 256 + 1024;
};

struct BwdParams {
  int B, S, hq, hkv;
  int causal;
  float scale, scale_log2;
  const float* lse;    // [B, hq, S] natural log
  const float* delta;  // [B, hq, S]
  int debug_skip_dq;   // diagnostics only (DSB200_BWD_SKIP_DQ=1): skip the dQ reduce-add to measure its cost
  long long* trace;    // debug: clock64 stamps of CTA (0,0,0), [step][16 events]; nullptr in production
};
#define BWD_TRACE(ev, t) \
  if (p.trace && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && (t) < 256) p.trace[(t) * 16 + (ev)] = clock64();
static long long* g_bwd_trace = nullptr;

// delta[b,h,s] = sum_d dO[b,s,h,d] * O[b,s,h,d]   (one warp per (token, head))
__global__ void attn_delta_kernel(const __nv_bfloat16* __restrict__ dout, const __nv_bfloat16* __restrict__ out,
                                  float* __restrict__ delta, int B, int S, int hq, int D) {
  const long long gw = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  const long long total = static_cast<long long>(B) * S * hq;
  if (gw >= total) return;
  const long long tok = gw / hq;
  const int h = static_cast<int>(gw - tok * hq);
  const __nv_bfloat16* a = dout + (tok * hq + h) * D;
  const __nv_bfloat16* b = out + (tok * hq + h) * D;
  float acc = 0.f;
  for (int i = lane * 2; i < D; i += 64) {
    float2 x = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(a + i));
    float2 y = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(b + i));
    acc += x.x * y.x + x.y * y.y;
  }
  acc = warp_sum(acc);
  if (lane == 0) {
    const long long bidx = tok / S;
    const long long s = tok - bidx * S;
    delta[(bidx * hq + h) * S + s] = acc;
  }
}

// dqkv[:, 0:hq*D] (bf16) <- dq_acc (fp32)
__global__ void attn_dq_convert_kernel(const float* __restrict__ dq_acc, __nv_bfloat16* __restrict__ dqkv, long long tokens,
                                       int qcols, int row_elems) {
  const long long total = tokens * (qcols / 4);
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long t = i / (qcols / 4);
    const int c = static_cast<int>(i - t * (qcols / 4)) * 4;
    float4 v = *reinterpret_cast<const float4*>(dq_acc + t * qcols + c);
    uint2 o;
    o.x = pack_bf16x2(v.x, v.y);
    o.y = pack_bf16x2(v.z, v.w);
    *reinterpret_cast<uint2*>(dqkv + t * row_elems + c) = o;
  }
}

template <int D>
__global__ void __launch_bounds__(BW_THREADS, 1)
flash_bwd_kernel(const __grid_constant__ CUtensorMap tmQKV, const __grid_constant__ CUtensorMap tmDO,
                 const __grid_constant__ CUtensorMap tmDQ, const __grid_constant__ CUtensorMap tmDQKV, const BwdParams p) {
  using SMEM = BwdSmem<D>;
  constexpr int NCH = SMEM::NCH;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + SMEM::BAR_OFF);
  uint64_t* kv_full = bars + 0;
  uint64_t* q_full = bars + 1;    // [2]
  uint64_t* q_empty = bars + 3;   // [2]
  uint64_t* do_full = bars + 5;
  uint64_t* do_empty = bars + 6;  // dO is dead once the dV MMAs retire: its reload overlaps the dK / dQ MMAs
  uint64_t* sdp_full = bars + 7;
  uint64_t* pds_full = bars + 8;
  uint64_t* mma2_done = bars + 9;
  uint64_t* dq_drained = bars + 10;
  uint64_t* q_stage_free = bars + 11;  // [2] the dQ staging that borrowed Q slot st has been read by the TMA engine
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 14);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int jb = blockIdx.x;  // key block
  const int kv_head = blockIdx.y, batch = blockIdx.z;
  const int group = p.hq / p.hkv;
  const int n_qb = p.S / 128;
  const int i_first = p.causal ? jb : 0;
  const int steps_per_head = n_qb - i_first;
  const int n_steps = group * steps_per_head;
  const int krow0 = batch * p.S + jb * 128;
  const int k_col = (p.hq + kv_head) * D;
  const int v_col = (p.hq + p.hkv + kv_head) * D;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmQKV);
    tma_prefetch_desc(&tmDO);
    tma_prefetch_desc(&tmDQ);
    tma_prefetch_desc(&tmDQKV);
    mbar_init(kv_full, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&q_full[i], 1);
      mbar_init(&q_empty[i], 1);
    }
    mbar_init(do_full, 1);
    mbar_init(do_empty, 1);
    mbar_init(sdp_full, 1);
    mbar_init(pds_full, BW_SM_THREADS);
    mbar_init(mma2_done, 1);
    mbar_init(dq_drained, BW_SM_THREADS);
    mbar_init(&q_stage_free[0], 1);
    mbar_init(&q_stage_free[1], 1);
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
  const uint32_t TM_S = tmem_base, TM_DP = tmem_base + 128, TM_DV = tmem_base + 256, TM_DK = tmem_base + 384;

  if (warp == 0) {
    if (lane == 0) {
      mbar_expect_tx(kv_full, 2 * SMEM::TILE);
#pragma unroll
      for (int c = 0; c < NCH; ++c) {
        tma_load_2d(&tmQKV, kv_full, smem + SMEM::K_OFF + c * BW_CHUNK, k_col + c * 64, krow0);
        tma_load_2d(&tmQKV, kv_full, smem + SMEM::V_OFF + c * BW_CHUNK, v_col + c * 64, krow0);
      }
      for (int t = 0; t < n_steps; ++t) {
        const int g = t / steps_per_head;
        const int i = i_first + (t - g * steps_per_head);
        const int head = kv_head * group + g;
        const int qrow0 = batch * p.S + i * 128;
        const int st = t & 1;
        mbar_wait(&q_empty[st], ((t >> 1) & 1) ^ 1);
        if (t >= 2) mbar_wait(&q_stage_free[st], ((t >> 1) - 1) & 1);  // dQ(t-2) was staged through this slot
        mbar_expect_tx(&q_full[st], SMEM::TILE);
#pragma unroll
        for (int c = 0; c < NCH; ++c)
          tma_load_2d(&tmQKV, &q_full[st], smem + SMEM::Q_OFF + st * SMEM::TILE + c * BW_CHUNK, head * D + c * 64, qrow0);
        mbar_wait(do_empty, (t & 1) ^ 1);
        mbar_expect_tx(do_full, SMEM::TILE);
#pragma unroll
        for (int c = 0; c < NCH; ++c) tma_load_2d(&tmDO, do_full, smem + SMEM::DO_OFF + c * BW_CHUNK, head * D + c * 64, qrow0);
      }
    }
  } else if (warp == 1) {
    // The whole warp walks the loop convergently and one elected lane issues; operands are formed from kernel-uniform
    // values with a low-word add per MMA (an `if (lane == 0)` region made every descriptor a per-thread value that had
    // to be broadcast into uniform registers: ~100 cycles of issue overhead per 64-cycle MMA).
    {
      constexpr uint32_t idesc_s = umma_idesc(128, 128, 1, 1, 0, 0);   // Q K^T, dO V^T : both K-major
      constexpr uint32_t idesc_dkv = umma_idesc(128, D, 1, 1, 1, 1);   // P^T dO, dS^T Q: both MN-major
      constexpr uint32_t idesc_dq = umma_idesc(128, D, 1, 1, 0, 1);    // dS K          : A K-major, B MN-major
      constexpr uint32_t HI = (1024u >> 4) | (1u << 14) | (2u << 29);  // SBO 1024, descriptor version, SWIZZLE_128B
      constexpr uint32_t KM = (16u >> 4) << 16;                         // K-major : LBO 16
      constexpr uint32_t MN = (uint32_t(BW_CHUNK) >> 4) << 16;          // MN-major: LBO = chunk stride
      auto desc = [](uint32_t lo) { return (static_cast<uint64_t>(HI) << 32) | lo; };
      const uint32_t tm = __shfl_sync(0xffffffffu, tmem_base, 0);
      const bool leader = elect_one();
      const uint32_t base = smem_u32(smem);
      const uint32_t k_a = (base + SMEM::K_OFF) >> 4, v_a = (base + SMEM::V_OFF) >> 4, do_a = (base + SMEM::DO_OFF) >> 4;
      const uint32_t p_a = (base + SMEM::P_OFF) >> 4, ds_a = (base + SMEM::DS_OFF) >> 4;
      bwd_wait(kv_full, 0);
      for (int t = 0; t < n_steps; ++t) {
        const int st = t & 1;
        const uint32_t q_a = (base + SMEM::Q_OFF + st * SMEM::TILE) >> 4;
        bwd_wait(&q_full[st], (t >> 1) & 1);
        if (t > 0) bwd_wait(dq_drained, (t - 1) & 1);  // dQ of the previous step has left TMEM [0,128)
        tc_fence_after();
        if (leader) {
          BWD_TRACE(0, t)
#pragma unroll
          for (int kk = 0; kk < D / 16; ++kk) {
            const uint32_t off = ((kk >> 2) * BW_CHUNK + (kk & 3) * 32) >> 4;
            umma_f16<1>(tm, desc(KM | (q_a + off)), desc(KM | (k_a + off)), idesc_s, kk != 0);
          }
        }
        bwd_wait(do_full, t & 1);
        tc_fence_after();
        if (leader) {
#pragma unroll
          for (int kk = 0; kk < D / 16; ++kk) {
            const uint32_t off = ((kk >> 2) * BW_CHUNK + (kk & 3) * 32) >> 4;
            umma_f16<1>(tm + 128, desc(KM | (do_a + off)), desc(KM | (v_a + off)), idesc_s, kk != 0);
          }
          umma_commit<1>(sdp_full);
          BWD_TRACE(1, t)
        }
        bwd_wait(pds_full, t & 1);
        tc_fence_after();
        if (leader) {
          BWD_TRACE(2, t)
          // reduction dimension = the 128 q rows: 8 k-steps of 16 rows (2048 B apart in every [rows x 64] chunk)
#pragma unroll
          for (int kk = 0; kk < 8; ++kk)
            umma_f16<1>(tm + 256, desc(MN | (p_a + kk * 128)), desc(MN | (do_a + kk * 128)), idesc_dkv, (t | kk) != 0);
          umma_commit<1>(do_empty);  // dO(t) fully consumed: the producer reloads it while dK / dQ run
#pragma unroll
          for (int kk = 0; kk < 8; ++kk)
            umma_f16<1>(tm + 384, desc(MN | (ds_a + kk * 128)), desc(MN | (q_a + kk * 128)), idesc_dkv, (t | kk) != 0);
          // dQ = dS K : reduction dimension = 128 keys
#pragma unroll
          for (int kk = 0; kk < 8; ++kk) {
            const uint32_t a_off = ((kk >> 2) * BW_CHUNK + (kk & 3) * 32) >> 4;
            umma_f16<1>(tm, desc(KM | (ds_a + a_off)), desc(MN | (k_a + kk * 128)), idesc_dq, kk != 0);
          }
          umma_commit<1>(mma2_done);
          umma_commit<1>(&q_empty[st]);
          BWD_TRACE(3, t)
        }
      }
    }
  } else {
    const int q4 = warp & 3;
    constexpr int half = 0, NSPLIT = 1;  // column split across softmax threads (kept general: NSPLIT = 2 needs 8 warps)
    const int row = q4 * 32 + lane;
    const uint32_t lane_off = static_cast<uint32_t>(q4 * 32) << 16;
    const int sw = row & 7;
    const int epi_tid = threadIdx.x - 64;
    // lse / delta of step t+1 are fetched one step ahead: with the dQ reduce-adds of every CTA in flight a global load
    // takes thousands of cycles, which used to land in the middle of the P / dS computation
    auto stat_index = [&](int t) {
      const int g = t / steps_per_head;
      const int i = i_first + (t - g * steps_per_head);
      return (static_cast<size_t>(batch) * p.hq + kv_head * group + g) * p.S + i * 128 + row;
    };
    float lse_next = p.lse[stat_index(0)], dlt_next = p.delta[stat_index(0)];
    for (int t = 0; t < n_steps; ++t) {
      const int g = t / steps_per_head;
      const int i = i_first + (t - g * steps_per_head);
      const int head = kv_head * group + g;
      const float lse2 = lse_next * 1.4426950408889634f;
      const float dlt = dlt_next;
      if (t + 1 < n_steps) {
        const size_t nxt = stat_index(t + 1);
        lse_next = p.lse[nxt];
        dlt_next = p.delta[nxt];
      }
      const bool diag = p.causal && (i == jb);
      mbar_wait(sdp_full, t & 1);
      tc_fence_after();
      if (row == 0) { BWD_TRACE(4, t) }
      if (t > 0) {
        // the first half of dQ(t-1) was staged in the P buffer: the TMA engine must have read it before P(t) is stored
        // (the younger group -- the half in the Q slot -- may still be pending)
        if (epi_tid == 0) tma_wait_group_read<1>();
        named_bar_sync(1, BW_SM_THREADS);
      }
      if (row == 0) { BWD_TRACE(9, t) }
      // P = exp2(S c - lse), dS = P (dP - delta) scale with packed fp32x2 math (FFMA2 / FMUL2: half the issue slots)
      const uint64_t sc2 = bw_pack2(p.scale_log2, p.scale_log2), nlse2 = bw_pack2(-lse2, -lse2);
      const uint64_t scl2 = bw_pack2(p.scale, p.scale), ndlt2 = bw_pack2(-dlt * p.scale, -dlt * p.scale);
#pragma unroll
      for (int cc = 0; cc < 4 / NSPLIT; ++cc) {
        const int c = half * (4 / NSPLIT) + cc;
        uint32_t sv[32], dv[32];
        tmem_ld_32x32(TM_S + c * 32 + lane_off, sv);
        tmem_ld_32x32(TM_DP + c * 32 + lane_off, dv);
        tmem_ld_wait();
        if (diag) {
#pragma unroll
          for (int k = 0; k < 32; ++k)
            if (c * 32 + k > row) sv[k] = 0xff800000u;  // -inf -> P = 0, dS = 0
        }
        uint8_t* prow = smem + SMEM::P_OFF + (c >> 1) * BW_CHUNK + row * 128;
        uint8_t* dsrow = smem + SMEM::DS_OFF + (c >> 1) * BW_CHUNK + row * 128;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          uint32_t pp[4], dd[4];
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const int e = u * 8 + 2 * k;
            const uint64_t x2 = bw_fma2(bw_pack2(__uint_as_float(sv[e]), __uint_as_float(sv[e + 1])), sc2, nlse2);
            float x0, x1;
            bw_unpack2(x2, x0, x1);
            const float p0 = bwd_ex2(x0), p1 = bwd_ex2(x1);
            const uint64_t g2 = bw_fma2(bw_pack2(__uint_as_float(dv[e]), __uint_as_float(dv[e + 1])), scl2, ndlt2);
            float d0, d1;
            bw_unpack2(bw_mul2(bw_pack2(p0, p1), g2), d0, d1);
            pp[k] = pack_bf16x2(p0, p1);
            dd[k] = pack_bf16x2(d0, d1);
          }
          const int unit = (c & 1) * 4 + u;
          *reinterpret_cast<uint4*>(prow + ((unit ^ sw) << 4)) = make_uint4(pp[0], pp[1], pp[2], pp[3]);
          *reinterpret_cast<uint4*>(dsrow + ((unit ^ sw) << 4)) = make_uint4(dd[0], dd[1], dd[2], dd[3]);
        }
      }
      if (row == 0) { BWD_TRACE(10, t) }
      fence_proxy_async_smem();
      tc_fence_before();
      mbar_arrive(pds_full);
      if (row == 0) { BWD_TRACE(5, t) }
      if (t > 0 && epi_tid == 0) {
        // dQ(t-1) went out through Q slot (t-1)&1 a whole softmax phase ago: its TMA reads are long done, so this wait is free;
        // the producer may now refill the slot with Q(t+1)
        tma_wait_group_read<0>();
        mbar_arrive(&q_stage_free[(t - 1) & 1]);
      }

      // ---- drain dQ: TMEM [0,D) fp32 -> registers; the TMEM region is then free, so the MMA warp starts S/dP of the next step
      // while this warp group sends dQ out as swizzled fp32 tiles summed into the fp32 scratch by TMA reduce-add.  The
      // reduce-adds of all CTAs are bound by the L2 atomic rate, so the engine reads its smem source slowly (~1300 cycles per
      // 32 KB): the staging must not be needed soon.  Half of the columns go through the P buffer -- read while the next
      // step's S/dP MMAs run, i.e. before the next softmax phase stores P -- and half through the Q slot of THIS step, dead
      // once mma2 retired and refilled a full step later (the producer waits for q_stage_free).
      // Measured alternatives: red.global.add.v4.f32 straight from the registers (no staging) spends 4200 cycles per step
      // issuing half-sector atomics against ~2600 for the TMA path -- the per-SM atomic rate of ~25 B/clk is the floor either
      // way; eight softmax warps (two threads per row) leave the P/dS phase unchanged at ~2300 cycles: it is bound by MUFU +
      // FMA-pipe throughput (which do not overlap), not by latency.
      mbar_wait(mma2_done, t & 1);
      tc_fence_after();
      if (row == 0) { BWD_TRACE(6, t) }
      uint8_t* stage = smem + SMEM::Q_OFF + (t & 1) * SMEM::TILE;
      constexpr int DQ_PER = D / 32;  // 32-column chunks of dQ per thread
      uint32_t dq[DQ_PER][32];
#pragma unroll
      for (int cc = 0; cc < DQ_PER; ++cc) tmem_ld_32x32(TM_S + cc * 32 + lane_off, dq[cc]);
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(dq_drained);
      if (row == 0) { BWD_TRACE(7, t) }
      const int qrow0 = batch * p.S + i * 128;
      // first half -> P buffer (its reads finish under the next step's S/dP MMAs, before the next softmax phase stores P),
      // second half -> Q slot of this step (refilled a full step later); the two reduce-adds are issued back to back
#pragma unroll
      for (int hh = 0; hh < 2; ++hh) {
        uint8_t* stg = hh == 0 ? smem + SMEM::P_OFF : stage;
#pragma unroll
        for (int cc = 0; cc < DQ_PER / 2; ++cc) {
          uint8_t* rowp = stg + cc * BW_CHUNK + row * 128;
          const int src = hh * (DQ_PER / 2) + cc;
#pragma unroll
          for (int u = 0; u < 8; ++u)
            *reinterpret_cast<uint4*>(rowp + ((u ^ sw) << 4)) = make_uint4(dq[src][4 * u], dq[src][4 * u + 1], dq[src][4 * u + 2], dq[src][4 * u + 3]);
        }
      }
      fence_proxy_async_smem();
      named_bar_sync(1, BW_SM_THREADS);
      if (epi_tid == 0) {
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
          uint8_t* stg = hh == 0 ? smem + SMEM::P_OFF : stage;
          if (!p.debug_skip_dq)
            for (int cc = 0; cc < DQ_PER / 2; ++cc)
              tma_reduce_add_2d(&tmDQ, stg + cc * BW_CHUNK, head * D + (hh * (DQ_PER / 2) + cc) * 32, qrow0);
          tma_commit_group();  // one group per half: the P-buffer half is awaited alone at the top of the next step
        }
      }
      if (row == 0) { BWD_TRACE(8, t) }
    }

    // ---- dK, dV epilogue: TMEM -> bf16 -> swizzled staging (Q / dO tiles are dead) -> TMA store into packed dqkv
    if (epi_tid == 0) tma_wait_group_read<0>();  // the last dQ halves have left the Q slots
    named_bar_sync(1, BW_SM_THREADS);
    tc_fence_after();
#pragma unroll
    for (int which = 0; which < 2; ++which) {
      const uint32_t tm = which == 0 ? TM_DK : TM_DV;
      uint8_t* stage = smem + (which == 0 ? SMEM::Q_OFF : SMEM::DO_OFF);
#pragma unroll
      for (int cc = 0; cc < D / 32 / NSPLIT; ++cc) {
        const int c = half * (D / 32 / NSPLIT) + cc;
        uint32_t v[32];
        tmem_ld_32x32(tm + c * 32 + lane_off, v);
        tmem_ld_wait();
        uint8_t* rowp = stage + (c >> 1) * BW_CHUNK + row * 128;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          uint4 val;
          val.x = pack_bf16x2(__uint_as_float(v[8 * u + 0]), __uint_as_float(v[8 * u + 1]));
          val.y = pack_bf16x2(__uint_as_float(v[8 * u + 2]), __uint_as_float(v[8 * u + 3]));
          val.z = pack_bf16x2(__uint_as_float(v[8 * u + 4]), __uint_as_float(v[8 * u + 5]));
          val.w = pack_bf16x2(__uint_as_float(v[8 * u + 6]), __uint_as_float(v[8 * u + 7]));
          const int unit = (c & 1) * 4 + u;
          *reinterpret_cast<uint4*>(rowp + ((unit ^ sw) << 4)) = val;
        }
      }
    }
    fence_proxy_async_smem();
    named_bar_sync(1, BW_SM_THREADS);
    if (epi_tid == 0) {
#pragma unroll
      for (int c = 0; c < NCH; ++c) {
        tma_store_2d(&tmDQKV, smem + SMEM::Q_OFF + c * BW_CHUNK, k_col + c * 64, krow0);
        tma_store_2d(&tmDQKV, smem + SMEM::DO_OFF + c * BW_CHUNK, v_col + c * 64, krow0);
      }
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

template <int D>
static int launch_bwd(const void* dout, const void* qkv, const void* out, const float* lse, void* dqkv, float* dq_acc,
                      float* delta, int B, int S, int hq, int hkv, int causal, float scale, cudaStream_t stream) {
  using SMEM = BwdSmem<D>;
  const long long tokens = static_cast<long long>(B) * S;
  const int row_elems = (hq + 2 * hkv) * D;
  {
    const long long warps = tokens * hq;
    const int threads = 256;
    const long long blocks = (warps * 32 + threads - 1) / threads;
    attn_delta_kernel<<<static_cast<unsigned>(blocks), threads, 0, stream>>>(
        static_cast<const __nv_bfloat16*>(dout), static_cast<const __nv_bfloat16*>(out), delta, B, S, hq, D);
  }
  CUtensorMap tmQKV, tmDO, tmDQ, tmDQKV;
  int rc = make_tensor_map_2d(&tmQKV, TMA_BF16, qkv, tokens, row_elems, row_elems, 128, 64, TMA_SW_128);
  if (rc) return rc;
  rc = make_tensor_map_2d(&tmDO, TMA_BF16, dout, tokens, hq * D, hq * D, 128, 64, TMA_SW_128);
  if (rc) return rc;
  rc = make_tensor_map_2d(&tmDQ, TMA_F32, dq_acc, tokens, hq * D, hq * D, 128, 32, TMA_SW_128);
  if (rc) return rc;
  rc = make_tensor_map_2d(&tmDQKV, TMA_BF16, dqkv, tokens, row_elems, row_elems, 128, 64, TMA_SW_128);
  if (rc) return rc;
  auto kern = flash_bwd_kernel<D>;
  static bool attr_set = false;
  if (!attr_set) {
    DSB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM::TOTAL));
    attr_set = true;
  }
  BwdParams p;
  p.B = B; p.S = S; p.hq = hq; p.hkv = hkv; p.causal = causal;
  p.scale = scale;
  p.scale_log2 = scale * 1.4426950408889634f;
  p.lse = lse;
  p.delta = delta;
  static const int skip_dq = [] { const char* e = getenv("DSB200_BWD_SKIP_DQ"); return e ? atoi(e) : 0; }();
  p.debug_skip_dq = skip_dq;
  p.trace = g_bwd_trace;
  dim3 grid(S / 128, hkv, B);
  kern<<<grid, BW_THREADS, SMEM::TOTAL, stream>>>(tmQKV, tmDO, tmDQ, tmDQKV, p);
  DSB_CHECK_CUDA(cudaGetLastError());
  const int qcols = hq * D;
  long long work = tokens * (qcols / 4);
  int blocks = static_cast<int>((work + 255) / 256 < 148 * 16 ? (work + 255) / 256 : 148 * 16);
  attn_dq_convert_kernel<<<blocks, 256, 0, stream>>>(dq_acc, static_cast<__nv_bfloat16*>(dqkv), tokens, qcols, row_elems);
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

}  // namespace attn
void flash_bwd_set_trace(long long* buf) { attn::g_bwd_trace = buf; }
}  // namespace dsb

// Gradient of dsb200_flash_attn_fwd.  dout/out: [B*S, hq*D] bf16; qkv, dqkv: packed [B*S, (hq+2*hkv)*D] bf16;
// lse: [B, hq, S] from the forward.  Scratch supplied by the caller: dq_acc [B*S, hq*D] fp32 ZERO-INITIALISED,
// delta [B, hq, S] fp32.
DSB_EXPORT int dsb200_flash_attn_bwd(const void* dout, const void* qkv, const void* out, const float* lse, void* dqkv,
                                     float* dq_acc, float* delta, int B, int S, int hq, int hkv, int head_dim, int causal,
                                     float scale, cudaStream_t stream) {
  using namespace dsb;
  DSB_REQUIRE(head_dim == 64 || head_dim == 128, "flash_attn_bwd: head_dim %d unsupported (64 or 128)", head_dim);
  DSB_REQUIRE(S % 128 == 0 && S > 0, "flash_attn_bwd: sequence length %d must be a multiple of 128", S);
  DSB_REQUIRE(hq % hkv == 0, "flash_attn_bwd: hq %d not a multiple of hkv %d", hq, hkv);
  if (head_dim == 128)
    return attn::launch_bwd<128>(dout, qkv, out, lse, dqkv, dq_acc, delta, B, S, hq, hkv, causal, scale, stream);
  return attn::launch_bwd<64>(dout, qkv, out, lse, dqkv, dq_acc, delta, B, S, hq, hkv, causal, scale, stream);
}

// Flash attention forward for sm_100a on tcgen05 tensor cores (training path: causal / full, MHA / GQA, head_dim 64/128).
//
// Reads Q, K, V straight out of the packed QKV projection [tokens, (hq + 2*hkv) * D] with TMA (no head transposes), one
// CTA per (batch, q-head, 128-row q block):
//   warp 0   TMA producer : Q once, then (K,V) blocks of 128 keys through a 2-stage ring
//   warp 1   MMA issuer   : S = Q K^T  (128x128xD, SS)  into TMEM S[j&1];  PV = P V (128xDx128, P from smem, V MN-major)
//                           into TMEM PV[j&1]
//   warps 2-5 softmax     : thread == row.  tcgen05.ld S row -> online max/sum -> P (bf16) to 128B-swizzled smem ->
//                           one block later: O = (O + PV) * alpha, O kept in registers (no TMEM read-modify-write)
// Epilogue: O / l -> bf16 -> swizzled smem -> TMA store into [tokens, hq*D]; LSE (natural log) saved for backward.
// Replaces csrc/transformer/softmax_kernels.cu + the strided-batched cuBLAS QK^T / PV of the reference's training
// kernel (ds_transformer_cuda.cpp) and the user-supplied FlashAttention of its Ulysses path (sequence/layer.py:84).
#include <stdlib.h>

#include "../common/host.h"
#include "../common/ptx.cuh"

namespace dsb {
namespace attn {

constexpr int BM = 128;  // q rows per CTA
constexpr int BN = 128;  // keys per block
constexpr int NUM_THREADS = 192;
constexpr int SM_THREADS = 128;  // softmax threads
constexpr int CHUNK_BYTES = 128 * 128;  // one [128 rows x 64 bf16] swizzled chunk

template <int D>
struct FwdSmem {
  static constexpr int NCH = D / 64;                    // 64-column chunks per Q/K/V tile
  static constexpr int TILE = NCH * CHUNK_BYTES;        // Q, K or V tile bytes
  static constexpr int P_BYTES = 2 * CHUNK_BYTES;       // P tile: 128 x 128 bf16
  static constexpr int Q_OFF = 0;
  static constexpr int K_OFF = Q_OFF + TILE;            // 2 stages
  static constexpr int V_OFF = K_OFF + 2 * TILE;        // 2 stages
  static constexpr int P_OFF = V_OFF + 2 * TILE;        // 2 buffers
  static constexpr int BAR_OFF = P_OFF + 2 * P_BYTES;
  static constexpr int TOTAL = BAR_OFF + 256 + 1024;
};

struct FwdParams {
  int B, S, hq, hkv;
  int causal;
  float scale_log2;  // softmax scale * log2(e)
  float* lse;        // [B, hq, S]
};

template <int D>
__global__ void __launch_bounds__(NUM_THREADS, 1)
flash_fwd_kernel(const __grid_constant__ CUtensorMap tmQKV, const __grid_constant__ CUtensorMap tmO, const FwdParams p) {
  using SMEM = FwdSmem<D>;
  constexpr int NCH = SMEM::NCH;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + SMEM::BAR_OFF);
  uint64_t* q_full = bars + 0;
  uint64_t* kv_full = bars + 1;    // [2]
  uint64_t* kv_empty = bars + 3;   // [2]
  uint64_t* s_full = bars + 5;     // [2]  MMA -> softmax
  uint64_t* p_full = bars + 7;     // [2]  softmax -> MMA (P written, S consumed)
  uint64_t* o_full = bars + 9;     // [2]  MMA -> softmax (PV complete)
  uint64_t* o_free = bars + 11;    // [2]  softmax -> MMA (PV buffer read)
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 13);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // heavy (late) q blocks first: causal work per block grows with the block index
  const int n_qb = p.S / BM;
  const int qb = n_qb - 1 - static_cast<int>(blockIdx.x);
  const int head = blockIdx.y, batch = blockIdx.z;
  const int kv_head = head / (p.hq / p.hkv);
  const int n_blocks = p.causal ? (qb + 1) : (p.S / BN);
  const int row0 = batch * p.S + qb * BM;           // first token row of this q block
  const int q_col = head * D;
  const int k_col = (p.hq + kv_head) * D;
  const int v_col = (p.hq + p.hkv + kv_head) * D;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmQKV);
    tma_prefetch_desc(&tmO);
    mbar_init(q_full, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&kv_full[i], 1);
      mbar_init(&kv_empty[i], 1);
      mbar_init(&s_full[i], 1);
      mbar_init(&p_full[i], SM_THREADS);
      mbar_init(&o_full[i], 1);
      mbar_init(&o_free[i], SM_THREADS);
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
  const uint32_t TM_S = tmem_base;           // S[b]  at columns b*128
  const uint32_t TM_PV = tmem_base + 256;    // PV[b] at columns 256 + b*128

  if (warp == 0) {
    if (lane == 0) {
      mbar_expect_tx(q_full, SMEM::TILE);
#pragma unroll
      for (int c = 0; c < NCH; ++c) tma_load_2d(&tmQKV, q_full, smem + SMEM::Q_OFF + c * CHUNK_BYTES, q_col + c * 64, row0);
      for (int j = 0; j < n_blocks; ++j) {
        const int st = j & 1;
        mbar_wait(&kv_empty[st], ((j >> 1) & 1) ^ 1);
        mbar_expect_tx(&kv_full[st], 2 * SMEM::TILE);
        const int krow = batch * p.S + j * BN;
#pragma unroll
        for (int c = 0; c < NCH; ++c) {
          tma_load_2d(&tmQKV, &kv_full[st], smem + SMEM::K_OFF + st * SMEM::TILE + c * CHUNK_BYTES, k_col + c * 64, krow);
          tma_load_2d(&tmQKV, &kv_full[st], smem + SMEM::V_OFF + st * SMEM::TILE + c * CHUNK_BYTES, v_col + c * 64, krow);
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc_qk = umma_idesc(BM, BN, 1, 1, 0, 0);
      constexpr uint32_t idesc_pv = umma_idesc(BM, D, 1, 1, 0, 1);  // A = P (K-major), B = V (MN-major)
      const uint32_t q_s = smem_u32(smem + SMEM::Q_OFF);
      auto issue_qk = [&](int j) {
        const int st = j & 1;
        mbar_wait(&kv_full[st], (j >> 1) & 1);
        tc_fence_after();
        const uint32_t k_s = smem_u32(smem + SMEM::K_OFF + st * SMEM::TILE);
#pragma unroll
        for (int kk = 0; kk < D / 16; ++kk) {
          const uint32_t off = (kk >> 2) * CHUNK_BYTES + (kk & 3) * 32;
          umma_f16<1>(TM_S + st * 128, umma_smem_desc(q_s + off, 16, 1024), umma_smem_desc(k_s + off, 16, 1024), idesc_qk,
                      kk != 0 ? 1u : 0u);
        }
        umma_commit<1>(&s_full[st]);
      };
      auto issue_pv = [&](int j) {
        const int st = j & 1;
        mbar_wait(&p_full[st], (j >> 1) & 1);            // P(j) in smem, S(j) consumed
        if (j >= 2) mbar_wait(&o_free[st], ((j >> 1) - 1) & 1);  // PV[st] of block j-2 has been read
        tc_fence_after();
        const uint32_t p_s = smem_u32(smem + SMEM::P_OFF + st * SMEM::P_BYTES);
        const uint32_t v_s = smem_u32(smem + SMEM::V_OFF + st * SMEM::TILE);
#pragma unroll
        for (int kk = 0; kk < BN / 16; ++kk) {
          const uint32_t a_off = (kk >> 2) * CHUNK_BYTES + (kk & 3) * 32;
          umma_f16<1>(TM_PV + st * 128, umma_smem_desc(p_s + a_off, 16, 1024),
                      umma_smem_desc(v_s + kk * 2048, CHUNK_BYTES, 1024), idesc_pv, kk != 0 ? 1u : 0u);
        }
        umma_commit<1>(&o_full[st]);
        umma_commit<1>(&kv_empty[st]);  // K(j) and V(j) are no longer needed once PV(j) retires
      };
      mbar_wait(q_full, 0);
      issue_qk(0);
      for (int j = 0; j < n_blocks; ++j) {
        if (j + 1 < n_blocks) issue_qk(j + 1);   // S[(j+1)&1] is free: P(j-1) was awaited before PV(j-1)
        issue_pv(j);
      }
    }
  } else {
    // ------------------------------------------------------------- softmax + output (thread == row)
    const int q4 = warp & 3;
    const int row = q4 * 32 + lane;
    const uint32_t lane_off = static_cast<uint32_t>(q4 * 32) << 16;
    float o_acc[D];
#pragma unroll
    for (int i = 0; i < D; ++i) o_acc[i] = 0.f;
    float m_run = -INFINITY, l_run = 0.f;

    auto accumulate_pv = [&](int j, float alpha) {
      const int st = j & 1;
      mbar_wait(&o_full[st], (j >> 1) & 1);
      tc_fence_after();
#pragma unroll
      for (int c = 0; c < D / 32; ++c) {
        uint32_t v[32];
        tmem_ld_32x32(TM_PV + st * 128 + c * 32 + lane_off, v);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) o_acc[c * 32 + i] = (o_acc[c * 32 + i] + __uint_as_float(v[i])) * alpha;
      }
      tc_fence_before();
      mbar_arrive(&o_free[st]);
    };

    for (int j = 0; j < n_blocks; ++j) {
      const int st = j & 1;
      mbar_wait(&s_full[st], (j >> 1) & 1);
      tc_fence_after();
      const bool diag = p.causal && (j == qb);
      // pass 1: row max of the scaled scores (S is re-read from TMEM in pass 2: keeps the register budget for O)
      float m_blk = -INFINITY;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        uint32_t sv[32];
        tmem_ld_32x32(TM_S + st * 128 + c * 32 + lane_off, sv);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          float v = __uint_as_float(sv[i]);
          if (diag && (c * 32 + i) > row) v = -INFINITY;
          m_blk = fmaxf(m_blk, v);
        }
      }
      const float m_new = fmaxf(m_run, m_blk * p.scale_log2);
      const float alpha = exp2f(m_run - m_new);  // m_run = -inf on the first block -> 0 (o_acc, l are 0 anyway)
      float l_blk = 0.f;
      uint8_t* pbuf = smem + SMEM::P_OFF + st * SMEM::P_BYTES;
      const int sw = row & 7;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        uint32_t sv[32];
        tmem_ld_32x32(TM_S + st * 128 + c * 32 + lane_off, sv);
        tmem_ld_wait();
        uint8_t* rowp = pbuf + (c >> 1) * CHUNK_BYTES + row * 128;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          float e[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const float v = fmaf(__uint_as_float(sv[u * 8 + i]), p.scale_log2, -m_new);
            e[i] = (diag && (c * 32 + u * 8 + i) > row) ? 0.f : exp2f(v);
            l_blk += e[i];
          }
          uint4 val;
          val.x = pack_bf16x2(e[0], e[1]);
          val.y = pack_bf16x2(e[2], e[3]);
          val.z = pack_bf16x2(e[4], e[5]);
          val.w = pack_bf16x2(e[6], e[7]);
          const int unit = (c & 1) * 4 + u;  // 16-byte unit inside the 128-byte row of this chunk
          *reinterpret_cast<uint4*>(rowp + ((unit ^ sw) << 4)) = val;
        }
      }
      l_run = l_run * alpha + l_blk;
      m_run = m_new;
      fence_proxy_async_smem();
      tc_fence_before();
      mbar_arrive(&p_full[st]);
      // fold in PV(j-1) (relative to the previous max) and move O to the new max
      if (j > 0) {
        accumulate_pv(j - 1, alpha);
      }
    }
    accumulate_pv(n_blocks - 1, 1.f);

    // ---- epilogue: O / l -> bf16 -> swizzled staging (Q tile region is free: every QK MMA has retired) -> TMA store
    const float inv_l = 1.f / l_run;
    uint8_t* obuf = smem + SMEM::Q_OFF;
    const int sw = row & 7;
#pragma unroll
    for (int c = 0; c < NCH; ++c) {
      uint8_t* rowp = obuf + c * CHUNK_BYTES + row * 128;
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        uint4 val;
        const float* o = &o_acc[c * 64 + u * 8];
        val.x = pack_bf16x2(o[0] * inv_l, o[1] * inv_l);
        val.y = pack_bf16x2(o[2] * inv_l, o[3] * inv_l);
        val.z = pack_bf16x2(o[4] * inv_l, o[5] * inv_l);
        val.w = pack_bf16x2(o[6] * inv_l, o[7] * inv_l);
        *reinterpret_cast<uint4*>(rowp + ((u ^ sw) << 4)) = val;
      }
    }
    if (p.lse) p.lse[(static_cast<size_t>(batch) * p.hq + head) * p.S + qb * BM + row] = (m_run + log2f(l_run)) * 0.6931471805599453f;
    fence_proxy_async_smem();
    named_bar_sync(1, SM_THREADS);
    if (threadIdx.x == 64) {
#pragma unroll
      for (int c = 0; c < NCH; ++c) tma_store_2d(&tmO, obuf + c * CHUNK_BYTES, head * D + c * 64, row0);
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
static int launch_fwd(const void* qkv, void* out, float* lse, int B, int S, int hq, int hkv, int causal, float scale,
                      cudaStream_t stream) {
  using SMEM = FwdSmem<D>;
  CUtensorMap tmQKV, tmO;
  const int row_elems = (hq + 2 * hkv) * D;
  int rc = make_tensor_map_2d(&tmQKV, TMA_BF16, qkv, static_cast<uint64_t>(B) * S, row_elems, row_elems, 128, 64, TMA_SW_128);
  if (rc) return rc;
  rc = make_tensor_map_2d(&tmO, TMA_BF16, out, static_cast<uint64_t>(B) * S, hq * D, hq * D, 128, 64, TMA_SW_128);
  if (rc) return rc;
  auto kern = flash_fwd_kernel<D>;
  static bool attr_set = false;
  if (!attr_set) {
    DSB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM::TOTAL));
    attr_set = true;
  }
  FwdParams p;
  p.B = B; p.S = S; p.hq = hq; p.hkv = hkv; p.causal = causal;
  p.scale_log2 = scale * 1.4426950408889634f;
  p.lse = lse;
  dim3 grid(S / BM, hq, B);
  kern<<<grid, NUM_THREADS, SMEM::TOTAL, stream>>>(tmQKV, tmO, p);
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

}  // namespace attn

// second-generation kernel (flash_fwd2.cu): two Q tiles per CTA, P / O in TMEM
int flash_fwd2_launch(const void* qkv, void* out, float* lse, int B, int S, int hq, int hkv, int head_dim, int causal, float scale,
                      cudaStream_t stream);
void flash_fwd2_set_trace(long long* buf);
void flash_bwd_set_trace(long long* buf);
}  // namespace dsb

// debug hook: device buffer (>= 8192 int64) receiving clock64 stamps of one CTA of the next forward launches; nullptr = off
DSB_EXPORT int dsb200_flash_fwd_trace(long long* buf) {
  dsb::flash_fwd2_set_trace(buf);
  dsb::flash_bwd_set_trace(buf);
  return DSB_OK;
}

// out[B*S, hq*D] = softmax(Q K^T * scale [+ causal mask]) V with Q,K,V read from packed qkv [B*S, (hq+2*hkv)*D] (bf16).
// lse: optional [B, hq, S] fp32 (natural-log sum-exp of the scaled scores) for the backward pass.
DSB_EXPORT int dsb200_flash_attn_fwd(const void* qkv, void* out, float* lse, int B, int S, int hq, int hkv, int head_dim,
                                     int causal, float scale, cudaStream_t stream) {
  using namespace dsb;
  DSB_REQUIRE(head_dim == 64 || head_dim == 128, "flash_attn: head_dim %d unsupported (64 or 128)", head_dim);
  DSB_REQUIRE(S % 128 == 0 && S > 0, "flash_attn: sequence length %d must be a multiple of 128", S);
  DSB_REQUIRE(hq % hkv == 0, "flash_attn: hq %d not a multiple of hkv %d", hq, hkv);
  // DSB200_FLASH_FWD=1 forces the first-generation kernel (one Q tile per CTA); sequences that are not a multiple of 256 use it too
  static const int variant = [] {
    const char* e = getenv("DSB200_FLASH_FWD");
    return e ? atoi(e) : 0;
  }();
  if (variant != 1 && S % 256 == 0) return flash_fwd2_launch(qkv, out, lse, B, S, hq, hkv, head_dim, causal, scale, stream);
  if (head_dim == 128) return attn::launch_fwd<128>(qkv, out, lse, B, S, hq, hkv, causal, scale, stream);
  return attn::launch_fwd<64>(qkv, out, lse, B, S, hq, hkv, causal, scale, stream);
}

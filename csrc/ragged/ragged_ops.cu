// FastGen ragged-batch device ops for sm_100a:
//   kv_rotary_append   RoPE on q,k of a ragged batch + append k,v into the paged KV cache (block table lookup)
//   paged_attention    attention of every ragged token over its sequence's paged KV history (decode + chunked prefill)
//   ragged_embed       token (+ optional position) embedding lookup
//   logits_gather      pick the last token's hidden state of every sequence
//   top_k_gating / moe_scatter / moe_gather   MoE routing for inference (expert-sorted token order, weighted combine)
// Reference counterparts: inference/v2/kernels/ragged_ops/{linear_blocked_kv_rotary/blocked_kv_rotary.cu:24-36,250-259,
// blocked_flash (closed dskernels flash_attn_by_atoms), embed/embed.cu, logits_gather/logits_gather.cu,
// top_k_gating/top_k_gating.cu:93-105, moe_scatter/moe_scatter.cu:183-184, moe_gather/moe_gather.cu}.
// KV cache layout: [num_blocks][2 (k,v)][block_size][n_kv_heads][head_dim] (bf16).
#include "../common/host.h"
#include "../common/ptx.cuh"

namespace dsb {
namespace ragged {

typedef __nv_bfloat16 bf16;

// one warp per (token, head); q and k heads are rotated in place, k and v rows are copied into the cache
__global__ void kv_rotary_append_kernel(bf16* __restrict__ qkv, bf16* __restrict__ kv_cache, const int* __restrict__ tok_seq,
                                        const int* __restrict__ tok_pos, const int* __restrict__ block_table,
                                        const float* __restrict__ inv_freq, int n_tokens, int hq, int hkv, int D, int block_size,
                                        int max_blocks_per_seq, int rot_dim) {
  const int warps_per_block = blockDim.x >> 5;
  const long long gw = static_cast<long long>(blockIdx.x) * warps_per_block + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  const int heads = hq + 2 * hkv;
  if (gw >= static_cast<long long>(n_tokens) * heads) return;
  const int t = static_cast<int>(gw / heads);
  const int h = static_cast<int>(gw - static_cast<long long>(t) * heads);
  const int pos = tok_pos[t];
  bf16* src = qkv + static_cast<size_t>(t) * heads * D + static_cast<size_t>(h) * D;
  const int half = rot_dim / 2;  // rotate-half pairing inside the first rot_dim elements (partial rotary: rot_dim < D)
  if (h < hq + hkv && rot_dim > 0) {
    for (int j = lane; j < half; j += 32) {
      float s, c;
      sincosf(pos * inv_freq[j], &s, &c);
      const float a = __bfloat162float(src[j]), b = __bfloat162float(src[j + half]);
      src[j] = __float2bfloat16_rn(a * c - b * s);
      src[j + half] = __float2bfloat16_rn(b * c + a * s);
    }
    __syncwarp();
  }
  if (h >= hq) {
    const int is_v = h >= hq + hkv;
    const int kvh = is_v ? h - hq - hkv : h - hq;
    const int seq = tok_seq[t];
    const int blk = block_table[seq * max_blocks_per_seq + pos / block_size];
    bf16* dst = kv_cache + ((((static_cast<size_t>(blk) * 2 + is_v) * block_size + pos % block_size) * hkv + kvh) * D);
    for (int j = lane; j < D; j += 32) dst[j] = src[j];
  }
}

// One CTA (128 threads = 4 warps) per (token, kv head): the `group` q heads sharing the kv head are processed together.
// Each warp strides over the kv positions; per position a warp computes the `group` dot products (lanes split D),
// online softmax per (warp, q head), final cross-warp merge in shared memory.
template <int D, int MAXG>
__global__ void __launch_bounds__(128)
paged_attention_kernel(const bf16* __restrict__ qkv, const bf16* __restrict__ kv_cache, bf16* __restrict__ out,
                       const int* __restrict__ tok_seq, const int* __restrict__ tok_pos, const int* __restrict__ block_table,
                       int hq, int hkv, int block_size, int max_blocks_per_seq, float scale) {
  constexpr int EPL = D / 32;  // elements per lane
  const int t = blockIdx.x, kvh = blockIdx.y;
  const int group = hq / hkv;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int seq = tok_seq[t];
  const int n_kv = tok_pos[t] + 1;  // causal: keys 0..pos
  const int heads = hq + 2 * hkv;
  float q[MAXG][EPL];
#pragma unroll
  for (int g = 0; g < MAXG; ++g) {
    if (g < group) {
      const bf16* qp = qkv + static_cast<size_t>(t) * heads * D + static_cast<size_t>(kvh * group + g) * D + lane * EPL;
#pragma unroll
      for (int e = 0; e < EPL; ++e) q[g][e] = __bfloat162float(qp[e]) * scale;
    }
  }
  float m[MAXG], l[MAXG], acc[MAXG][EPL];
#pragma unroll
  for (int g = 0; g < MAXG; ++g) {
    m[g] = -INFINITY;
    l[g] = 0.f;
#pragma unroll
    for (int e = 0; e < EPL; ++e) acc[g][e] = 0.f;
  }
  const int* bt = block_table + seq * max_blocks_per_seq;
  for (int p = warp; p < n_kv; p += 4) {
    const int blk = bt[p / block_size];
    const size_t base = ((static_cast<size_t>(blk) * 2 * block_size + p % block_size) * hkv + kvh) * D + lane * EPL;
    const size_t vbase = base + static_cast<size_t>(block_size) * hkv * D;
    float k[EPL], v[EPL];
#pragma unroll
    for (int e = 0; e < EPL; ++e) {
      k[e] = __bfloat162float(kv_cache[base + e]);
      v[e] = __bfloat162float(kv_cache[vbase + e]);
    }
#pragma unroll
    for (int g = 0; g < MAXG; ++g) {
      if (g < group) {
        float s = 0.f;
#pragma unroll
        for (int e = 0; e < EPL; ++e) s = fmaf(q[g][e], k[e], s);
        s = warp_sum(s);
        const float mn = fmaxf(m[g], s);
        const float a = __expf(m[g] - mn), pe = __expf(s - mn);
        l[g] = l[g] * a + pe;
#pragma unroll
        for (int e = 0; e < EPL; ++e) acc[g][e] = acc[g][e] * a + pe * v[e];
        m[g] = mn;
      }
    }
  }
  // merge the 4 warps
  __shared__ float s_m[4][MAXG], s_l[4][MAXG];
  __shared__ float s_acc[4][MAXG][D];
#pragma unroll
  for (int g = 0; g < MAXG; ++g) {
    if (g < group) {
      if (lane == 0) { s_m[warp][g] = m[g]; s_l[warp][g] = l[g]; }
#pragma unroll
      for (int e = 0; e < EPL; ++e) s_acc[warp][g][lane * EPL + e] = acc[g][e];
    }
  }
  __syncthreads();
  for (int idx = threadIdx.x; idx < group * D; idx += 128) {
    const int g = idx / D, d = idx - g * D;
    float mm = fmaxf(fmaxf(s_m[0][g], s_m[1][g]), fmaxf(s_m[2][g], s_m[3][g]));
    float ll = 0.f, o = 0.f;
#pragma unroll
    for (int w = 0; w < 4; ++w) {
      const float a = (s_m[w][g] == -INFINITY) ? 0.f : __expf(s_m[w][g] - mm);
      ll += s_l[w][g] * a;
      o += s_acc[w][g][d] * a;
    }
    out[static_cast<size_t>(t) * hq * D + static_cast<size_t>(kvh * group + g) * D + d] = __float2bfloat16_rn(o / ll);
  }
}

__global__ void ragged_embed_kernel(const long long* __restrict__ ids, const bf16* __restrict__ wte, const bf16* __restrict__ wpe,
                                    const int* __restrict__ tok_pos, bf16* __restrict__ out, int n_tokens, int dim, int pos_offset) {
  const int t = blockIdx.y;
  const long long id = ids[t];
  const int p = tok_pos ? tok_pos[t] + pos_offset : 0;
  for (int c = (blockIdx.x * blockDim.x + threadIdx.x) * 8; c < dim; c += gridDim.x * blockDim.x * 8) {
    uint4 v = *reinterpret_cast<const uint4*>(wte + id * dim + c);
    if (wpe) {
      uint4 pv = *reinterpret_cast<const uint4*>(wpe + static_cast<size_t>(p) * dim + c);
      __nv_bfloat162* a = reinterpret_cast<__nv_bfloat162*>(&v);
      const __nv_bfloat162* b = reinterpret_cast<const __nv_bfloat162*>(&pv);
#pragma unroll
      for (int k = 0; k < 4; ++k) a[k] = __hadd2(a[k], b[k]);
    }
    *reinterpret_cast<uint4*>(out + static_cast<size_t>(t) * dim + c) = v;
  }
}

__global__ void logits_gather_kernel(const bf16* __restrict__ hidden, const int* __restrict__ last_tok, bf16* __restrict__ out,
                                     int dim) {
  const int s = blockIdx.y;
  const int t = last_tok[s];
  for (int c = (blockIdx.x * blockDim.x + threadIdx.x) * 8; c < dim; c += gridDim.x * blockDim.x * 8)
    *reinterpret_cast<uint4*>(out + static_cast<size_t>(s) * dim + c) = *reinterpret_cast<const uint4*>(hidden + static_cast<size_t>(t) * dim + c);
}

// One warp per token: softmax over E logits, top-k (k <= 8) by repeated arg-max, renormalised scores,
// expert_counts[e] += 1 and the token's slot inside the expert = the returned old count.
__global__ void top_k_gating_kernel(const float* __restrict__ logits, int* __restrict__ expert_counts, float* __restrict__ scores,
                                    int* __restrict__ assignments, int* __restrict__ offsets, int n_tokens, int E, int K,
                                    int normalize) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= n_tokens) return;
  const float* lg = logits + static_cast<size_t>(warp) * E;
  float mx = -INFINITY;
  for (int e = lane; e < E; e += 32) mx = fmaxf(mx, lg[e]);
  mx = warp_max(mx);
  float sum = 0.f;
  for (int e = lane; e < E; e += 32) sum += __expf(lg[e] - mx);
  sum = warp_sum(sum);
  unsigned long long taken = 0ull;  // experts < 64 tracked per lane-local ids below (E <= 2048 via per-lane mask)
  float top_sum = 0.f;
  float my_scores[8];
  int my_experts[8];
  for (int k = 0; k < K; ++k) {
    float best = -INFINITY;
    int best_e = -1;
    for (int e = lane; e < E; e += 32) {
      const int slot = e / 32;
      if ((taken >> slot) & 1ull) continue;
      if (lg[e] > best) { best = lg[e]; best_e = e; }
    }
    // warp arg-max (ties -> lowest expert id)
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ob = __shfl_xor_sync(0xffffffffu, best, o);
      const int oe = __shfl_xor_sync(0xffffffffu, best_e, o);
      if (ob > best || (ob == best && oe >= 0 && (best_e < 0 || oe < best_e))) { best = ob; best_e = oe; }
    }
    if (best_e >= 0 && (best_e & 31) == lane) taken |= 1ull << (best_e / 32);
    my_experts[k] = best_e;
    my_scores[k] = __expf(best - mx) / sum;
    top_sum += my_scores[k];
  }
  if (lane == 0) {
    for (int k = 0; k < K; ++k) {
      const int e = my_experts[k];
      assignments[warp * K + k] = e;
      scores[warp * K + k] = normalize ? my_scores[k] / top_sum : my_scores[k];
      offsets[warp * K + k] = atomicAdd(&expert_counts[e], 1);
    }
  }
}

// expert_cumsum = inclusive cumsum of expert_counts (single small CTA), then rows are copied into expert-sorted order
__global__ void expert_cumsum_kernel(const int* __restrict__ counts, long long* __restrict__ cumsum, int E) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    long long s = 0;
    for (int e = 0; e < E; ++e) { s += counts[e]; cumsum[e] = s; }
  }
}
__global__ void moe_scatter_kernel(const bf16* __restrict__ x, const int* __restrict__ assignments, const int* __restrict__ offsets,
                                   const long long* __restrict__ cumsum, bf16* __restrict__ moe_in, int* __restrict__ mapped_slots,
                                   int dim, int K) {
  const int t = blockIdx.x;
  for (int k = 0; k < K; ++k) {
    const int e = assignments[t * K + k];
    const long long start = e == 0 ? 0 : cumsum[e - 1];
    const long long slot = start + offsets[t * K + k];
    if (threadIdx.x == 0) mapped_slots[t * K + k] = static_cast<int>(slot);
    for (int c = threadIdx.x * 8; c < dim; c += blockDim.x * 8)
      *reinterpret_cast<uint4*>(moe_in + slot * dim + c) = *reinterpret_cast<const uint4*>(x + static_cast<size_t>(t) * dim + c);
  }
}
__global__ void moe_gather_kernel(const bf16* __restrict__ moe_out, const float* __restrict__ scores, const int* __restrict__ mapped_slots,
                                  bf16* __restrict__ out, int dim, int K) {
  const int t = blockIdx.x;
  for (int c = threadIdx.x * 8; c < dim; c += blockDim.x * 8) {
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int k = 0; k < K; ++k) {
      const float w = scores[t * K + k];
      uint4 v = *reinterpret_cast<const uint4*>(moe_out + static_cast<size_t>(mapped_slots[t * K + k]) * dim + c);
      const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        float2 f = __bfloat1622float2(h[i]);
        acc[2 * i] = fmaf(w, f.x, acc[2 * i]);
        acc[2 * i + 1] = fmaf(w, f.y, acc[2 * i + 1]);
      }
    }
    uint4 o;
    o.x = pack_bf16x2(acc[0], acc[1]); o.y = pack_bf16x2(acc[2], acc[3]); o.z = pack_bf16x2(acc[4], acc[5]); o.w = pack_bf16x2(acc[6], acc[7]);
    *reinterpret_cast<uint4*>(out + static_cast<size_t>(t) * dim + c) = o;
  }
}

}  // namespace ragged
}  // namespace dsb

using namespace dsb;
using namespace dsb::ragged;

DSB_EXPORT int dsb200_kv_rotary_append(void* qkv, void* kv_cache, const int* tok_seq, const int* tok_pos, const int* block_table,
                                       const float* inv_freq, int n_tokens, int hq, int hkv, int head_dim, int block_size,
                                       int max_blocks_per_seq, int rot_dim, cudaStream_t stream) {
  if (n_tokens == 0) return DSB_OK;
  DSB_REQUIRE(rot_dim % 2 == 0 && rot_dim <= head_dim, "kv_rotary_append: bad rotary dim %d", rot_dim);
  const long long warps = static_cast<long long>(n_tokens) * (hq + 2 * hkv);
  const int wpb = 8;
  kv_rotary_append_kernel<<<static_cast<unsigned>((warps + wpb - 1) / wpb), wpb * 32, 0, stream>>>(
      (bf16*)qkv, (bf16*)kv_cache, tok_seq, tok_pos, block_table, inv_freq, n_tokens, hq, hkv, head_dim, block_size, max_blocks_per_seq,
      rot_dim);
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

DSB_EXPORT int dsb200_paged_attention(const void* qkv, const void* kv_cache, void* out, const int* tok_seq, const int* tok_pos,
                                      const int* block_table, int n_tokens, int hq, int hkv, int head_dim, int block_size,
                                      int max_blocks_per_seq, float scale, cudaStream_t stream) {
  if (n_tokens == 0) return DSB_OK;
  DSB_REQUIRE(hq % hkv == 0 && hq / hkv <= 8, "paged_attention: q/kv head ratio %d unsupported (<= 8)", hq / hkv);
  dim3 grid(n_tokens, hkv);
#define DSB_PA(DD)                                                                                                         \
  paged_attention_kernel<DD, 8><<<grid, 128, 0, stream>>>((const bf16*)qkv, (const bf16*)kv_cache, (bf16*)out, tok_seq, tok_pos, \
                                                          block_table, hq, hkv, block_size, max_blocks_per_seq, scale)
  if (head_dim == 128) DSB_PA(128);
  else if (head_dim == 64) DSB_PA(64);
  else if (head_dim == 96) DSB_PA(96);
  else if (head_dim == 256) DSB_PA(256);
  else {
    set_last_error("paged_attention: head_dim %d unsupported (64, 96, 128, 256)", head_dim);
    return DSB_ERR_UNSUPPORTED;
  }
#undef DSB_PA
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

DSB_EXPORT int dsb200_ragged_embed(const long long* ids, const void* wte, const void* wpe, const int* tok_pos, void* out,
                                   int n_tokens, int dim, int pos_offset, cudaStream_t stream) {
  if (n_tokens == 0) return DSB_OK;
  DSB_REQUIRE(dim % 8 == 0, "ragged_embed: dim must be a multiple of 8");
  dim3 grid((dim / 8 + 127) / 128, n_tokens);
  ragged_embed_kernel<<<grid, 128, 0, stream>>>(ids, (const bf16*)wte, (const bf16*)wpe, tok_pos, (bf16*)out, n_tokens, dim, pos_offset);
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

DSB_EXPORT int dsb200_logits_gather(const void* hidden, const int* last_tok, void* out, int n_seqs, int dim, cudaStream_t stream) {
  if (n_seqs == 0) return DSB_OK;
  DSB_REQUIRE(dim % 8 == 0, "logits_gather: dim must be a multiple of 8");
  dim3 grid((dim / 8 + 127) / 128, n_seqs);
  logits_gather_kernel<<<grid, 128, 0, stream>>>((const bf16*)hidden, last_tok, (bf16*)out, dim);
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

// expert_counts must be zeroed by the caller.  K <= 8, E <= 2048.
DSB_EXPORT int dsb200_top_k_gating(const float* logits, int* expert_counts, float* scores, int* assignments, int* offsets,
                                   int n_tokens, int n_experts, int top_k, int normalize, cudaStream_t stream) {
  if (n_tokens == 0) return DSB_OK;
  DSB_REQUIRE(top_k >= 1 && top_k <= 8 && n_experts <= 2048, "top_k_gating: k=%d (<=8), experts=%d (<=2048)", top_k, n_experts);
  const int wpb = 4;
  top_k_gating_kernel<<<(n_tokens + wpb - 1) / wpb, wpb * 32, 0, stream>>>(logits, expert_counts, scores, assignments, offsets, n_tokens,
                                                                         n_experts, top_k, normalize);
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

DSB_EXPORT int dsb200_moe_scatter(const void* x, const int* expert_counts, const int* assignments, const int* offsets,
                                  long long* expert_cumsum, void* moe_in, int* mapped_slots, int n_tokens, int n_experts, int dim,
                                  int top_k, cudaStream_t stream) {
  if (n_tokens == 0) return DSB_OK;
  DSB_REQUIRE(dim % 8 == 0, "moe_scatter: dim must be a multiple of 8");
  expert_cumsum_kernel<<<1, 32, 0, stream>>>(expert_counts, expert_cumsum, n_experts);
  moe_scatter_kernel<<<n_tokens, 128, 0, stream>>>((const bf16*)x, assignments, offsets, expert_cumsum, (bf16*)moe_in, mapped_slots, dim,
                                                  top_k);
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

DSB_EXPORT int dsb200_moe_gather(const void* moe_out, const float* scores, const int* mapped_slots, void* out, int n_tokens, int dim,
                                 int top_k, cudaStream_t stream) {
  if (n_tokens == 0) return DSB_OK;
  DSB_REQUIRE(dim % 8 == 0, "moe_gather: dim must be a multiple of 8");
  moe_gather_kernel<<<n_tokens, 128, 0, stream>>>((const bf16*)moe_out, scores, mapped_slots, (bf16*)out, dim, top_k);
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

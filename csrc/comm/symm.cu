// Peer-memory ("symmetric memory") runtime and collective kernels over NVLink-5 / NVSwitch.
//
// Every rank allocates the same-sized buffer with cudaMalloc, exports a CUDA IPC handle, and maps all peers' buffers;
// kernels then address peer memory directly with ld/st.global (NVSwitch gives every pair full bandwidth, so all
// algorithms here are single-hop "pull" or "push" -- no rings, no hierarchy).  Cross-GPU ordering uses a flag pad in
// the same symmetric allocation: writers publish with st.release.sys after their data stores, readers spin (bounded)
// with ld.acquire.sys.  The reference has no counterpart: its only backend is torch.distributed/NCCL (oneCCL + a
// POSIX-shm all-reduce on CPU, csrc/cpu/comm/ccl.cpp:540-611); these kernels take the role of the collectives at
// call sites C1/C3/C7/C8/C11/C16 of SURVEY.md 2.5 and feed the fused paths (RS -> cast/scale -> Adam, AG -> GEMM).
#include "../common/host.h"
#include "../common/ptx.cuh"

namespace dsb {
namespace symm {

constexpr int MAX_RANKS = 16;

struct Peers {
  void* ptr[MAX_RANKS];  // base of every rank's symmetric buffer (own rank included), as mapped in THIS process
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint4 ld_nc_v4(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
// peer data may change between kernels: relaxed system-scope loads (never the non-coherent path)
__device__ __forceinline__ uint4 ld_peer_v4(const void* p) {
  uint4 r;
  asm volatile("ld.relaxed.sys.global.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p)
               : "memory");
  return r;
}

// ---------------------------------------------------------------- device-wide barrier across ranks
// flags: W uint32 slots at byte offset `flag_off` of every rank's symmetric buffer.  Rank r writes `epoch` into slot r of
// every peer, then waits until all W local slots reached `epoch`.  One CTA, W threads.
__global__ void barrier_kernel(Peers peers, size_t flag_off, int rank, int world, uint32_t epoch, int* error_flag) {
  __threadfence_system();
  const int t = threadIdx.x;
  if (t < world) {
    uint32_t* remote = reinterpret_cast<uint32_t*>(static_cast<uint8_t*>(peers.ptr[t]) + flag_off) + rank;
    // max, not store: barriers issued on different streams may execute out of epoch order on a rank, and a late, lower epoch
    // must never lower a flag a peer is already waiting on
    asm volatile("red.release.sys.global.max.u32 [%0], %1;" ::"l"(remote), "r"(epoch) : "memory");
    const uint32_t* local = reinterpret_cast<const uint32_t*>(static_cast<uint8_t*>(peers.ptr[rank]) + flag_off) + t;
    uint32_t spins = 0;
    while (static_cast<int32_t>(ld_acquire_sys(local) - epoch) < 0) {
      if (++spins > DSB_SPIN_LIMIT) {
        if (error_flag) *error_flag = 1;
        printf("dsb200: cross-GPU barrier timed out on rank %d waiting for rank %d (epoch %u)\n", rank, t, epoch);
        __trap();
      }
      __nanosleep(64);
    }
  }
  __syncthreads();
  __threadfence_system();
}

// Graph-capturable form: the epoch is a counter in this rank's LOCAL memory, advanced by the kernel itself (one CTA: every thread
// reads it before the block barrier, thread 0 writes it after), so a captured barrier replays correctly.  Every rank must issue
// these barriers in the same order on one stream (or capture them in one graph); their flag pad is separate from the host-epoch one.
__global__ void barrier_dev_kernel(Peers peers, size_t flag_off, int rank, int world, uint32_t* epoch_ctr) {
  __threadfence_system();
  const int t = threadIdx.x;
  const uint32_t epoch = *epoch_ctr + 1;
  if (t < world) {
    uint32_t* remote = reinterpret_cast<uint32_t*>(static_cast<uint8_t*>(peers.ptr[t]) + flag_off) + rank;
    asm volatile("red.release.sys.global.max.u32 [%0], %1;" ::"l"(remote), "r"(epoch) : "memory");
    const uint32_t* local = reinterpret_cast<const uint32_t*>(static_cast<uint8_t*>(peers.ptr[rank]) + flag_off) + t;
    uint32_t spins = 0;
    while (static_cast<int32_t>(ld_acquire_sys(local) - epoch) < 0) {
      if (++spins > DSB_SPIN_LIMIT) {
        printf("dsb200: cross-GPU barrier (device epoch) timed out on rank %d waiting for rank %d (epoch %u)\n", rank, t, epoch);
        __trap();
      }
      __nanosleep(64);
    }
  }
  __syncthreads();
  if (t == 0) *epoch_ctr = epoch;
  __threadfence_system();
}

// ---------------------------------------------------------------- all-gather (pull): out[p*n .. (p+1)*n) = peer p's shard
// `src_off` = byte offset of the shard inside each rank's symmetric buffer; dst is local memory (any allocation).
__global__ void allgather_pull_kernel(Peers peers, size_t src_off, uint8_t* __restrict__ dst, size_t bytes_per_rank, int rank,
                                      int world) {
  const size_t vecs = bytes_per_rank / 16;
  const size_t total = vecs * world;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride) {
    // start each rank on a different peer so the NVSwitch ports are loaded evenly
    const int pi = static_cast<int>(i / vecs);
    const int peer = (pi + rank) % world;
    const size_t v = i - static_cast<size_t>(pi) * vecs;
    const uint4 val = ld_peer_v4(static_cast<const uint8_t*>(peers.ptr[peer]) + src_off + v * 16);
    *reinterpret_cast<uint4*>(dst + static_cast<size_t>(peer) * bytes_per_rank + v * 16) = val;
  }
}

// ---------------------------------------------------------------- all-to-all (pull), equal splits:
// every rank holds world chunks of `bytes_per_chunk` at `src_off` (chunk j = what it sends to rank j); rank r gathers
// out[p] = peer p's chunk r.  Used for MoE dispatch / combine and the Ulysses head <-> sequence exchanges
// (reference: all_to_all_single call sites in moe/sharded_moe.py:95-111 and sequence/layer.py:15-41, NCCL there).
__global__ void alltoall_pull_kernel(Peers peers, size_t src_off, uint8_t* __restrict__ dst, size_t bytes_per_chunk, int rank,
                                     int world) {
  const size_t vecs = bytes_per_chunk / 16;
  const size_t total = vecs * world;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int pi = static_cast<int>(i / vecs);
    const int peer = (pi + rank) % world;  // stagger the peers so every NVSwitch port is busy
    const size_t v = i - static_cast<size_t>(pi) * vecs;
    const uint4 val = ld_peer_v4(static_cast<const uint8_t*>(peers.ptr[peer]) + src_off + static_cast<size_t>(rank) * bytes_per_chunk + v * 16);
    *reinterpret_cast<uint4*>(dst + static_cast<size_t>(peer) * bytes_per_chunk + v * 16) = val;
  }
}

// ---- NVLS (multimem) primitives; the kernels that use them are further down
__device__ __forceinline__ uint4 multimem_ld_reduce_bf16x8(const void* mc) {
  uint4 r;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(mc)
               : "memory");
  return r;
}
__device__ __forceinline__ float4 multimem_ld_reduce_f32x4(const void* mc) {
  float4 r;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(mc)
               : "memory");
  return r;
}
__device__ __forceinline__ void multimem_st_v4(void* mc, const uint4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "f"(__uint_as_float(v.x)),
               "f"(__uint_as_float(v.y)), "f"(__uint_as_float(v.z)), "f"(__uint_as_float(v.w))
               : "memory");
}

// scale, (accumulate into) store and sum-of-squares of one reduced 16-byte vector: shared by the P2P and NVLS reduce-scatters
template <int EPV, bool DST_BF16>
__device__ __forceinline__ float rs_finish(float (&acc)[EPV], void* __restrict__ dst, size_t v, float scale, int accumulate, bool want_sq) {
#pragma unroll
  for (int k = 0; k < EPV; ++k) acc[k] *= scale;
  if constexpr (DST_BF16) {
    __nv_bfloat16* d = static_cast<__nv_bfloat16*>(dst) + v * EPV;
    if (accumulate) {
#pragma unroll
      for (int k = 0; k < EPV; ++k) acc[k] += __bfloat162float(d[k]);
    }
#pragma unroll
    for (int k = 0; k < EPV; k += 2) *reinterpret_cast<uint32_t*>(d + k) = pack_bf16x2(acc[k], acc[k + 1]);
  } else {
    float* d = static_cast<float*>(dst) + v * EPV;
    if (accumulate) {
#pragma unroll
      for (int k = 0; k < EPV; ++k) acc[k] += d[k];
    }
#pragma unroll
    for (int k = 0; k < EPV; k += 4) *reinterpret_cast<float4*>(d + k) = make_float4(acc[k], acc[k + 1], acc[k + 2], acc[k + 3]);
  }
  float sq = 0.f;
  if (want_sq) {
#pragma unroll
    for (int k = 0; k < EPV; ++k) sq = fmaf(acc[k], acc[k], sq);
  }
  return sq;
}

// one atomicAdd per CTA
__device__ __forceinline__ void block_atomic_sum(float local, float* __restrict__ out) {
  __shared__ float red[32];
  local = warp_sum(local);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  if (l == 0) red[w] = local;
  __syncthreads();
  if (w == 0) {
    float t = l < (blockDim.x >> 5) ? red[l] : 0.f;
    t = warp_sum(t);
    if (l == 0) atomicAdd(out, t);
  }
}

// ---------------------------------------------------------------- reduce-scatter (pull) fused with scale / cast / norm
// Every rank holds a full gradient buffer (bf16 or fp32) of world*n elements at `src_off` of its symmetric buffer.
// Rank r sums slice r over all peers in fp32, multiplies by `scale`, and either stores or accumulates into `dst`
// (fp32 or bf16).  Optionally adds sum(g^2) of the result into `sumsq` (one atomicAdd per CTA) for the gradient norm.
template <bool SRC_BF16, bool DST_BF16>
__global__ void reduce_scatter_pull_kernel(Peers peers, size_t src_off, void* __restrict__ dst, size_t n, int rank, int world,
                                           float scale, int accumulate, float* __restrict__ sumsq) {
  constexpr int EPV = SRC_BF16 ? 8 : 4;  // elements per 16-byte vector
  constexpr int BATCH = 8;               // peer loads issued back to back before the first add (a peer load is a 2-3 us round trip)
  const size_t vecs = n / EPV;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  const size_t esz = SRC_BF16 ? 2 : 4;
  float local_sq = 0.f;
  for (size_t v = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; v < vecs; v += stride) {
    float acc[EPV];
#pragma unroll
    for (int k = 0; k < EPV; ++k) acc[k] = 0.f;
    const size_t byte_off = src_off + (static_cast<size_t>(rank) * n + v * EPV) * esz;
    for (int p0 = 0; p0 < world; p0 += BATCH) {
      uint4 raw[BATCH];
#pragma unroll
      for (int b = 0; b < BATCH; ++b) {
        const int pi = p0 + b;
        if (pi < world) raw[b] = ld_peer_v4(static_cast<const uint8_t*>(peers.ptr[(pi + rank) % world]) + byte_off);
      }
#pragma unroll
      for (int b = 0; b < BATCH; ++b) {
        if (p0 + b >= world) continue;
        if constexpr (SRC_BF16) {
          const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw[b]);
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            float2 f = __bfloat1622float2(h[k]);
            acc[2 * k] += f.x;
            acc[2 * k + 1] += f.y;
          }
        } else {
          acc[0] += __uint_as_float(raw[b].x); acc[1] += __uint_as_float(raw[b].y);
          acc[2] += __uint_as_float(raw[b].z); acc[3] += __uint_as_float(raw[b].w);
        }
      }
    }
    local_sq += rs_finish<EPV, DST_BF16>(acc, dst, v, scale, accumulate, sumsq != nullptr);
  }
  if (sumsq) block_atomic_sum(local_sq, sumsq);
}

// ---------------------------------------------------------------- qgZ: reduce-scatter of block-scaled fp8 gradients (pull)
// Every rank has quantised its W*n-element gradient buffer into e4m3 payload (1 byte/element at q_off) + one fp32 scale per
// 128 elements (at s_off) inside its symmetric buffer (csrc/quant/quantizer.cu fp8_quant_kernel).  Rank r pulls slice r of
// payload + scales from every peer -- half the NVLink bytes of the bf16 reduce-scatter --, dequantises and sums in fp32,
// then scale / cast / accumulate / sum-of-squares as above.  16 payload bytes per thread per peer, W loads in flight.
__device__ __forceinline__ float ld_peer_f32(const void* p) {
  float r;
  asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(r) : "l"(p) : "memory");
  return r;
}
__device__ __forceinline__ void fp8x4_to_float(uint32_t w, float s, float* acc) {
  // e4m3 x4 -> half2 x2 (cvt.rn.f16x2.e4m3x2), then fp32 multiply-add with the block scale
  uint32_t lo, hi;
  asm("{\n\t.reg .b16 a, b;\n\tmov.b32 {a, b}, %2;\n\tcvt.rn.f16x2.e4m3x2 %0, a;\n\tcvt.rn.f16x2.e4m3x2 %1, b;\n\t}"
      : "=r"(lo), "=r"(hi)
      : "r"(w));
  const __half2 h0 = *reinterpret_cast<const __half2*>(&lo), h1 = *reinterpret_cast<const __half2*>(&hi);
  const float2 f0 = __half22float2(h0), f1 = __half22float2(h1);
  acc[0] = fmaf(f0.x, s, acc[0]); acc[1] = fmaf(f0.y, s, acc[1]);
  acc[2] = fmaf(f1.x, s, acc[2]); acc[3] = fmaf(f1.y, s, acc[3]);
}
template <bool DST_BF16>
__global__ void fp8_reduce_scatter_pull_kernel(Peers peers, size_t q_off, size_t s_off, void* __restrict__ dst, size_t n, int rank,
                                               int world, float scale, int accumulate, float* __restrict__ sumsq) {
  constexpr int BATCH = 8;
  const size_t vecs = n / 16;  // 16 fp8 elements per 16-byte vector
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  float local_sq = 0.f;
  for (size_t v = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; v < vecs; v += stride) {
    const size_t elem = static_cast<size_t>(rank) * n + v * 16;
    const size_t qb = q_off + elem, sb = s_off + (elem >> 7) * 4;
    float acc[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) acc[k] = 0.f;
    for (int p0 = 0; p0 < world; p0 += BATCH) {
      uint4 raw[BATCH];
      float sc[BATCH];
#pragma unroll
      for (int b = 0; b < BATCH; ++b) {
        if (p0 + b < world) {
          const uint8_t* base = static_cast<const uint8_t*>(peers.ptr[(p0 + b + rank) % world]);
          raw[b] = ld_peer_v4(base + qb);
          sc[b] = ld_peer_f32(base + sb);
        }
      }
#pragma unroll
      for (int b = 0; b < BATCH; ++b) {
        if (p0 + b >= world) continue;
        fp8x4_to_float(raw[b].x, sc[b], acc);
        fp8x4_to_float(raw[b].y, sc[b], acc + 4);
        fp8x4_to_float(raw[b].z, sc[b], acc + 8);
        fp8x4_to_float(raw[b].w, sc[b], acc + 12);
      }
    }
    float lo8[8], hi8[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) { lo8[k] = acc[k]; hi8[k] = acc[8 + k]; }
    if constexpr (DST_BF16) {
      local_sq += rs_finish<8, true>(lo8, dst, v * 2, scale, accumulate, sumsq != nullptr);
      local_sq += rs_finish<8, true>(hi8, dst, v * 2 + 1, scale, accumulate, sumsq != nullptr);
    } else {
      float q4[4][4];
#pragma unroll
      for (int k = 0; k < 16; ++k) q4[k >> 2][k & 3] = acc[k];
#pragma unroll
      for (int j = 0; j < 4; ++j) local_sq += rs_finish<4, false>(q4[j], dst, v * 4 + j, scale, accumulate, sumsq != nullptr);
    }
  }
  if (sumsq) block_atomic_sum(local_sq, sumsq);
}

// ---------------------------------------------------------------- reduce-scatter fused with the partitioned AdamW update
// One pass: g = scale * sum_peers grad[peer][rank*n + i]; Adam on the fp32 master shard; bf16 parameter shard written.
// Valid when no global-norm clipping is needed before the update (gradient_clipping == 0) -- otherwise use the
// reduce-scatter above (which produces the norm) followed by the fused Adam kernel.
struct RsAdamArgs {
  float lr, beta1, beta2, eps, weight_decay, bc1, bc2_rsqrt, scale;
  int adamw;
};
// `peers` (pull: every rank's buffer) or `mc` (NVLS: the switch adds) delivers the gradients of elements
// [elem_lo, elem_lo + n) of this rank's slice, whose first element sits at byte `slice_off` of the symmetric buffers.
// p / m / v / lp / g_out point at element elem_lo of the rank-local shard.  g_out (bf16, optional) keeps the reduced,
// scaled gradient for inspection; sumsq (optional) accumulates its squared norm.
template <bool NVLS>
__global__ void rs_adam_kernel(Peers peers, const uint8_t* __restrict__ mc, size_t slice_off, float* __restrict__ p,
                               float* __restrict__ m, float* __restrict__ v, __nv_bfloat16* __restrict__ lp,
                               __nv_bfloat16* __restrict__ g_out, float* __restrict__ sumsq, size_t n, int rank, int world,
                               RsAdamArgs a) {
  const size_t vecs = n / 8;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  float local_sq = 0.f;
  for (size_t vi = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; vi < vecs; vi += stride) {
    float g[8];
    const size_t byte_off = slice_off + vi * 16;
    if constexpr (NVLS) {
      const uint4 raw = multimem_ld_reduce_bf16x8(mc + byte_off);
      const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        float2 f = __bfloat1622float2(h[k]);
        g[2 * k] = f.x;
        g[2 * k + 1] = f.y;
      }
    } else {
#pragma unroll
      for (int k = 0; k < 8; ++k) g[k] = 0.f;
      constexpr int BATCH = 8;
      for (int p0 = 0; p0 < world; p0 += BATCH) {
        uint4 raw[BATCH];
#pragma unroll
        for (int b = 0; b < BATCH; ++b)
          if (p0 + b < world) raw[b] = ld_peer_v4(static_cast<const uint8_t*>(peers.ptr[(p0 + b + rank) % world]) + byte_off);
#pragma unroll
        for (int b = 0; b < BATCH; ++b) {
          if (p0 + b >= world) continue;
          const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw[b]);
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            float2 f = __bfloat1622float2(h[k]);
            g[2 * k] += f.x;
            g[2 * k + 1] += f.y;
          }
        }
      }
    }
    const size_t e = vi * 8;
    float pp[8], mm[8], vv[8];
    *reinterpret_cast<float4*>(pp) = *reinterpret_cast<const float4*>(p + e);
    *reinterpret_cast<float4*>(pp + 4) = *reinterpret_cast<const float4*>(p + e + 4);
    *reinterpret_cast<float4*>(mm) = *reinterpret_cast<const float4*>(m + e);
    *reinterpret_cast<float4*>(mm + 4) = *reinterpret_cast<const float4*>(m + e + 4);
    *reinterpret_cast<float4*>(vv) = *reinterpret_cast<const float4*>(v + e);
    *reinterpret_cast<float4*>(vv + 4) = *reinterpret_cast<const float4*>(v + e + 4);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      float gg = g[k] * a.scale;
      g[k] = gg;
      local_sq = fmaf(gg, gg, local_sq);
      if (!a.adamw) gg = fmaf(a.weight_decay, pp[k], gg);
      mm[k] = fmaf(a.beta1, mm[k], (1.f - a.beta1) * gg);
      vv[k] = fmaf(a.beta2, vv[k], (1.f - a.beta2) * gg * gg);
      float upd = (mm[k] * a.bc1) / (sqrtf(vv[k]) * a.bc2_rsqrt + a.eps);
      if (a.adamw) upd = fmaf(a.weight_decay, pp[k], upd);
      pp[k] = fmaf(-a.lr, upd, pp[k]);
    }
    *reinterpret_cast<float4*>(p + e) = *reinterpret_cast<float4*>(pp);
    *reinterpret_cast<float4*>(p + e + 4) = *reinterpret_cast<float4*>(pp + 4);
    *reinterpret_cast<float4*>(m + e) = *reinterpret_cast<float4*>(mm);
    *reinterpret_cast<float4*>(m + e + 4) = *reinterpret_cast<float4*>(mm + 4);
    *reinterpret_cast<float4*>(v + e) = *reinterpret_cast<float4*>(vv);
    *reinterpret_cast<float4*>(v + e + 4) = *reinterpret_cast<float4*>(vv + 4);
    uint4 o;
    o.x = pack_bf16x2(pp[0], pp[1]); o.y = pack_bf16x2(pp[2], pp[3]); o.z = pack_bf16x2(pp[4], pp[5]); o.w = pack_bf16x2(pp[6], pp[7]);
    *reinterpret_cast<uint4*>(lp + e) = o;
    if (g_out) {
      o.x = pack_bf16x2(g[0], g[1]); o.y = pack_bf16x2(g[2], g[3]); o.z = pack_bf16x2(g[4], g[5]); o.w = pack_bf16x2(g[6], g[7]);
      *reinterpret_cast<uint4*>(g_out + e) = o;
    }
  }
  if (sumsq) block_atomic_sum(local_sq, sumsq);
}

// ---------------------------------------------------------------- one-shot all-reduce (small messages: TP inference)
// in place on a region of the symmetric buffer: every rank reads all peers' copies, sums in fp32, writes `dst` (local,
// may be ordinary memory).  Optional fused residual add + RMSNorm is done by the caller's next kernel.
__global__ void allreduce_oneshot_kernel(Peers peers, size_t src_off, __nv_bfloat16* __restrict__ dst, size_t n, int rank,
                                         int world) {
  const size_t vecs = n / 8;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  for (size_t vi = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; vi < vecs; vi += stride) {
    float acc[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k] = 0.f;
    for (int pi = 0; pi < world; ++pi) {
      const int peer = (pi + rank) % world;
      const uint4 raw = ld_peer_v4(static_cast<const uint8_t*>(peers.ptr[peer]) + src_off + vi * 16);
      const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        float2 f = __bfloat1622float2(h[k]);
        acc[2 * k] += f.x;
        acc[2 * k + 1] += f.y;
      }
    }
    uint4 o;
    o.x = pack_bf16x2(acc[0], acc[1]); o.y = pack_bf16x2(acc[2], acc[3]); o.z = pack_bf16x2(acc[4], acc[5]); o.w = pack_bf16x2(acc[6], acc[7]);
    *reinterpret_cast<uint4*>(dst + vi * 8) = o;
  }
}

// ---------------------------------------------------------------- MoE expert parallelism: dispatch push / combine pull
// Capacity layout of the reference's MoE (moe/sharded_moe.py:477-554): token s routed to expert e lands in slot c of that
// expert's [C, M] buffer; expert e lives on rank e / n_local.  The receive buffer of a rank is [W src ranks][n_local][C][M]
// inside its symmetric buffer.  Dispatch = every (token, k) row is STORED straight into the owner's receive buffer at
// [my rank][e % n_local][slot] (replaces scatter + all-to-all); combine = every token PULLS its k expert-output rows from the
// owners' buffers and sums them with the gate weights (replaces all-to-all + gather).  One warp per (token, k) / per token.
__device__ __forceinline__ void st_peer_v4(void* p, const uint4& v) {
  asm volatile("st.relaxed.sys.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint8_t* moe_row_ptr(const Peers& peers, size_t buf_off, int rank, int e, int slot, int n_local,
                                                int capacity, int M) {
  const int owner = e / n_local, le = e - owner * n_local;
  const size_t row = (static_cast<size_t>(rank) * n_local + le) * capacity + slot;
  return static_cast<uint8_t*>(peers.ptr[owner]) + buf_off + row * static_cast<size_t>(M) * 2;
}
// x [S][M] bf16; token_expert / token_slot [S][K] (expert < 0 = dropped); row_scale optional [S][K] fp32 (backward of combine:
// rows are scaled by the gate weight on the way out)
__global__ void moe_dispatch_push_kernel(Peers peers, size_t buf_off, const __nv_bfloat16* __restrict__ x, const int* __restrict__ token_expert,
                                         const int* __restrict__ token_slot, const float* __restrict__ row_scale, int S, int K, int M,
                                         int rank, int n_local, int capacity) {
  const int wid = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (wid >= S * K) return;
  const int s = wid / K;
  const int e = token_expert[wid];
  if (e < 0) return;
  uint8_t* dst = moe_row_ptr(peers, buf_off, rank, e, token_slot[wid], n_local, capacity, M);
  const uint4* src = reinterpret_cast<const uint4*>(x + static_cast<size_t>(s) * M);
  const float sc = row_scale ? row_scale[wid] : 1.f;
  for (int v = lane; v < M / 8; v += 32) {
    uint4 val = src[v];
    if (row_scale) {
      __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&val);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        float2 f = __bfloat1622float2(h[i]);
        h[i] = __floats2bfloat162_rn(f.x * sc, f.y * sc);
      }
    }
    st_peer_v4(dst + v * 16, val);
  }
}
// out[s] = sum_k w[s][k] * row(e_k, slot_k) pulled from the owner; weights NULL = plain sum (backward of dispatch).
// dot_out optional [S][K]: <pulled row, dot_with[s]> (gate-weight gradient in the backward of combine; out may be NULL then)
__global__ void moe_combine_pull_kernel(Peers peers, size_t buf_off, __nv_bfloat16* __restrict__ out, const int* __restrict__ token_expert,
                                        const int* __restrict__ token_slot, const float* __restrict__ weights,
                                        const __nv_bfloat16* __restrict__ dot_with, float* __restrict__ dot_out, int S, int K, int M, int rank,
                                        int n_local, int capacity) {
  const int s = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (s >= S) return;
  for (int v0 = lane; v0 < M / 8 || (dot_out && v0 < ((M / 8 + 31) / 32) * 32); v0 += 32) {
    const bool in = v0 < M / 8;
    float acc[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = 0.f;
    uint4 dw = make_uint4(0, 0, 0, 0);
    if (dot_out && in) dw = *reinterpret_cast<const uint4*>(dot_with + static_cast<size_t>(s) * M + v0 * 8);
    for (int k = 0; k < K; ++k) {
      const int e = token_expert[s * K + k];
      float part = 0.f;
      if (e >= 0 && in) {
        const uint8_t* src = moe_row_ptr(peers, buf_off, rank, e, token_slot[s * K + k], n_local, capacity, M);
        const uint4 raw = ld_peer_v4(src + v0 * 16);
        const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw);
        const __nv_bfloat162* d = reinterpret_cast<const __nv_bfloat162*>(&dw);
        const float w = weights ? weights[s * K + k] : 1.f;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float2 f = __bfloat1622float2(h[i]);
          acc[2 * i] = fmaf(w, f.x, acc[2 * i]);
          acc[2 * i + 1] = fmaf(w, f.y, acc[2 * i + 1]);
          if (dot_out) {
            const float2 g = __bfloat1622float2(d[i]);
            part = fmaf(f.x, g.x, fmaf(f.y, g.y, part));
          }
        }
      }
      if (dot_out) {
        part = warp_sum(part);
        if (lane == 0 && e >= 0) dot_out[s * K + k] += part;  // one warp owns token s: plain accumulate over the v0 passes
      }
    }
    if (out && in) {
      uint4 o;
      o.x = pack_bf16x2(acc[0], acc[1]); o.y = pack_bf16x2(acc[2], acc[3]); o.z = pack_bf16x2(acc[4], acc[5]); o.w = pack_bf16x2(acc[6], acc[7]);
      *reinterpret_cast<uint4*>(out + static_cast<size_t>(s) * M + v0 * 8) = o;
    }
  }
}

// ---------------------------------------------------------------- Ulysses: sequence <-> head exchange fused with its layout change
// "Source layout"  (sequence-sharded): rows = B x S_local tokens of this rank, every row holds ALL heads of up to three
//                   segments (q | k | v of a packed QKV projection, or the attention output alone).
// "Dest layout"    (head-sharded): rows = B x S tokens of the WHOLE sequence (S = P * S_local, rank-major), every row holds
//                   this rank's heads of each segment, packed.
// push: local source rows are STORED straight into the owners' dest-layout buffers (the all-to-all before attention, with
//       its reshapes and the q/k/v merge folded in); pull: local source rows are LOADED from the owners' dest-layout buffers
//       (the all-to-all after attention).  Each is the other's backward.  One 16-byte vector per thread per step.
struct HeadSeg {
  int src_col, dst_col, heads, heads_per_rank;  // columns in elements; heads of the segment in total / per rank
};
struct HeadSegs {
  HeadSeg s[3];
  int n;
  int vec_per_row;  // sum over segments of heads * D / 8
};
template <bool PUSH>
__global__ void ulysses_heads_kernel(Peers peers, size_t buf_off, __nv_bfloat16* __restrict__ local, int B, int S_local, int P, int rank,
                                     int D, int src_ld, int dst_ld, HeadSegs segs) {
  const long long total = static_cast<long long>(B) * S_local * segs.vec_per_row;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  const int dv = D / 8;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += stride) {
    const long long t = i / segs.vec_per_row;
    int v = static_cast<int>(i - t * segs.vec_per_row);
    int si = 0;
    while (si + 1 < segs.n && v >= segs.s[si].heads * dv) {
      v -= segs.s[si].heads * dv;
      ++si;
    }
    const HeadSeg sg = segs.s[si];
    const int head = v / dv, d8 = v - head * dv;
    const int owner = head / sg.heads_per_rank, lh = head - owner * sg.heads_per_rank;
    const int b = static_cast<int>(t / S_local), sl = static_cast<int>(t - static_cast<long long>(b) * S_local);
    const long long dst_row = static_cast<long long>(b) * S_local * P + static_cast<long long>(rank) * S_local + sl;
    __nv_bfloat16* lp = local + t * src_ld + sg.src_col + head * D + d8 * 8;
    uint8_t* rp = static_cast<uint8_t*>(peers.ptr[owner]) + buf_off + (dst_row * dst_ld + sg.dst_col + lh * D + d8 * 8) * 2;
    if constexpr (PUSH) st_peer_v4(rp, *reinterpret_cast<const uint4*>(lp));
    else *reinterpret_cast<uint4*>(lp) = ld_peer_v4(rp);
  }
}

// ---------------------------------------------------------------- NVLS (NVSwitch multicast) collectives
// `mc` is the multicast mapping of the SAME symmetric buffer: a multimem.ld_reduce at offset x returns the sum over every
// rank's bytes at x, added inside the switch (each GPU receives 1x the slice instead of reading W-1 peers); a multimem.st
// writes the value into every rank's copy.  Callers order these against the producers with the device barrier above.
// reduce-scatter: rank r reduces slice r of the W*n-element buffer at `src_off` in the switch, then scale / cast /
// accumulate / sum-of-squares exactly like the pull kernel.  U loads in flight per thread.
template <bool SRC_BF16, bool DST_BF16>
__global__ void nvls_reduce_scatter_kernel(const uint8_t* __restrict__ mc, size_t src_off, void* __restrict__ dst, size_t n, int rank,
                                           float scale, int accumulate, float* __restrict__ sumsq) {
  constexpr int EPV = SRC_BF16 ? 8 : 4;
  constexpr int U = 4;
  const size_t vecs = n / EPV;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  const uint8_t* base = mc + src_off + static_cast<size_t>(rank) * n * (SRC_BF16 ? 2 : 4);
  float local_sq = 0.f;
  for (size_t v0 = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; v0 < vecs; v0 += stride * U) {
    uint4 raw[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t v = v0 + u * stride;
      if (v < vecs) {
        if constexpr (SRC_BF16) {
          raw[u] = multimem_ld_reduce_bf16x8(base + v * 16);
        } else {
          const float4 f = multimem_ld_reduce_f32x4(base + v * 16);
          raw[u] = make_uint4(__float_as_uint(f.x), __float_as_uint(f.y), __float_as_uint(f.z), __float_as_uint(f.w));
        }
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t v = v0 + u * stride;
      if (v >= vecs) continue;
      float acc[EPV];
      if constexpr (SRC_BF16) {
        const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw[u]);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          float2 f = __bfloat1622float2(h[k]);
          acc[2 * k] = f.x;
          acc[2 * k + 1] = f.y;
        }
      } else {
        acc[0] = __uint_as_float(raw[u].x); acc[1] = __uint_as_float(raw[u].y);
        acc[2] = __uint_as_float(raw[u].z); acc[3] = __uint_as_float(raw[u].w);
      }
      local_sq += rs_finish<EPV, DST_BF16>(acc, dst, v, scale, accumulate, sumsq != nullptr);
    }
  }
  if (sumsq) block_atomic_sum(local_sq, sumsq);
}

// all-reduce, in place in the symmetric buffer: rank r reduces vectors [r*per, (r+1)*per) in the switch and multicasts the
// sums back into every rank's copy (one NVLink crossing each way per element).  bf16.
__global__ void nvls_allreduce_kernel(uint8_t* __restrict__ mc, size_t off, size_t vecs_per_rank, int rank) {
  constexpr int U = 4;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  uint8_t* base = mc + off + static_cast<size_t>(rank) * vecs_per_rank * 16;
  for (size_t v0 = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; v0 < vecs_per_rank; v0 += stride * U) {
    uint4 raw[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t v = v0 + u * stride;
      if (v < vecs_per_rank) raw[u] = multimem_ld_reduce_bf16x8(base + v * 16);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t v = v0 + u * stride;
      if (v < vecs_per_rank) multimem_st_v4(base + v * 16, raw[u]);
    }
  }
}

// all-gather (push): this rank's `bytes` at local `src` are multicast into every rank's buffer at dst_off + rank*bytes.
__global__ void nvls_allgather_kernel(uint8_t* __restrict__ mc, size_t dst_off, const uint8_t* __restrict__ src, size_t bytes, int rank) {
  const size_t vecs = bytes / 16;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  uint8_t* base = mc + dst_off + static_cast<size_t>(rank) * bytes;
  for (size_t v = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; v < vecs; v += stride)
    multimem_st_v4(base + v * 16, ld_nc_v4(src + v * 16));
}

static int comm_grid(size_t vecs, int max_ctas) {
  size_t b = (vecs + 255) / 256;
  size_t cap = max_ctas > 0 ? static_cast<size_t>(max_ctas) : static_cast<size_t>(num_sms()) * 4;
  if (b < 1) b = 1;
  return static_cast<int>(b < cap ? b : cap);
}

static int make_peers(Peers* out, void* const* peer_ptrs, int world) {
  if (world < 1 || world > MAX_RANKS) {
    set_last_error("symm: world size %d unsupported (max %d)", world, MAX_RANKS);
    return DSB_ERR_ARG;
  }
  for (int i = 0; i < MAX_RANKS; ++i) out->ptr[i] = i < world ? peer_ptrs[i] : nullptr;
  return DSB_OK;
}

}  // namespace symm
}  // namespace dsb

using namespace dsb;
using namespace dsb::symm;

// ---- allocation / mapping (host pointers unless noted)
DSB_EXPORT int dsb200_symm_alloc(long long bytes, void** ptr, unsigned char* handle64) {
  DSB_CHECK_CUDA(cudaMalloc(ptr, static_cast<size_t>(bytes)));
  DSB_CHECK_CUDA(cudaMemset(*ptr, 0, static_cast<size_t>(bytes)));
  cudaIpcMemHandle_t h;
  DSB_CHECK_CUDA(cudaIpcGetMemHandle(&h, *ptr));
  static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
  memcpy(handle64, &h, 64);
  return DSB_OK;
}
DSB_EXPORT int dsb200_symm_open(const unsigned char* handle64, void** ptr) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  DSB_CHECK_CUDA(cudaIpcOpenMemHandle(ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return DSB_OK;
}
DSB_EXPORT int dsb200_symm_close(void* ptr) {
  DSB_CHECK_CUDA(cudaIpcCloseMemHandle(ptr));
  return DSB_OK;
}
DSB_EXPORT int dsb200_symm_free(void* ptr) {
  DSB_CHECK_CUDA(cudaFree(ptr));
  return DSB_OK;
}

// peer_ptrs: HOST array of `world` device pointers (this process's mappings of every rank's buffer, own included)
DSB_EXPORT int dsb200_symm_barrier(void* const* peer_ptrs, long long flag_off, int rank, int world, unsigned int epoch,
                                   cudaStream_t stream) {
  Peers P;
  int rc = make_peers(&P, peer_ptrs, world);
  if (rc) return rc;
  barrier_kernel<<<1, 32, 0, stream>>>(P, static_cast<size_t>(flag_off), rank, world, epoch, nullptr);
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

// flag_off: a pad of `world` uint32 used ONLY by device-epoch barriers (zero-initialised); epoch_ctr: local device uint32
DSB_EXPORT int dsb200_symm_barrier_dev(void* const* peer_ptrs, long long flag_off, int rank, int world, unsigned int* epoch_ctr,
                                       cudaStream_t stream) {
  Peers P;
  int rc = make_peers(&P, peer_ptrs, world);
  if (rc) return rc;
  DSB_REQUIRE(epoch_ctr != nullptr, "symm_barrier_dev: epoch counter missing");
  barrier_dev_kernel<<<1, 32, 0, stream>>>(P, static_cast<size_t>(flag_off), rank, world, epoch_ctr);
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

DSB_EXPORT int dsb200_p2p_allgather(void* const* peer_ptrs, long long src_off, void* dst, long long bytes_per_rank, int rank,
                                    int world, int max_ctas, cudaStream_t stream) {
  DSB_REQUIRE(bytes_per_rank % 16 == 0 && src_off % 16 == 0, "p2p_allgather: sizes/offsets must be multiples of 16 bytes");
  Peers P;
  int rc = make_peers(&P, peer_ptrs, world);
  if (rc) return rc;
  const size_t vecs = static_cast<size_t>(bytes_per_rank) / 16 * world;
  allgather_pull_kernel<<<comm_grid(vecs, max_ctas), 256, 0, stream>>>(P, static_cast<size_t>(src_off), static_cast<uint8_t*>(dst),
                                                                         static_cast<size_t>(bytes_per_rank), rank, world);
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

// dst[p * bytes_per_chunk ...] = chunk `rank` of peer p's buffer at src_off (equal-split all_to_all_single); dst is local memory.
DSB_EXPORT int dsb200_p2p_alltoall(void* const* peer_ptrs, long long src_off, void* dst, long long bytes_per_chunk, int rank, int world,
                                   int max_ctas, cudaStream_t stream) {
  DSB_REQUIRE(bytes_per_chunk % 16 == 0 && src_off % 16 == 0, "p2p_alltoall: sizes/offsets must be multiples of 16 bytes");
  Peers P;
  int rc = make_peers(&P, peer_ptrs, world);
  if (rc) return rc;
  const size_t vecs = static_cast<size_t>(bytes_per_chunk) / 16 * world;
  alltoall_pull_kernel<<<comm_grid(vecs, max_ctas), 256, 0, stream>>>(P, static_cast<size_t>(src_off), static_cast<uint8_t*>(dst),
                                                                        static_cast<size_t>(bytes_per_chunk), rank, world);
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

// src dtype / dst dtype: 0 bf16, 2 fp32.  n = elements per rank (multiple of 8).  sumsq: optional device float (+= sum g^2).
DSB_EXPORT int dsb200_p2p_reduce_scatter(void* const* peer_ptrs, long long src_off, int src_dtype, void* dst, int dst_dtype,
                                         long long n, int rank, int world, float scale, int accumulate, float* sumsq,
                                         int max_ctas, cudaStream_t stream) {
  DSB_REQUIRE(n % 8 == 0 && src_off % 16 == 0, "p2p_reduce_scatter: n must be a multiple of 8 and the offset 16-byte aligned");
  Peers P;
  int rc = make_peers(&P, peer_ptrs, world);
  if (rc) return rc;
  const size_t vecs = static_cast<size_t>(n) / (src_dtype == 0 ? 8 : 4);
  const int grid = comm_grid(vecs, max_ctas);
  const size_t off = static_cast<size_t>(src_off);
  if (src_dtype == 0 && dst_dtype == 0)
    reduce_scatter_pull_kernel<true, true><<<grid, 256, 0, stream>>>(P, off, dst, (size_t)n, rank, world, scale, accumulate, sumsq);
  else if (src_dtype == 0 && dst_dtype == 2)
    reduce_scatter_pull_kernel<true, false><<<grid, 256, 0, stream>>>(P, off, dst, (size_t)n, rank, world, scale, accumulate, sumsq);
  else if (src_dtype == 2 && dst_dtype == 2)
    reduce_scatter_pull_kernel<false, false><<<grid, 256, 0, stream>>>(P, off, dst, (size_t)n, rank, world, scale, accumulate, sumsq);
  else {
    set_last_error("p2p_reduce_scatter: unsupported dtype combination %d -> %d", src_dtype, dst_dtype);
    return DSB_ERR_UNSUPPORTED;
  }
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

// Range form: Adam on elements [elem_lo, elem_lo + n) of this rank's slice (slice length n_per_rank) of the W*n_per_rank
// bf16 gradient buffer at src_off.  p/m/v/lp/g_out point at element elem_lo of the local shard.  mc_base != NULL: the sum
// comes from the NVSwitch (multimem.ld_reduce); NULL: pulled from the peers.  g_out / sumsq optional (may be NULL).
DSB_EXPORT int dsb200_p2p_rs_adam_range(void* const* peer_ptrs, void* mc_base, long long src_off, long long n_per_rank,
                                        long long elem_lo, long long n, float* p, float* m, float* v, void* lp, void* g_out,
                                        float* sumsq, int rank, int world, float scale, float lr, float beta1, float beta2,
                                        float eps, float weight_decay, int adamw_mode, int bias_correction, int step,
                                        int max_ctas, cudaStream_t stream) {
  DSB_REQUIRE(n % 8 == 0 && elem_lo % 8 == 0 && n_per_rank % 8 == 0 && src_off % 16 == 0,
              "p2p_rs_adam: element counts must be multiples of 8 and the offset 16-byte aligned");
  DSB_REQUIRE(elem_lo >= 0 && elem_lo + n <= n_per_rank, "p2p_rs_adam: range outside the rank's slice");
  Peers P;
  int rc = make_peers(&P, peer_ptrs, world);
  if (rc) return rc;
  RsAdamArgs a;
  a.lr = lr; a.beta1 = beta1; a.beta2 = beta2; a.eps = eps; a.weight_decay = weight_decay; a.adamw = adamw_mode; a.scale = scale;
  a.bc1 = bias_correction ? 1.f / (1.f - powf(beta1, (float)step)) : 1.f;
  a.bc2_rsqrt = bias_correction ? 1.f / sqrtf(1.f - powf(beta2, (float)step)) : 1.f;
  const size_t slice_off = static_cast<size_t>(src_off) + (static_cast<size_t>(rank) * n_per_rank + elem_lo) * 2;
  const int grid = comm_grid(static_cast<size_t>(n) / 8, max_ctas);
  if (mc_base)
    rs_adam_kernel<true><<<grid, 256, 0, stream>>>(P, static_cast<const uint8_t*>(mc_base), slice_off, p, m, v,
                                                    static_cast<__nv_bfloat16*>(lp), static_cast<__nv_bfloat16*>(g_out), sumsq,
                                                    static_cast<size_t>(n), rank, world, a);
  else
    rs_adam_kernel<false><<<grid, 256, 0, stream>>>(P, nullptr, slice_off, p, m, v, static_cast<__nv_bfloat16*>(lp),
                                                     static_cast<__nv_bfloat16*>(g_out), sumsq, static_cast<size_t>(n), rank, world, a);
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

DSB_EXPORT int dsb200_p2p_rs_adam(void* const* peer_ptrs, long long src_off, float* p, float* m, float* v, void* lp, long long n,
                                  int rank, int world, float scale, float lr, float beta1, float beta2, float eps,
                                  float weight_decay, int adamw_mode, int bias_correction, int step, int max_ctas,
                                  cudaStream_t stream) {
  return dsb200_p2p_rs_adam_range(peer_ptrs, nullptr, src_off, n, 0, n, p, m, v, lp, nullptr, nullptr, rank, world, scale, lr, beta1,
                                  beta2, eps, weight_decay, adamw_mode, bias_correction, step, max_ctas, stream);
}

DSB_EXPORT int dsb200_p2p_allreduce_bf16(void* const* peer_ptrs, long long src_off, void* dst, long long n, int rank, int world,
                                         int max_ctas, cudaStream_t stream) {
  DSB_REQUIRE(n % 8 == 0 && src_off % 16 == 0, "p2p_allreduce: n must be a multiple of 8 and the offset 16-byte aligned");
  Peers P;
  int rc = make_peers(&P, peer_ptrs, world);
  if (rc) return rc;
  allreduce_oneshot_kernel<<<comm_grid(static_cast<size_t>(n) / 8, max_ctas), 256, 0, stream>>>(
      P, static_cast<size_t>(src_off), static_cast<__nv_bfloat16*>(dst), static_cast<size_t>(n), rank, world);
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

// All-gather on the copy engines: one peer -> local cudaMemcpyAsync per rank (NVLink DMA, zero SMs), staggered so that at any
// moment the ranks read from different peers.  dst is local memory; the peers' shards sit at src_off of their symmetric buffers.
DSB_EXPORT int dsb200_p2p_allgather_ce(void* const* peer_ptrs, long long src_off, void* dst, long long bytes_per_rank, int rank,
                                       int world, cudaStream_t stream) {
  DSB_REQUIRE(world >= 1 && world <= MAX_RANKS, "p2p_allgather_ce: world size %d unsupported", world);
  for (int pi = 0; pi < world; ++pi) {
    const int peer = (pi + rank + 1) % world;  // own shard last: it is the only copy that does not cross NVLink
    DSB_CHECK_CUDA(cudaMemcpyAsync(static_cast<uint8_t*>(dst) + static_cast<size_t>(peer) * bytes_per_rank,
                                   static_cast<const uint8_t*>(peer_ptrs[peer]) + src_off, static_cast<size_t>(bytes_per_rank),
                                   cudaMemcpyDeviceToDevice, stream));
  }
  return DSB_OK;
}

// ---- NVLS entry points.  `mc_base` = multicast mapping of the symmetric buffer (every rank bound), DEVICE pointer.
// src dtype / dst dtype: 0 bf16, 2 fp32.  n = elements per rank (multiple of 8).
DSB_EXPORT int dsb200_nvls_reduce_scatter(void* mc_base, long long src_off, int src_dtype, void* dst, int dst_dtype, long long n,
                                          int rank, int world, float scale, int accumulate, float* sumsq, int max_ctas,
                                          cudaStream_t stream) {
  DSB_REQUIRE(mc_base != nullptr, "nvls_reduce_scatter: no multicast mapping");
  DSB_REQUIRE(n % 8 == 0 && src_off % 16 == 0 && (n * (src_dtype == 0 ? 2 : 4)) % 16 == 0,
              "nvls_reduce_scatter: n must be a multiple of 8 and the offset 16-byte aligned");
  (void)world;
  const size_t vecs = static_cast<size_t>(n) / (src_dtype == 0 ? 8 : 4);
  const int grid = comm_grid((vecs + 3) / 4, max_ctas > 0 ? max_ctas : 64);
  const uint8_t* mc = static_cast<const uint8_t*>(mc_base);
  const size_t off = static_cast<size_t>(src_off);
  if (src_dtype == 0 && dst_dtype == 0)
    nvls_reduce_scatter_kernel<true, true><<<grid, 256, 0, stream>>>(mc, off, dst, (size_t)n, rank, scale, accumulate, sumsq);
  else if (src_dtype == 0 && dst_dtype == 2)
    nvls_reduce_scatter_kernel<true, false><<<grid, 256, 0, stream>>>(mc, off, dst, (size_t)n, rank, scale, accumulate, sumsq);
  else if (src_dtype == 2 && dst_dtype == 2)
    nvls_reduce_scatter_kernel<false, false><<<grid, 256, 0, stream>>>(mc, off, dst, (size_t)n, rank, scale, accumulate, sumsq);
  else {
    set_last_error("nvls_reduce_scatter: unsupported dtype combination %d -> %d", src_dtype, dst_dtype);
    return DSB_ERR_UNSUPPORTED;
  }
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

// In-place bf16 all-reduce of n elements at `off` (n % (8*world) == 0).  The caller brackets it with dsb200_symm_barrier.
DSB_EXPORT int dsb200_nvls_allreduce_bf16(void* mc_base, long long off, long long n, int rank, int world, int max_ctas,
                                          cudaStream_t stream) {
  DSB_REQUIRE(mc_base != nullptr, "nvls_allreduce: no multicast mapping");
  DSB_REQUIRE(world > 0 && n % (8LL * world) == 0 && off % 16 == 0, "nvls_allreduce: n must be a multiple of 8*world, offset 16-byte aligned");
  const size_t per = static_cast<size_t>(n) / 8 / world;
  nvls_allreduce_kernel<<<comm_grid((per + 3) / 4, max_ctas > 0 ? max_ctas : 64), 256, 0, stream>>>(
      static_cast<uint8_t*>(mc_base), static_cast<size_t>(off), per, rank);
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

DSB_EXPORT int dsb200_nvls_allgather(void* mc_base, long long dst_off, const void* src, long long bytes, int rank, int max_ctas,
                                     cudaStream_t stream) {
  DSB_REQUIRE(mc_base != nullptr, "nvls_allgather: no multicast mapping");
  DSB_REQUIRE(bytes % 16 == 0 && dst_off % 16 == 0, "nvls_allgather: sizes/offsets must be multiples of 16 bytes");
  nvls_allgather_kernel<<<comm_grid(static_cast<size_t>(bytes) / 16, max_ctas > 0 ? max_ctas : 64), 256, 0, stream>>>(
      static_cast<uint8_t*>(mc_base), static_cast<size_t>(dst_off), static_cast<const uint8_t*>(src), static_cast<size_t>(bytes), rank);
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

// qgZ consumer: q_off / s_off = byte offsets of the fp8 payload (W*n bytes) and the fp32 block scales (W*n/128 floats) in
// every rank's symmetric buffer; n = elements per rank (multiple of 128); dst dtype 0 bf16 / 2 fp32.
DSB_EXPORT int dsb200_p2p_fp8_reduce_scatter(void* const* peer_ptrs, long long q_off, long long s_off, void* dst, int dst_dtype,
                                             long long n, int rank, int world, float scale, int accumulate, float* sumsq,
                                             int max_ctas, cudaStream_t stream) {
  DSB_REQUIRE(n % 128 == 0 && q_off % 16 == 0 && s_off % 4 == 0, "p2p_fp8_reduce_scatter: n must be a multiple of 128 (whole scale blocks)");
  Peers P;
  int rc = make_peers(&P, peer_ptrs, world);
  if (rc) return rc;
  const int grid = comm_grid(static_cast<size_t>(n) / 16, max_ctas);
  if (dst_dtype == 0)
    fp8_reduce_scatter_pull_kernel<true><<<grid, 256, 0, stream>>>(P, (size_t)q_off, (size_t)s_off, dst, (size_t)n, rank, world, scale, accumulate, sumsq);
  else if (dst_dtype == 2)
    fp8_reduce_scatter_pull_kernel<false><<<grid, 256, 0, stream>>>(P, (size_t)q_off, (size_t)s_off, dst, (size_t)n, rank, world, scale, accumulate, sumsq);
  else {
    set_last_error("p2p_fp8_reduce_scatter: unsupported destination dtype %d", dst_dtype);
    return DSB_ERR_UNSUPPORTED;
  }
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

// ---- MoE expert-parallel dispatch / combine over peer memory.  buf_off: byte offset of the [W][n_local][capacity][M] bf16
// receive buffer inside every rank's symmetric buffer; token_expert / token_slot: [S][K] int32 (expert < 0 = dropped token).
DSB_EXPORT int dsb200_moe_dispatch_push(void* const* peer_ptrs, long long buf_off, const void* x, const int* token_expert,
                                        const int* token_slot, const float* row_scale, int S, int K, int M, int rank, int world,
                                        int n_local, int capacity, cudaStream_t stream) {
  DSB_REQUIRE(M % 8 == 0 && buf_off % 16 == 0 && S >= 0 && K >= 1, "moe_dispatch_push: model dim must be a multiple of 8");
  if (S == 0) return DSB_OK;
  Peers P;
  int rc = make_peers(&P, peer_ptrs, world);
  if (rc) return rc;
  const long long warps = static_cast<long long>(S) * K;
  moe_dispatch_push_kernel<<<static_cast<unsigned>((warps * 32 + 255) / 256), 256, 0, stream>>>(
      P, static_cast<size_t>(buf_off), static_cast<const __nv_bfloat16*>(x), token_expert, token_slot, row_scale, S, K, M, rank, n_local, capacity);
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

// out [S][M] bf16 (may be NULL when only dot_out is wanted); weights [S][K] fp32 or NULL; dot_with [S][M] bf16 + dot_out [S][K]
// fp32 (zero-initialised by the caller) optional.
DSB_EXPORT int dsb200_moe_combine_pull(void* const* peer_ptrs, long long buf_off, void* out, const int* token_expert,
                                       const int* token_slot, const float* weights, const void* dot_with, float* dot_out, int S, int K,
                                       int M, int rank, int world, int n_local, int capacity, cudaStream_t stream) {
  DSB_REQUIRE(M % 8 == 0 && buf_off % 16 == 0 && S >= 0 && K >= 1, "moe_combine_pull: model dim must be a multiple of 8");
  DSB_REQUIRE((dot_with == nullptr) == (dot_out == nullptr), "moe_combine_pull: dot_with and dot_out go together");
  if (S == 0) return DSB_OK;
  Peers P;
  int rc = make_peers(&P, peer_ptrs, world);
  if (rc) return rc;
  moe_combine_pull_kernel<<<static_cast<unsigned>((static_cast<long long>(S) * 32 + 255) / 256), 256, 0, stream>>>(
      P, static_cast<size_t>(buf_off), static_cast<__nv_bfloat16*>(out), token_expert, token_slot, weights,
      static_cast<const __nv_bfloat16*>(dot_with), dot_out, S, K, M, rank, n_local, capacity);
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

// Ulysses head <-> sequence exchange over peer memory.  `local`: [B * S_local][src_ld] bf16 (source layout); the dest-layout buffer
// [B * S_local * world][dst_ld] lives at buf_off of every rank's symmetric buffer.  segs: n_seg x {src_col, dst_col, heads,
// heads_per_rank} (ints, host).  push = 1: local -> owners' buffers; push = 0: owners' buffers -> local.
DSB_EXPORT int dsb200_ulysses_heads_exchange(void* const* peer_ptrs, long long buf_off, void* local, int B, int S_local, int rank, int world,
                                             int D, int src_ld, int dst_ld, const int* segs, int n_seg, int push, int max_ctas,
                                             cudaStream_t stream) {
  DSB_REQUIRE(D % 8 == 0 && src_ld % 8 == 0 && dst_ld % 8 == 0 && buf_off % 16 == 0 && n_seg >= 1 && n_seg <= 3,
              "ulysses_heads_exchange: head size and row strides must be multiples of 8, 1..3 segments");
  Peers P;
  int rc = make_peers(&P, peer_ptrs, world);
  if (rc) return rc;
  HeadSegs hs;
  hs.n = n_seg;
  hs.vec_per_row = 0;
  for (int i = 0; i < n_seg; ++i) {
    hs.s[i].src_col = segs[4 * i]; hs.s[i].dst_col = segs[4 * i + 1]; hs.s[i].heads = segs[4 * i + 2]; hs.s[i].heads_per_rank = segs[4 * i + 3];
    DSB_REQUIRE(hs.s[i].heads == hs.s[i].heads_per_rank * world && hs.s[i].src_col % 8 == 0 && hs.s[i].dst_col % 8 == 0,
                "ulysses_heads_exchange: segment %d: heads must split evenly over the ranks", i);
    hs.vec_per_row += hs.s[i].heads * (D / 8);
  }
  const long long total = static_cast<long long>(B) * S_local * hs.vec_per_row;
  if (total == 0) return DSB_OK;
  const int grid = comm_grid(static_cast<size_t>(total), max_ctas);
  if (push)
    ulysses_heads_kernel<true><<<grid, 256, 0, stream>>>(P, static_cast<size_t>(buf_off), static_cast<__nv_bfloat16*>(local), B, S_local, world, rank, D, src_ld, dst_ld, hs);
  else
    ulysses_heads_kernel<false><<<grid, 256, 0, stream>>>(P, static_cast<size_t>(buf_off), static_cast<__nv_bfloat16*>(local), B, S_local, world, rank, D, src_ld, dst_ld, hs);
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

// Skinny GEMM for decode steps: out[M, N] = x[M, K] w[N, K]^T with few rows (M <= 128 tokens) and a large weight matrix.
//
// The op is a read of the weights; what matters is that all 148 SMs stream them.  The square-tile kernels put 128 tokens x
// 256 features on a CTA: a 5120-feature projection of a 16-token batch is 20 CTAs, 1/7 of the machine.  Here the roles are
// swapped and K is split:
//     D^T [128 weight rows x NT tokens] += W_tile [128 x 64] (A, smem, K-major)  x  X_tile^T [NT x 64] (B, smem, K-major)
// Up to 8 K slices are the CTAs of one thread-block cluster and reduce through distributed shared memory (one kernel); more
// slices (or partials that do not fit the leader's stages) go through an fp32 workspace and a second kernel.
// grid = (ceil(N / 128), splits): every CTA streams a [128 x K/splits] slab of W through a 6-8 stage TMA ring (16 KB of W per
// stage -> ~100 KB in flight per SM) and issues 4 tcgen05.mma (M = 128, N = NT, K = 16) per stage; the accumulator is
// [128 lanes x NT columns] of TMEM.  splits == 1: the epilogue writes bf16 out[token][feature] directly; otherwise fp32
// partials go to a workspace [split][token][feature] and a second kernel sums them.  NT = 16 / 32 / 64 / 128 (rows past M are
// TMA zero fill: they cost no HBM traffic).
#include "../common/host.h"
#include "../common/ptx.cuh"

namespace dsb {
namespace skinny {

constexpr int BW = 128;  // weight rows per CTA (MMA M)
constexpr int BK = 64;   // k elements per stage (one 128-byte swizzle row)
constexpr int NUM_THREADS = 192;  // TMA warp, MMA warp, 4 epilogue warps

template <int NT>
struct Smem {
  static constexpr int STAGES = NT == 128 ? 6 : 8;  // 6 x 32 KB / 8 x (16 KB + the token tile)
  static constexpr int W_BYTES = BW * BK * 2;
  static constexpr int X_BYTES = NT * BK * 2;
  static constexpr int X_STRIDE = (X_BYTES + 1023) / 1024 * 1024;  // tiles stay 1024-byte aligned (128-byte swizzle atoms)
  static constexpr int STAGE = W_BYTES + X_STRIDE;
  static constexpr int BAR_OFF = STAGES * STAGE;
  static constexpr int TOTAL = BAR_OFF + 256 + 1024;
};

struct Params {
  int M, N, K;
  int kb_per_split, kb_total;
  __nv_bfloat16* out;  // [M][ldo] (splits == 1)
  int ldo;
  float* ws;           // [splits][M][N] (splits > 1)
};

template <int NT, bool CLUSTER>
__global__ void __launch_bounds__(NUM_THREADS, 1)
skinny_kernel(const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmX, const Params p) {
  using SM = Smem<NT>;
  constexpr int STAGES = SM::STAGES;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + SM::BAR_OFF);
  uint64_t* empty = full + STAGES;
  uint64_t* done = empty + STAGES;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(done + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n0 = blockIdx.x * BW;
  const int split = blockIdx.y;
  const int kb0 = split * p.kb_per_split;
  const int kb1 = min(p.kb_total, kb0 + p.kb_per_split);
  const int nkb = max(0, kb1 - kb0);

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmW);
    tma_prefetch_desc(&tmX);
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    mbar_init(done, 1);
    fence_barrier_init();
  }
  constexpr int TM_COLS = NT < 32 ? 32 : NT;
  if (warp == 1) {
    tmem_alloc<1>(tmem_ptr, TM_COLS);
    tmem_relinquish<1>();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = *tmem_ptr;

  if (nkb > 0) {
    if (warp == 0) {
      if (lane == 0) {
        int st = 0, ph = 0;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&empty[st], ph ^ 1);
          mbar_expect_tx(&full[st], SM::W_BYTES + SM::X_BYTES);
          uint8_t* base = smem + st * SM::STAGE;
          tma_load_2d(&tmW, &full[st], base, kb * BK, n0);
          tma_load_2d(&tmX, &full[st], base + SM::W_BYTES, kb * BK, 0);
          if (++st == STAGES) { st = 0; ph ^= 1; }
        }
      }
      __syncwarp();
      if constexpr (CLUSTER) { cluster_sync_all(); cluster_sync_all(); }
    } else if (warp == 1) {
      const bool leader = elect_one();
      constexpr uint32_t idesc = umma_idesc(BW, NT, 1, 1, 0, 0);
      const uint32_t tmu = __shfl_sync(0xffffffffu, tm, 0);
      const uint64_t w_desc0 = umma_smem_desc(smem_u32(smem), 16, 1024);
      const uint64_t x_desc0 = umma_smem_desc(smem_u32(smem + SM::W_BYTES), 16, 1024);
      int st = 0, ph = 0;
      for (int i = 0; i < nkb; ++i) {
        mbar_wait_lean(&full[st], ph);
        tc_fence_after();
        if (leader) {
          const uint32_t so = static_cast<uint32_t>((st * SM::STAGE) >> 4);
#pragma unroll
          for (int kk = 0; kk < BK / 16; ++kk)
            umma_f16<1>(tmu, w_desc0 + so + static_cast<uint32_t>((kk * 32) >> 4), x_desc0 + so + static_cast<uint32_t>((kk * 32) >> 4), idesc,
                        (i | kk) != 0 ? 1u : 0u);
          umma_commit<1>(&empty[st]);
          if (i == nkb - 1) umma_commit<1>(done);
        }
        if (++st == STAGES) { st = 0; ph ^= 1; }
      }
      if constexpr (CLUSTER) { cluster_sync_all(); cluster_sync_all(); }
    } else {
      // epilogue: thread == weight row (TMEM lane); a warp's 32 rows are 32 consecutive output features of one token
      const int q4 = warp & 3;
      const int r = q4 * 32 + lane;
      const int n = n0 + r;
      mbar_wait(done, 0);
      tc_fence_after();
      const uint32_t lane_off = static_cast<uint32_t>(q4 * 32) << 16;
      float acc[NT];
#pragma unroll
      for (int c = 0; c < NT / 16; ++c) {
        uint32_t v[16];
        tmem_ld_32x16(tm + c * 16 + lane_off, v);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 16; ++i) acc[c * 16 + i] = __uint_as_float(v[i]);
      }
      if constexpr (CLUSTER) {
        // split-K inside a thread-block cluster: the K slices of one row tile are the CTAs of a cluster (cluster rank ==
        // split); the non-leaders hand their fp32 accumulators to the leader through distributed shared memory -- its
        // pipeline stages are idle by then -- laid out [rank - 1][token][weight row] (a warp stores 32 consecutive floats)
        cluster_sync_all();  // every CTA's tensor core and TMA are done with their stages
        float* red = reinterpret_cast<float*>(smem);
        if (split != 0) {
          uint32_t remote;
          asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(smem_u32(red + (static_cast<size_t>(split - 1) * NT) * BW + r)), "r"(0));
#pragma unroll
          for (int t = 0; t < NT; ++t)
            asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(remote + static_cast<uint32_t>(t * BW * 4)), "f"(acc[t]) : "memory");
        }
        cluster_sync_all();  // partials have landed in the leader's shared memory
        if (split == 0) {
          const int S = static_cast<int>(gridDim.y);
#pragma unroll
          for (int t = 0; t < NT; ++t) {
            float a = acc[t];
            for (int q = 0; q < S - 1; ++q) a += red[(static_cast<size_t>(q) * NT + t) * BW + r];
            if (n < p.N && t < p.M) p.out[static_cast<size_t>(t) * p.ldo + n] = __float2bfloat16_rn(a);
          }
        }
      } else if (n < p.N) {
#pragma unroll
        for (int t = 0; t < NT; ++t) {
          if (t < p.M) {
            if (p.ws != nullptr) p.ws[(static_cast<size_t>(split) * p.M + t) * p.N + n] = acc[t];
            else p.out[static_cast<size_t>(t) * p.ldo + n] = __float2bfloat16_rn(acc[t]);
          }
        }
      }
    }
  } else if (CLUSTER) {
    // a split without k-blocks still joins the cluster protocol (its partial is zero)
    cluster_sync_all();
    if (warp >= 2 && split != 0) {
      const int r = (warp & 3) * 32 + lane;
      float* red = reinterpret_cast<float*>(smem);
      uint32_t remote;
      asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(smem_u32(red + (static_cast<size_t>(split - 1) * NT) * BW + r)), "r"(0));
      for (int t = 0; t < NT; ++t)
        asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(remote + static_cast<uint32_t>(t * BW * 4)), "f"(0.f) : "memory");
    }
    cluster_sync_all();
  } else if (warp >= 2 && p.ws != nullptr) {
    // this split has no k-blocks (K shorter than the split grid): its partial is zero
    const int n = n0 + (warp & 3) * 32 + lane;
    if (n < p.N)
      for (int t = 0; t < p.M; ++t) p.ws[(static_cast<size_t>(split) * p.M + t) * p.N + n] = 0.f;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<1>(tm, TM_COLS);
  }
}

// out[t][n] = sum_s ws[s][t][n]; 4 features per thread
__global__ void skinny_reduce_kernel(const float* __restrict__ ws, __nv_bfloat16* __restrict__ out, int M, int N, int ldo, int splits) {
  const long long i4 = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) * 4;
  if (i4 >= static_cast<long long>(M) * N) return;
  const int t = static_cast<int>(i4 / N), n = static_cast<int>(i4 - static_cast<long long>(t) * N);
  float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int s = 0; s < splits; ++s) {
    const float4 v = *reinterpret_cast<const float4*>(ws + (static_cast<size_t>(s) * M + t) * N + n);
    a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w;
  }
  __nv_bfloat16* o = out + static_cast<size_t>(t) * ldo + n;
  *reinterpret_cast<uint2*>(o) = make_uint2(pack_bf16x2(a.x, a.y), pack_bf16x2(a.z, a.w));
}

template <int NT>
static int launch(const void* x, const void* w, void* out, float* ws, int M, int N, int K, int ldx, int ldw, int ldo, int splits,
                  cudaStream_t stream) {
  using SM = Smem<NT>;
  CUtensorMap tmW, tmX;
  int rc = make_tensor_map_2d(&tmW, TMA_BF16, w, N, K, ldw, BW, BK, TMA_SW_128);
  if (rc) return rc;
  rc = make_tensor_map_2d(&tmX, TMA_BF16, x, M, K, ldx, NT, BK, TMA_SW_128);
  if (rc) return rc;
  Params p;
  p.M = M; p.N = N; p.K = K;
  p.kb_total = (K + BK - 1) / BK;
  p.kb_per_split = (p.kb_total + splits - 1) / splits;
  p.out = static_cast<__nv_bfloat16*>(out);
  p.ldo = ldo;
  dim3 grid((N + BW - 1) / BW, splits);
  // up to 8 K slices whose fp32 partials fit the leader's pipeline stages: one kernel, reduction over distributed shared memory
  const bool cluster = splits > 1 && splits <= 8 && static_cast<size_t>(splits - 1) * NT * BW * 4 <= static_cast<size_t>(SM::STAGES) * SM::STAGE;
  if (cluster) {
    auto kern = skinny_kernel<NT, true>;
    static bool attr_set = false;
    if (!attr_set) {
      DSB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SM::TOTAL));
      attr_set = true;
    }
    p.ws = nullptr;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = dim3(NUM_THREADS);
    cfg.dynamicSmemBytes = SM::TOTAL;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 1;
    attr[0].val.clusterDim.y = static_cast<unsigned>(splits);
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    DSB_CHECK_CUDA(cudaLaunchKernelEx(&cfg, kern, tmW, tmX, p));
    return DSB_OK;
  }
  auto kern = skinny_kernel<NT, false>;
  static bool attr_set = false;
  if (!attr_set) {
    DSB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SM::TOTAL));
    attr_set = true;
  }
  p.ws = splits > 1 ? ws : nullptr;
  kern<<<grid, NUM_THREADS, SM::TOTAL, stream>>>(tmW, tmX, p);
  DSB_CHECK_CUDA(cudaGetLastError());
  if (splits > 1) {
    const long long quads = static_cast<long long>(M) * N / 4;
    skinny_reduce_kernel<<<static_cast<unsigned>((quads + 255) / 256), 256, 0, stream>>>(ws, static_cast<__nv_bfloat16*>(out), M, N, ldo, splits);
    DSB_CHECK_CUDA(cudaGetLastError());
  }
  return DSB_OK;
}

}  // namespace skinny
}  // namespace dsb

// out[M][N] (bf16, row stride ldo) = x[M][K] w[N][K]^T, bf16 operands, M <= 128.  splits >= 1 cuts K; with splits > 1 the caller
// passes a workspace of splits * M * N floats and N must be a multiple of 4.  K, ldx, ldw multiples of 8 (16-byte rows for TMA).
DSB_EXPORT int dsb200_gemm_skinny_bf16(const void* x, const void* w, void* out, void* workspace, int M, int N, int K, int ldx, int ldw,
                                       int ldo, int splits, cudaStream_t stream) {
  using namespace dsb;
  DSB_REQUIRE(M >= 1 && M <= 128 && N >= 1 && K >= 1, "gemm_skinny: M=%d (1..128) N=%d K=%d", M, N, K);
  DSB_REQUIRE(K % 8 == 0 && ldx % 8 == 0 && ldw % 8 == 0, "gemm_skinny: K, ldx, ldw must be multiples of 8");
  DSB_REQUIRE(splits >= 1 && (splits == 1 || (workspace != nullptr && N % 4 == 0 && ldo % 4 == 0)),
              "gemm_skinny: splits=%d needs a workspace and N, ldo multiples of 4", splits);
  float* ws = static_cast<float*>(workspace);
  if (M <= 16) return skinny::launch<16>(x, w, out, ws, M, N, K, ldx, ldw, ldo, splits, stream);
  if (M <= 32) return skinny::launch<32>(x, w, out, ws, M, N, K, ldx, ldw, ldo, splits, stream);
  if (M <= 64) return skinny::launch<64>(x, w, out, ws, M, N, K, ldx, ldw, ldo, splits, stream);
  return skinny::launch<128>(x, w, out, ws, M, N, K, ldx, ldw, ldo, splits, stream);
}

// ================================================================================================ weight-only quantised weights
// Mixed-input form (16-bit activations x 8- / 4-bit weights with group-wise scales; reference: inference/v2/kernels/cutlass_ops/
// mixed_gemm, an external CUTLASS kernel there).  Same weight-streaming structure; the weight tile arrives as raw integers
// (TMA, 8 or 4 KB per stage), the four epilogue warps dequantise it -- thread == weight row: 64 values x the row's group scale ->
// bf16 in the 128-byte-swizzled K-major layout the MMA reads -- and hand it to the MMA warp.  HBM traffic is the quantised bytes.
//   8-bit: int8 [N][K];  4-bit: uint8 [N][K/2], low nibble first, value = nibble - 8;  scales fp32 [N][K / group], group % 64 == 0.
namespace dsb {
namespace skinny {

template <int NT, int BITS>
struct MixSmem {
  static constexpr int STAGES = NT == 128 ? 5 : 6;
  static constexpr int RAW_BYTES = BW * BK * BITS / 8;
  static constexpr int W_BYTES = BW * BK * 2;
  static constexpr int X_BYTES = NT * BK * 2;
  static constexpr int X_STRIDE = (X_BYTES + 1023) / 1024 * 1024;
  static constexpr int STAGE = W_BYTES + X_STRIDE + RAW_BYTES;  // [bf16 W tile][X tile][raw tile]
  static constexpr int BAR_OFF = STAGES * STAGE;
  static constexpr int TOTAL = BAR_OFF + 256 + 1024;
};

struct MixParams {
  Params g;
  const float* scales;  // [N][K / group]
  int group, n_groups;
};

template <int NT, int BITS>
__global__ void __launch_bounds__(NUM_THREADS, 1)
skinny_mixed_kernel(const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmX, const MixParams mp) {
  using SM = MixSmem<NT, BITS>;
  constexpr int STAGES = SM::STAGES;
  const Params& p = mp.g;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* raw_full = reinterpret_cast<uint64_t*>(smem + SM::BAR_OFF);  // TMA -> dequant warps (raw W) and MMA warp (X)
  uint64_t* w_full = raw_full + STAGES;                                   // dequant warps -> MMA warp
  uint64_t* empty = w_full + STAGES;
  uint64_t* done = empty + STAGES;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(done + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n0 = blockIdx.x * BW;
  const int split = blockIdx.y;
  const int kb0 = split * p.kb_per_split;
  const int kb1 = min(p.kb_total, kb0 + p.kb_per_split);
  const int nkb = max(0, kb1 - kb0);

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmW);
    tma_prefetch_desc(&tmX);
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&raw_full[i], 1);
      mbar_init(&w_full[i], 128);
      mbar_init(&empty[i], 1);
    }
    mbar_init(done, 1);
    fence_barrier_init();
  }
  constexpr int TM_COLS = NT < 32 ? 32 : NT;
  if (warp == 1) {
    tmem_alloc<1>(tmem_ptr, TM_COLS);
    tmem_relinquish<1>();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = *tmem_ptr;

  if (nkb > 0) {
    if (warp == 0) {
      if (lane == 0) {
        int st = 0, ph = 0;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&empty[st], ph ^ 1);
          mbar_expect_tx(&raw_full[st], SM::RAW_BYTES + SM::X_BYTES);
          uint8_t* base = smem + st * SM::STAGE;
          tma_load_2d(&tmW, &raw_full[st], base + SM::W_BYTES + SM::X_STRIDE, kb * (BK * BITS / 8), n0);
          tma_load_2d(&tmX, &raw_full[st], base + SM::W_BYTES, kb * BK, 0);
          if (++st == STAGES) { st = 0; ph ^= 1; }
        }
      }
    } else if (warp == 1) {
      const bool leader = elect_one();
      constexpr uint32_t idesc = umma_idesc(BW, NT, 1, 1, 0, 0);
      const uint32_t tmu = __shfl_sync(0xffffffffu, tm, 0);
      const uint64_t w_desc0 = umma_smem_desc(smem_u32(smem), 16, 1024);
      const uint64_t x_desc0 = umma_smem_desc(smem_u32(smem + SM::W_BYTES), 16, 1024);
      int st = 0, ph = 0;
      for (int i = 0; i < nkb; ++i) {
        mbar_wait_lean(&raw_full[st], ph);  // X tile (and the raw weights) have landed
        mbar_wait_lean(&w_full[st], ph);    // the dequantised weight tile is written
        tc_fence_after();
        if (leader) {
          const uint32_t so = static_cast<uint32_t>((st * SM::STAGE) >> 4);
#pragma unroll
          for (int kk = 0; kk < BK / 16; ++kk)
            umma_f16<1>(tmu, w_desc0 + so + static_cast<uint32_t>((kk * 32) >> 4), x_desc0 + so + static_cast<uint32_t>((kk * 32) >> 4), idesc,
                        (i | kk) != 0 ? 1u : 0u);
          umma_commit<1>(&empty[st]);
          if (i == nkb - 1) umma_commit<1>(done);
        }
        if (++st == STAGES) { st = 0; ph ^= 1; }
      }
    } else {
      const int q4 = warp & 3;
      const int r = q4 * 32 + lane;  // weight row of the tile == TMEM lane
      const int n = n0 + r;
      const float* srow = mp.scales + static_cast<size_t>(min(n, p.N - 1)) * mp.n_groups;
      int st = 0, ph = 0;
      for (int i = 0; i < nkb; ++i) {
        const float sc = srow[((kb0 + i) * BK) / mp.group];  // issued before the wait: its latency hides behind the TMA's
        mbar_wait(&raw_full[st], ph);
        uint8_t* base = smem + st * SM::STAGE;
        const uint8_t* raw = base + SM::W_BYTES + SM::X_STRIDE + r * (BK * BITS / 8);
        uint8_t* wrow = base + r * 128;
        const int sw = r & 7;
#pragma unroll
        for (int u = 0; u < 8; ++u) {  // 8 values -> one 16-byte unit of the bf16 row
          float f[8];
          if constexpr (BITS == 8) {
            const uint2 q = *reinterpret_cast<const uint2*>(raw + u * 8);
            const uint32_t w2[2] = {q.x, q.y};
#pragma unroll
            for (int j = 0; j < 8; ++j) f[j] = static_cast<float>(static_cast<int8_t>((w2[j >> 2] >> ((j & 3) * 8)) & 0xff));
          } else {
            const uint32_t q = *reinterpret_cast<const uint32_t*>(raw + u * 4);
#pragma unroll
            for (int j = 0; j < 8; ++j) f[j] = static_cast<float>(static_cast<int>((q >> (j * 4)) & 0xf) - 8);
          }
          uint4 v;
          v.x = pack_bf16x2(f[0] * sc, f[1] * sc);
          v.y = pack_bf16x2(f[2] * sc, f[3] * sc);
          v.z = pack_bf16x2(f[4] * sc, f[5] * sc);
          v.w = pack_bf16x2(f[6] * sc, f[7] * sc);
          *reinterpret_cast<uint4*>(wrow + ((u ^ sw) << 4)) = v;
        }
        fence_proxy_async_smem();
        mbar_arrive(&w_full[st]);
        if (++st == STAGES) { st = 0; ph ^= 1; }
      }
      // epilogue (as in skinny_kernel)
      mbar_wait(done, 0);
      tc_fence_after();
      const uint32_t lane_off = static_cast<uint32_t>(q4 * 32) << 16;
#pragma unroll
      for (int c = 0; c < NT / 16; ++c) {
        uint32_t v[16];
        tmem_ld_32x16(tm + c * 16 + lane_off, v);
        tmem_ld_wait();
        if (n < p.N) {
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            const int t = c * 16 + i;
            if (t < p.M) {
              if (p.ws != nullptr) p.ws[(static_cast<size_t>(split) * p.M + t) * p.N + n] = __uint_as_float(v[i]);
              else p.out[static_cast<size_t>(t) * p.ldo + n] = __float2bfloat16_rn(__uint_as_float(v[i]));
            }
          }
        }
      }
    }
  } else if (warp >= 2 && p.ws != nullptr) {
    const int n = n0 + (warp & 3) * 32 + lane;
    if (n < p.N)
      for (int t = 0; t < p.M; ++t) p.ws[(static_cast<size_t>(split) * p.M + t) * p.N + n] = 0.f;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<1>(tm, TM_COLS);
  }
}

template <int NT, int BITS>
static int launch_mixed(const void* x, const void* wq, const float* scales, void* out, float* ws, int M, int N, int K, int ldx, int ldo,
                        int group, int splits, cudaStream_t stream) {
  using SM = MixSmem<NT, BITS>;
  CUtensorMap tmW, tmX;
  const int row_bytes = K * BITS / 8;
  int rc = make_tensor_map_2d(&tmW, TMA_U8, wq, N, row_bytes, row_bytes, BW, BK * BITS / 8, TMA_SW_NONE);
  if (rc) return rc;
  rc = make_tensor_map_2d(&tmX, TMA_BF16, x, M, K, ldx, NT, BK, TMA_SW_128);
  if (rc) return rc;
  auto kern = skinny_mixed_kernel<NT, BITS>;
  static bool attr_set = false;
  if (!attr_set) {
    DSB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SM::TOTAL));
    attr_set = true;
  }
  MixParams mp;
  Params& p = mp.g;
  p.M = M; p.N = N; p.K = K;
  p.kb_total = (K + BK - 1) / BK;
  p.kb_per_split = (p.kb_total + splits - 1) / splits;
  p.out = static_cast<__nv_bfloat16*>(out);
  p.ldo = ldo;
  p.ws = splits > 1 ? ws : nullptr;
  mp.scales = scales;
  mp.group = group;
  mp.n_groups = K / group;
  dim3 grid((N + BW - 1) / BW, splits);
  kern<<<grid, NUM_THREADS, SM::TOTAL, stream>>>(tmW, tmX, mp);
  DSB_CHECK_CUDA(cudaGetLastError());
  if (splits > 1) {
    const long long quads = static_cast<long long>(M) * N / 4;
    skinny_reduce_kernel<<<static_cast<unsigned>((quads + 255) / 256), 256, 0, stream>>>(ws, static_cast<__nv_bfloat16*>(out), M, N, ldo, splits);
    DSB_CHECK_CUDA(cudaGetLastError());
  }
  return DSB_OK;
}

template <int BITS>
static int dispatch_mixed(const void* x, const void* wq, const float* scales, void* out, float* ws, int M, int N, int K, int ldx, int ldo,
                          int group, int splits, cudaStream_t stream) {
  if (M <= 16) return launch_mixed<16, BITS>(x, wq, scales, out, ws, M, N, K, ldx, ldo, group, splits, stream);
  if (M <= 32) return launch_mixed<32, BITS>(x, wq, scales, out, ws, M, N, K, ldx, ldo, group, splits, stream);
  if (M <= 64) return launch_mixed<64, BITS>(x, wq, scales, out, ws, M, N, K, ldx, ldo, group, splits, stream);
  return launch_mixed<128, BITS>(x, wq, scales, out, ws, M, N, K, ldx, ldo, group, splits, stream);
}

}  // namespace skinny
}  // namespace dsb

// out[M][N] (bf16) = x[M][K] (bf16) * dequant(wq)[N][K]^T, M <= 128.  bits = 8: wq int8 [N][K]; bits = 4: wq uint8 [N][K/2] (low nibble
// first, value = nibble - 8).  scales fp32 [N][K / group], group a multiple of 64 dividing K; K a multiple of 64.
// splits > 1 needs a workspace of splits * M * N floats and N, ldo multiples of 4.
DSB_EXPORT int dsb200_gemm_skinny_mixed(const void* x, const void* wq, const void* scales, void* out, void* workspace, int M, int N, int K,
                                        int ldx, int ldo, int bits, int group, int splits, cudaStream_t stream) {
  using namespace dsb;
  DSB_REQUIRE(M >= 1 && M <= 128 && N >= 1 && K >= 64, "gemm_skinny_mixed: M=%d (1..128) N=%d K=%d", M, N, K);
  DSB_REQUIRE(bits == 8 || bits == 4, "gemm_skinny_mixed: bits=%d (4 or 8)", bits);
  DSB_REQUIRE(K % 64 == 0 && ldx % 8 == 0 && group >= 64 && group % 64 == 0 && K % group == 0,
              "gemm_skinny_mixed: K=%d must be a multiple of 64 and of the group size %d (itself a multiple of 64)", K, group);
  DSB_REQUIRE(splits >= 1 && (splits == 1 || (workspace != nullptr && N % 4 == 0 && ldo % 4 == 0)),
              "gemm_skinny_mixed: splits=%d needs a workspace and N, ldo multiples of 4", splits);
  float* ws = static_cast<float*>(workspace);
  const float* sc = static_cast<const float*>(scales);
  if (bits == 8) return skinny::dispatch_mixed<8>(x, wq, sc, out, ws, M, N, K, ldx, ldo, group, splits, stream);
  return skinny::dispatch_mixed<4>(x, wq, sc, out, ws, M, N, K, ldx, ldo, group, splits, stream);
}

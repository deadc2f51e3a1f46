// MXFP8 GEMM on the hardware block-scaled tensor-core path (tcgen05.mma.kind::mxf8f6f4.block_scale):
//     C[m][n] = sum_k (A[m][k] * 2^(sfa[m][k/32] - 127)) * (B[n][k] * 2^(sfb[n][k/32] - 127))
// A [M][K], B [N][K] e4m3 payloads (K contiguous, K % 128 == 0); one UE8M0 scale byte per 32 consecutive k-elements of every row,
// delivered in the packed tile layout the tensor core wants (written by mxfp8_quantize_kernel below): for every block of 128
// rows and every 128-wide K block, 512 bytes with byte (row % 32) * 16 + (row % 128 / 32) * 4 + (k / 32 % 4) holding the scale of
// (row, k).  One pipeline stage = one K block: the TMA producer brings the A / B tiles and the 512-byte scale chunks (plain bulk
// copies), the MMA lane copies the chunks into tensor memory with tcgen05.cp (32x128b.warpx4: each of the four lane quadrants
// gets the 32 x 16-byte pattern) and issues four K = 32 block-scaled MMAs whose scale-factor id selects the byte.  Scaling
// costs no instruction anywhere else: the accumulator in TMEM is final, the epilogue only converts to bf16.
// Tiles 128 x 256, four stages, accumulator 256 TMEM columns + two alternating scale areas of 12 columns.
// Reference counterpart: none (no fp8 path at all); role: SURVEY K2, BASELINE config #5 "fp8 block-scaled".
#include <cuda_fp8.h>

#include "gemm_core.cuh"

namespace dsb {
namespace gemm8mx {

constexpr int BM = 128;
constexpr int BN = 256;
constexpr int BKB = 128;
constexpr int STAGES = 4;
constexpr int A_BYTES = BM * BKB;
constexpr int B_BYTES = BN * BKB;
constexpr int TILE_BYTES = A_BYTES + B_BYTES;        // 48 KB, keeps 1024-byte alignment per stage
constexpr int SF_CHUNK = 512;                        // scales of 128 rows x one K block
constexpr int SF_STAGE = 3 * SF_CHUNK;               // A chunk + two B chunks
constexpr int SF_OFFSET = STAGES * TILE_BYTES;
constexpr int EPI_OFFSET = SF_OFFSET + STAGES * SF_STAGE;
constexpr int EPI_BUF = 128 * 128;                   // 128 rows x 64 bf16
constexpr int BAR_OFFSET = EPI_OFFSET + EPI_BUF;
constexpr int SMEM_TOTAL = BAR_OFFSET + 256 + 1024;
constexpr int NUM_THREADS = 192;
constexpr int EPI_THREADS = 128;
constexpr uint32_t SF_COL = 256;                     // TMEM columns: [0, 256) accumulator, then 2 x (4 SFA + 8 SFB)

struct Params {
  int M, N, K, KB;
  int tiles_m, tiles_n;
  const uint8_t* sfa;  // packed: [ceil(M / 128)][KB][512]
  const uint8_t* sfb;  // packed: [2 * ceil(N / 256)][KB][512]
};

__device__ __forceinline__ void bulk_load(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(smem_dst)),
               "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
// 32 rows x 16 bytes in shared memory (rows 16 bytes apart: 8-row core matrices 128 bytes apart, no swizzle) -> 4 TMEM columns in
// every lane quadrant
__device__ __forceinline__ void sf_to_tmem(uint32_t tmem_addr, uint32_t smem_addr) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr >> 4) & 0x3FFF);
  d |= static_cast<uint64_t>((128u >> 4) & 0x3FFF) << 32;  // SBO = 8 rows * 16 B; LBO unused (one 16-byte column)
  d |= 1ull << 46;                                         // descriptor version
  asm volatile("tcgen05.cp.cta_group::1.32x128b.warpx4 [%0], %1;" ::"r"(tmem_addr), "l"(d) : "memory");
}
__device__ __forceinline__ void umma_mxf8(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t tmem_sfa, uint32_t tmem_sfb,
                                          uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::mxf8f6f4.block_scale [%0], %1, %2, %3, [%5], [%6], p;\n\t}\n" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate), "r"(tmem_sfa), "r"(tmem_sfb)
      : "memory");
}
// block-scaled instruction descriptor: e4m3 x e4m3 (formats 0), K-major operands, UE8M0 scales; sf ids select the K = 32 sub-block
__host__ __device__ constexpr uint32_t idesc_mx(int M, int N, uint32_t sf_id) {
  return (sf_id << 4) | (static_cast<uint32_t>(N >> 3) << 17) | (1u << 23) | (static_cast<uint32_t>(M >> 4) << 24) | (sf_id << 29);
}

__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_mxfp8_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                  const __grid_constant__ CUtensorMap tmC, const Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sf_smem = smem + SF_OFFSET;
  uint8_t* epi_buf = smem + EPI_OFFSET;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + BAR_OFFSET);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full = empty_bar + STAGES;
  uint64_t* tmem_empty = tmem_full + 1;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tmem_empty + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int num_tiles = p.tiles_m * p.tiles_n;
  const int num_kb = p.KB;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    tma_prefetch_desc(&tmC);
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    mbar_init(tmem_full, 1);
    mbar_init(tmem_empty, EPI_THREADS);
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
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        int m_blk, n_blk;
        gemm::tile_coords(tile, p.tiles_m, p.tiles_n, m_blk, n_blk);
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait_lean(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * TILE_BYTES;
          uint8_t* sf = sf_smem + stage * SF_STAGE;
          mbar_expect_tx(&full_bar[stage], TILE_BYTES + SF_STAGE);
          tma_load_2d(&tmA, &full_bar[stage], sa, kb * BKB, m_blk * BM);
          tma_load_2d(&tmB, &full_bar[stage], sa + A_BYTES, kb * BKB, n_blk * BN);
          bulk_load(sf, p.sfa + (static_cast<size_t>(m_blk) * num_kb + kb) * SF_CHUNK, SF_CHUNK, &full_bar[stage]);
          bulk_load(sf + SF_CHUNK, p.sfb + (static_cast<size_t>(2 * n_blk) * num_kb + kb) * SF_CHUNK, SF_CHUNK, &full_bar[stage]);
          bulk_load(sf + 2 * SF_CHUNK, p.sfb + (static_cast<size_t>(2 * n_blk + 1) * num_kb + kb) * SF_CHUNK, SF_CHUNK, &full_bar[stage]);
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    constexpr uint32_t HI = (1024u >> 4) | (1u << 14) | (2u << 29);
    constexpr uint32_t KM = (16u >> 4) << 16;
    const uint32_t tm = __shfl_sync(0xffffffffu, tmem_base, 0);
    const bool leader = elect_one();
    const uint32_t base = smem_u32(smem);
    const uint32_t sf_base = smem_u32(sf_smem);
    int stage = 0;
    uint32_t phase = 0, acc_phase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      mbar_wait_lean(tmem_empty, acc_phase ^ 1);  // the epilogue has drained the accumulator of the previous tile
      tc_fence_after();
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait_lean(&full_bar[stage], phase);
        tc_fence_after();
        if (leader) {
          const uint32_t sfc = tm + SF_COL + (kb & 1) * 16;  // alternate two scale areas; copies and MMAs retire in issue order
          const uint32_t sfs = sf_base + stage * SF_STAGE;
          sf_to_tmem(sfc, sfs);
          sf_to_tmem(sfc + 4, sfs + SF_CHUNK);
          sf_to_tmem(sfc + 8, sfs + 2 * SF_CHUNK);
          const uint32_t a_lo = KM | ((base + stage * TILE_BYTES) >> 4);
          const uint32_t b_lo = KM | ((base + stage * TILE_BYTES + A_BYTES) >> 4);
#pragma unroll
          for (uint32_t kk = 0; kk < BKB / 32; ++kk)
            umma_mxf8(tm, (static_cast<uint64_t>(HI) << 32) | (a_lo + kk * 2), (static_cast<uint64_t>(HI) << 32) | (b_lo + kk * 2),
                      idesc_mx(BM, BN, kk), sfc, sfc + 4, (kb | static_cast<int>(kk)) != 0 ? 1u : 0u);
          umma_commit<1>(&empty_bar[stage]);
        }
        if (++stage == STAGES) {
          stage = 0;
          phase ^= 1;
        }
      }
      if (leader) umma_commit<1>(tmem_full);
      acc_phase ^= 1;
    }
  } else {
    const int q = warp & 3;
    const int row = q * 32 + lane;
    const int epi_tid = threadIdx.x - 64;
    constexpr int CW = 64;
    uint32_t acc_phase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      int m_blk, n_blk;
      gemm::tile_coords(tile, p.tiles_m, p.tiles_n, m_blk, n_blk);
      const int m0 = m_blk * BM, n0 = n_blk * BN;
      mbar_wait_lean(tmem_full, acc_phase);
      tc_fence_after();
      const uint32_t t_acc = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
#pragma unroll 1
      for (int c = 0; c < BN / CW; ++c) {
        const bool live = (n0 + c * CW) < p.N;
        uint32_t v[32], w[32];
        if (live) {
          tmem_ld_32x32(t_acc + c * CW, v);
          tmem_ld_32x32(t_acc + c * CW + 32, w);
          tmem_ld_wait();
        }
        if (c == BN / CW - 1) {
          tc_fence_before();
          mbar_arrive(tmem_empty);
        }
        if (!live) continue;
        if (epi_tid == 0) tma_wait_group_read<0>();  // single staging buffer: the previous store has read it
        named_bar_sync(1, EPI_THREADS);
        uint8_t* rowp = epi_buf + row * 128;
        const int sw = row & 7;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const uint32_t* src = j < 4 ? &v[8 * j] : &w[8 * (j - 4)];
          uint4 val;
          val.x = pack_bf16x2(__uint_as_float(src[0]), __uint_as_float(src[1]));
          val.y = pack_bf16x2(__uint_as_float(src[2]), __uint_as_float(src[3]));
          val.z = pack_bf16x2(__uint_as_float(src[4]), __uint_as_float(src[5]));
          val.w = pack_bf16x2(__uint_as_float(src[6]), __uint_as_float(src[7]));
          *reinterpret_cast<uint4*>(rowp + ((j ^ sw) << 4)) = val;
        }
        fence_proxy_async_smem();
        named_bar_sync(1, EPI_THREADS);
        if (epi_tid == 0) {
          tma_store_2d(&tmC, epi_buf, n0 + c * CW, m0);
          tma_commit_group();
        }
      }
      acc_phase ^= 1;
    }
    if (epi_tid == 0) tma_wait_group<0>();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<1>(tmem_base, 512);
  }
}

// ---------------------------------------------------------------- MXFP8 quantiser: payload + packed UE8M0 scale tiles
// One warp per (row, 128-wide K block): 4 elements per lane, 8 lanes per 32-element scale group.  Scale = the smallest power of
// two >= amax / 448, stored as its biased exponent.
template <typename T>
__global__ void mxfp8_quantize_kernel(const T* __restrict__ x, uint8_t* __restrict__ q, uint8_t* __restrict__ sf, int rows, int K, long long ldx) {
  const long long wid = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  const int KB = K / 128;
  if (wid >= static_cast<long long>(rows) * KB) return;
  const int row = static_cast<int>(wid / KB), kb = static_cast<int>(wid - static_cast<long long>(row) * KB);
  const T* src = x + static_cast<size_t>(row) * ldx + kb * 128 + lane * 4;
  float v[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) v[i] = static_cast<float>(src[i]);
  float amax = fmaxf(fmaxf(fabsf(v[0]), fabsf(v[1])), fmaxf(fabsf(v[2]), fabsf(v[3])));
#pragma unroll
  for (int o = 1; o < 8; o <<= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
  const uint32_t bits = __float_as_uint(amax * (1.f / 448.f));
  uint32_t ue = ((bits >> 23) & 0xffu) + ((bits & 0x7fffffu) != 0 ? 1u : 0u);  // ceil(log2) as a biased exponent
  ue = min(max(ue, 1u), 253u);
  const float inv = __uint_as_float((254u - ue) << 23);  // 2^-(ue - 127)
  uint32_t packed = 0;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    __nv_fp8_e4m3 f(fminf(fmaxf(v[i] * inv, -448.f), 448.f));
    packed |= static_cast<uint32_t>(f.__x) << (8 * i);
  }
  *reinterpret_cast<uint32_t*>(q + static_cast<size_t>(row) * K + kb * 128 + lane * 4) = packed;
  if ((lane & 7) == 0) {
    const int r = row & 127;
    sf[(static_cast<size_t>(row >> 7) * KB + kb) * 512 + (r & 31) * 16 + (r >> 5) * 4 + (lane >> 3)] = static_cast<uint8_t>(ue);
  }
}

}  // namespace gemm8mx
}  // namespace dsb

// Bytes of the packed scale tensor for a [rows][K] operand (rows padded to a multiple of 256 so either operand role works).
DSB_EXPORT long long dsb200_mxfp8_sf_bytes(int rows, int K) {
  return static_cast<long long>((rows + 255) / 256) * 2 * (K / 128) * 512;
}

// x [rows][K] (row stride ldx elements; dtype 0 bf16, 1 fp16, 2 fp32; K % 128 == 0) -> q [rows][K] e4m3 + sf (dsb200_mxfp8_sf_bytes,
// ZERO-INITIALISED by the caller: rows beyond `rows` keep scale byte 0).
DSB_EXPORT int dsb200_mxfp8_quantize(const void* x, void* q, void* sf, int rows, int K, long long ldx, int dtype, cudaStream_t stream) {
  using namespace dsb;
  DSB_REQUIRE(K > 0 && K % 128 == 0 && rows >= 0 && ldx >= K, "mxfp8_quantize: K (%d) must be a positive multiple of 128", K);
  if (rows == 0) return DSB_OK;
  const long long warps = static_cast<long long>(rows) * (K / 128);
  const unsigned grid = static_cast<unsigned>((warps * 32 + 255) / 256);
  uint8_t* qq = static_cast<uint8_t*>(q);
  uint8_t* ss = static_cast<uint8_t*>(sf);
  if (dtype == 0) gemm8mx::mxfp8_quantize_kernel<__nv_bfloat16><<<grid, 256, 0, stream>>>(static_cast<const __nv_bfloat16*>(x), qq, ss, rows, K, ldx);
  else if (dtype == 1) gemm8mx::mxfp8_quantize_kernel<__half><<<grid, 256, 0, stream>>>(static_cast<const __half*>(x), qq, ss, rows, K, ldx);
  else if (dtype == 2) gemm8mx::mxfp8_quantize_kernel<float><<<grid, 256, 0, stream>>>(static_cast<const float*>(x), qq, ss, rows, K, ldx);
  else {
    set_last_error("mxfp8_quantize: unsupported dtype %d", dtype);
    return DSB_ERR_UNSUPPORTED;
  }
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

// C[M,N] (bf16) = dequant(A) * dequant(B)^T with MXFP8 operands (payload + packed UE8M0 scales from dsb200_mxfp8_quantize).
DSB_EXPORT int dsb200_gemm_mxfp8(const void* A, const void* B, void* C, const void* sfa, const void* sfb, int M, int N, int K, int lda,
                                 int ldb, int ldc, int max_ctas, cudaStream_t stream) {
  using namespace dsb;
  DSB_REQUIRE(M > 0 && N > 0 && K > 0 && K % 128 == 0, "gemm_mxfp8: K (%d) must be a positive multiple of 128", K);
  DSB_REQUIRE(sfa != nullptr && sfb != nullptr, "gemm_mxfp8: both scale tensors are required");
  DSB_REQUIRE((reinterpret_cast<uintptr_t>(A) & 15) == 0 && (reinterpret_cast<uintptr_t>(B) & 15) == 0 && (reinterpret_cast<uintptr_t>(C) & 15) == 0 &&
                  (reinterpret_cast<uintptr_t>(sfa) & 15) == 0 && (reinterpret_cast<uintptr_t>(sfb) & 15) == 0,
              "gemm_mxfp8: operands and scale tensors must be 16-byte aligned");
  DSB_REQUIRE(lda % 16 == 0 && ldb % 16 == 0 && ldc % 8 == 0, "gemm_mxfp8: lda/ldb must be multiples of 16, ldc of 8");
  CUtensorMap tmA, tmB, tmC;
  int rc = make_tensor_map_2d(&tmA, TMA_U8, A, M, K, lda, gemm8mx::BM, gemm8mx::BKB, TMA_SW_128);
  if (rc) return rc;
  rc = make_tensor_map_2d(&tmB, TMA_U8, B, N, K, ldb, gemm8mx::BN, gemm8mx::BKB, TMA_SW_128);
  if (rc) return rc;
  rc = make_tensor_map_2d(&tmC, TMA_BF16, C, M, N, ldc, gemm8mx::BM, 64, TMA_SW_128);
  if (rc) return rc;
  static bool attr_set = false;
  if (!attr_set) {
    DSB_CHECK_CUDA(cudaFuncSetAttribute(gemm8mx::gemm_mxfp8_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, gemm8mx::SMEM_TOTAL));
    attr_set = true;
  }
  gemm8mx::Params p;
  p.M = M; p.N = N; p.K = K; p.KB = K / 128;
  p.tiles_m = (M + gemm8mx::BM - 1) / gemm8mx::BM;
  p.tiles_n = (N + gemm8mx::BN - 1) / gemm8mx::BN;
  p.sfa = static_cast<const uint8_t*>(sfa);
  p.sfb = static_cast<const uint8_t*>(sfb);
  int grid = p.tiles_m * p.tiles_n;
  int cap = num_sms();
  if (max_ctas > 0 && max_ctas < cap) cap = max_ctas;
  if (grid > cap) grid = cap;
  gemm8mx::gemm_mxfp8_kernel<<<grid, gemm8mx::NUM_THREADS, gemm8mx::SMEM_TOTAL, stream>>>(tmA, tmB, tmC, p);
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

// All-gather fused into the consuming GEMM: Y[M,N] = X[M,K] * W[N,K]^T where W is ZeRO-3 / TP sharded
// across ranks as a slice of one flat unit buffer.  ONE kernel: two copy warps per CTA pull W's row blocks from the
// owning peers over NVLink (ld.relaxed.sys on IPC-mapped pointers) into the local gathered buffer and publish
// per-block counters; the TMA producer of every CTA waits for the counter of the n-block it is about to consume, so
// the transfer of block j+1.. overlaps the tcgen05 math on block j.  The gathered copy stays in `b_local` for the
// backward pass.  Baseline it replaces: NCCL all_gather_into_tensor on a side stream + cuBLAS GEMM
// (reference partition_parameters.py:1064-1098 + zero/linear.py:43-112).
#include "gemm_2cta.cuh"

namespace dsb {
namespace gemm {

// CTA-pair core (256 x 256 pair tiles): used for the large shapes, where it is ~10% faster than the single-CTA core
static int launch_ag2(const void* A, void* b_local, void* C, int M, int N, int K, int lda, int ldc, void* const* peer_shards,
                      long long w_off, long long shard_numel, int* flags, int world, cudaStream_t stream) {
  constexpr int STAGES = 6;
  using S = Smem2<STAGES>;
  CUtensorMap tmA, tmB, tmC;
  int rc = make_tensor_map_2d(&tmA, TMA_BF16, A, M, K, lda, BM, BK, TMA_SW_128);
  if (rc) return rc;
  rc = make_tensor_map_2d(&tmB, TMA_BF16, b_local, N, K, K, BN2_CTA, BK, TMA_SW_128);
  if (rc) return rc;
  rc = make_tensor_map_2d(&tmC, TMA_BF16, C, M, N, ldc, BM, 64, TMA_SW_128);
  if (rc) return rc;
  auto kern = gemm2_ag_bf16_kernel<STAGES>;
  static bool attr_set = false;
  if (!attr_set) {
    DSB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, S::TOTAL));
    attr_set = true;
  }
  Params p;
  p.M = M; p.N = N; p.K = K;
  p.tiles_m = (M + BM2 - 1) / BM2;
  p.tiles_n = (N + BN2 - 1) / BN2;
  p.reduce_add = 0;
  AgArgs ag;
  for (int i = 0; i < AG_MAX_RANKS; ++i) ag.peer_shard[i] = i < world ? peer_shards[i] : nullptr;
  ag.w_off = w_off;
  ag.shard_numel = shard_numel;
  ag.b_local = static_cast<__nv_bfloat16*>(b_local);
  ag.flags = flags;
  ag.world = world;
  DSB_CHECK_CUDA(cudaMemsetAsync(flags, 0, sizeof(int) * (p.tiles_n + 1), stream));  // per-block counters + the claim counter
  int clusters = p.tiles_m * p.tiles_n;
  if (clusters > num_sms() / 2) clusters = num_sms() / 2;
  kern<<<2 * clusters, NUM_THREADS_AG, S::TOTAL, stream>>>(tmA, tmB, tmC, p, ag);
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

static int launch_ag(const void* A, void* b_local, void* C, int M, int N, int K, int lda, int ldc, void* const* peer_shards,
                     long long w_off, long long shard_numel, int* flags, int world, cudaStream_t stream) {
  constexpr int BN = 256, STAGES = 4;
  using S = Smem<BN, STAGES>;
  CUtensorMap tmA, tmB, tmC;
  int rc = make_tensor_map_2d(&tmA, TMA_BF16, A, M, K, lda, BM, BK, TMA_SW_128);
  if (rc) return rc;
  rc = make_tensor_map_2d(&tmB, TMA_BF16, b_local, N, K, K, BN, BK, TMA_SW_128);
  if (rc) return rc;
  rc = make_tensor_map_2d(&tmC, TMA_BF16, C, M, N, ldc, BM, 64, TMA_SW_128);
  if (rc) return rc;
  auto kern = gemm_ag_bf16_kernel<BN, STAGES>;
  static bool attr_set = false;
  if (!attr_set) {
    DSB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, S::TOTAL));
    attr_set = true;
  }
  Params p;
  p.M = M; p.N = N; p.K = K;
  p.tiles_m = (M + BM - 1) / BM;
  p.tiles_n = (N + BN - 1) / BN;
  p.reduce_add = 0;
  AgArgs ag;
  for (int i = 0; i < AG_MAX_RANKS; ++i) ag.peer_shard[i] = i < world ? peer_shards[i] : nullptr;
  ag.w_off = w_off;
  ag.shard_numel = shard_numel;
  ag.b_local = static_cast<__nv_bfloat16*>(b_local);
  ag.flags = flags;
  ag.world = world;
  DSB_CHECK_CUDA(cudaMemsetAsync(flags, 0, sizeof(int) * (p.tiles_n + 1), stream));  // per-block counters + the claim counter
  int grid = p.tiles_m * p.tiles_n;
  const int cap = num_sms();
  if (grid > cap) grid = cap;
  kern<<<grid, NUM_THREADS_AG, S::TOTAL, stream>>>(tmA, tmB, tmC, p, ag);
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

}  // namespace gemm
}  // namespace dsb

// C[M,N] (bf16) = A[M,K] * W^T with W[N,K] gathered from peers INSIDE the kernel.
//   b_local     : [N,K] bf16 destination of the gather (kept for backward); also the GEMM's B operand
//   peer_shards : HOST array of `world` device pointers -- each rank's shard region of the unit (this process's mappings)
//   w_off       : flat element offset of W[0][0] inside the unit; shard_numel: elements per rank shard
//   flags       : device scratch, >= ceil(N/256) + 1 ints (zeroed by this call): per n-block arrival counters + one work-claim counter
DSB_EXPORT int dsb200_gemm_ag_bf16(const void* A, void* b_local, void* C, int M, int N, int K, int lda, int ldc,
                                   void* const* peer_shards, long long w_off, long long shard_numel, int* flags, int world,
                                   cudaStream_t stream) {
  using namespace dsb;
  DSB_REQUIRE(M > 0 && N > 0 && K > 0, "gemm_ag: empty problem");
  DSB_REQUIRE(world >= 1 && world <= gemm::AG_MAX_RANKS, "gemm_ag: world %d unsupported", world);
  DSB_REQUIRE(K % 8 == 0 && w_off % 8 == 0 && shard_numel % 8 == 0, "gemm_ag: K, w_off and shard_numel must be multiples of 8");
  DSB_REQUIRE(lda % 8 == 0 && ldc % 8 == 0, "gemm_ag: lda/ldc must be multiples of 8");
  const long long tiles2 = static_cast<long long>((M + 255) / 256) * ((N + 255) / 256);
  if (M > 128 && N > 128 && tiles2 >= 37)
    return gemm::launch_ag2(A, b_local, C, M, N, K, lda, ldc, peer_shards, w_off, shard_numel, flags, world, stream);
  return gemm::launch_ag(A, b_local, C, M, N, K, lda, ldc, peer_shards, w_off, shard_numel, flags, world, stream);
}

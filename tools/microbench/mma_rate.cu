// Issue-rate probe for tcgen05.mma (cta_group::1, kind::f16, bf16 in / fp32 out): R back-to-back instructions of shape
// 128 x N x 16 on one SM, operands = zeroed shared memory (SS) or TMEM (TS), timed with clock64 around issue + commit + wait.
//   nvcc -gencode arch=compute_100a,code=sm_100a -std=c++17 -O3 -I csrc tools/microbench/mma_rate.cu -o tools/microbench/mma_rate
#include <cstdio>
#include <cuda_runtime.h>
#include "common/ptx.cuh"
using namespace dsb;

template <int N, bool TS, bool B_MN>
__global__ void __launch_bounds__(128, 1) probe(long long* out, int reps) {
  extern __shared__ uint8_t raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 96 * 1024);
  uint32_t* tptr = reinterpret_cast<uint32_t*>(bar + 2);
  for (int i = threadIdx.x; i < 96 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (threadIdx.x == 0) { mbar_init(bar, 1); fence_barrier_init(); }
  if (threadIdx.x < 32) { tmem_alloc<1>(tptr, 512); tmem_relinquish<1>(); }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = *tptr;
  if (threadIdx.x == 0) {
    constexpr uint32_t idesc = umma_idesc(128, N, 1, 1, 0, B_MN ? 1 : 0);
    const uint32_t a_s = smem_u32(smem), b_s = smem_u32(smem + 32 * 1024);
    const long long t0 = clock64();
    for (int r = 0; r < reps; ++r) {
      const int kk = r & 7;
      const uint32_t off = (kk >> 2) * 16384 + (kk & 3) * 32;
      const uint64_t bdesc = B_MN ? umma_smem_desc(b_s + kk * 2048, 16384, 1024) : umma_smem_desc(b_s + off, 16, 1024);
      if (TS) umma_f16_ts(tm, tm + 256 + kk * 8, bdesc, idesc, r != 0);
      else umma_f16<1>(tm, umma_smem_desc(a_s + off, 16, 1024), bdesc, idesc, r != 0);
    }
    const long long t1 = clock64();
    umma_commit<1>(bar);
    mbar_wait(bar, 0);
    const long long t2 = clock64();
    out[0] = t1 - t0;
    out[1] = t2 - t0;
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) { tc_fence_after(); tmem_dealloc<1>(tm, 512); }
}

template <int N, bool TS, bool B_MN>
void run(const char* name, long long* d) {
  auto k = probe<N, TS, B_MN>;
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  for (int reps : {8, 64, 512}) {
    long long h[2];
    k<<<1, 128, 100 * 1024>>>(d, reps);
    k<<<1, 128, 100 * 1024>>>(d, reps);
    cudaError_t e = cudaDeviceSynchronize();
    cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
    printf("%-34s reps %4d : issue %8lld cyc, done %8lld cyc  -> %6.1f cyc/MMA  (%s)\n", name, reps, h[0], h[1], double(h[1]) / reps,
           cudaGetErrorString(e));
  }
}

// Variant: the whole warp stays converged, one elected lane issues; descriptors = precomputed base + a 32-bit add per step.
template <int N, bool TS>
__global__ void __launch_bounds__(128, 1) probe_warp(long long* out, int reps) {
  extern __shared__ uint8_t raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 96 * 1024);
  uint32_t* tptr = reinterpret_cast<uint32_t*>(bar + 2);
  for (int i = threadIdx.x; i < 96 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (threadIdx.x == 0) { mbar_init(bar, 1); fence_barrier_init(); }
  if (threadIdx.x < 32) { tmem_alloc<1>(tptr, 512); tmem_relinquish<1>(); }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = __shfl_sync(0xffffffffu, *tptr, 0);
  if (threadIdx.x < 32) {
    const bool leader = elect_one();
    constexpr uint32_t idesc = umma_idesc(128, N, 1, 1, 0, 0);
    const uint64_t a0 = umma_smem_desc(smem_u32(smem), 16, 1024), b0 = umma_smem_desc(smem_u32(smem + 32 * 1024), 16, 1024);
    const long long t0 = clock64();
    for (int r = 0; r < reps; r += 8) {
#pragma unroll
      for (int kk = 0; kk < 8; ++kk) {
        const uint32_t off = ((kk >> 2) * 16384 + (kk & 3) * 32) >> 4;
        if (leader) {
          if (TS) umma_f16_ts(tm, tm + 256 + kk * 8, b0 + off, idesc, (r | kk) != 0);
          else umma_f16<1>(tm, a0 + off, b0 + off, idesc, (r | kk) != 0);
        }
      }
    }
    const long long t1 = clock64();
    if (leader) umma_commit<1>(bar);
    mbar_wait(bar, 0);
    const long long t2 = clock64();
    if (leader) { out[0] = t1 - t0; out[1] = t2 - t0; }
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) { tc_fence_after(); tmem_dealloc<1>(tm, 512); }
}

template <int N, bool TS>
void run_warp(const char* name, long long* d) {
  auto k = probe_warp<N, TS>;
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  for (int reps : {8, 64, 512}) {
    long long h[2];
    k<<<1, 128, 100 * 1024>>>(d, reps);
    k<<<1, 128, 100 * 1024>>>(d, reps);
    cudaError_t e = cudaDeviceSynchronize();
    cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
    printf("%-34s reps %4d : issue %8lld cyc, done %8lld cyc  -> %6.1f cyc/MMA  (%s)\n", name, reps, h[0], h[1], double(h[1]) / reps,
           cudaGetErrorString(e));
  }
}

int main() {
  long long* d;
  cudaMalloc(&d, 64);
  run<128, false, false>("SS 128x128x16 B K-major", d);
  run<128, true, false>("TS 128x128x16 B K-major", d);
  run<128, true, true>("TS 128x128x16 B MN-major", d);
  run<256, false, false>("SS 128x256x16 B K-major", d);
  run<256, true, false>("TS 128x256x16 B K-major", d);
  run<64, false, false>("SS 128x64x16 B K-major", d);
  run<16, true, false>("TS 128x16x16 B K-major", d);
  run_warp<128, false>("warp-converged SS 128x128x16", d);
  run_warp<128, true>("warp-converged TS 128x128x16", d);
  run_warp<64, true>("warp-converged TS 128x64x16", d);
  run_warp<16, true>("warp-converged TS 128x16x16", d);
  return 0;
}

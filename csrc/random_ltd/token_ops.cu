// Random layer-token-drop (random-LTD) data movement: sort the sampled token indices of every (layer, batch) row,
// gather the kept tokens into a dense [B, R, H] / [R, B, H] tensor and scatter the layer output back.
// Reference kernels: csrc/random_ltd/{token_sort.cu,gather_scatter.cu,pt_binding.cpp}.  Design here:
//   * sort    : indices are unique positions in [0, S) -> one CTA per row builds a bitmap of the row in shared memory
//               (S/32 words), block-scans the popcounts and emits the set bits in order: O(S/32 + R) per row,
//               no comparison sort, any S up to 1M positions.
//   * gather / scatter: pure row copies, dtype-agnostic (rows are moved as 16-byte vectors when the row size allows);
//               one warp per token row so a 4096 x bf16 row is 16 x 16 B per lane, fully coalesced.
#include "../common/host.h"

namespace dsb {
namespace ltd {

constexpr int SORT_THREADS = 256;

__global__ void token_sort_kernel(int* __restrict__ idx, int rows, int reserved, int seq) {
  extern __shared__ unsigned bitmap[];
  __shared__ int partial[SORT_THREADS];
  const int words = (seq + 31) / 32;
  for (int row = blockIdx.x; row < rows; row += gridDim.x) {
    int* r = idx + (long long)row * reserved;
    for (int w = threadIdx.x; w < words; w += blockDim.x) bitmap[w] = 0u;
    __syncthreads();
    for (int i = threadIdx.x; i < reserved; i += blockDim.x) {
      const int v = r[i];
      if (v >= 0 && v < seq) atomicOr(&bitmap[v >> 5], 1u << (v & 31));
    }
    __syncthreads();
    const int per = (words + blockDim.x - 1) / blockDim.x;
    const int w0 = threadIdx.x * per, w1 = min(words, w0 + per);
    int cnt = 0;
    for (int w = w0; w < w1; ++w) cnt += __popc(bitmap[w]);
    partial[threadIdx.x] = cnt;
    __syncthreads();
    // exclusive scan over the per-thread counts (256 entries: a serial pass by warp 0 lanes over 8 entries each + shuffle)
    if (threadIdx.x < 32) {
      int local[SORT_THREADS / 32];
      int s = 0;
#pragma unroll
      for (int k = 0; k < SORT_THREADS / 32; ++k) {
        local[k] = s;
        s += partial[threadIdx.x * (SORT_THREADS / 32) + k];
      }
      int incl = s;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, o);
        if ((int)threadIdx.x >= o) incl += t;
      }
      const int base = incl - s;
#pragma unroll
      for (int k = 0; k < SORT_THREADS / 32; ++k) partial[threadIdx.x * (SORT_THREADS / 32) + k] = base + local[k];
    }
    __syncthreads();
    int out = partial[threadIdx.x];
    for (int w = w0; w < w1; ++w) {
      unsigned bits = bitmap[w];
      while (bits) {
        const int b = __ffs(bits) - 1;
        bits &= bits - 1;
        if (out < reserved) r[out] = (w << 5) + b;
        ++out;
      }
    }
    __syncthreads();
  }
}

// one warp per (batch, kept-token) row.  GATHER: part[b, j] = all[b, idx[b, j]] ; else all[b, idx[b, j]] = part[b, j]
template <bool GATHER, typename V>
__global__ void token_move_kernel(char* __restrict__ all, char* __restrict__ part, const int* __restrict__ idx, int B,
                                  int S, int R, long long row_bytes, int batch_first) {
  const int lane = threadIdx.x & 31;
  const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
  const long long nvec = row_bytes / (long long)sizeof(V);
  for (long long t = warp; t < (long long)B * R; t += nwarps) {
    const int b = (int)(t / R), j = (int)(t % R);
    const int s = idx[(long long)b * R + j];
    const long long all_row = batch_first ? ((long long)b * S + s) : ((long long)s * B + b);
    const long long part_row = batch_first ? ((long long)b * R + j) : ((long long)j * B + b);
    V* a = reinterpret_cast<V*>(all + all_row * row_bytes);
    V* p = reinterpret_cast<V*>(part + part_row * row_bytes);
    for (long long v = lane; v < nvec; v += 32) {
      if (GATHER)
        p[v] = a[v];
      else
        a[v] = p[v];
    }
  }
}

template <bool GATHER>
static int launch_move(void* all, void* part, const void* idx, int B, int S, int R, long long row_bytes, int batch_first,
                       cudaStream_t stream) {
  const long long rows = (long long)B * R;
  const int threads = 256;
  long long want = (rows * 32 + threads - 1) / threads;
  const long long cap = (long long)num_sms() * 16;
  const int grid = (int)(want < cap ? (want > 0 ? want : 1) : cap);
  const bool vec16 = row_bytes % 16 == 0 && ((uintptr_t)all & 15) == 0 && ((uintptr_t)part & 15) == 0;
  if (vec16)
    token_move_kernel<GATHER, uint4><<<grid, threads, 0, stream>>>((char*)all, (char*)part, (const int*)idx, B, S, R,
                                                                  row_bytes, batch_first);
  else if (row_bytes % 2 == 0)
    token_move_kernel<GATHER, unsigned short><<<grid, threads, 0, stream>>>((char*)all, (char*)part, (const int*)idx, B,
                                                                           S, R, row_bytes, batch_first);
  else
    token_move_kernel<GATHER, unsigned char><<<grid, threads, 0, stream>>>((char*)all, (char*)part, (const int*)idx, B, S,
                                                                          R, row_bytes, batch_first);
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

}  // namespace ltd
}  // namespace dsb

// idx: int32 [rows, reserved] of unique positions in [0, seq); sorted ascending in place, row by row.
DSB_EXPORT int dsb200_token_sort(void* idx, int rows, int reserved, int seq, cudaStream_t stream) {
  using namespace dsb;
  DSB_REQUIRE(rows > 0 && reserved > 0 && seq >= reserved, "token_sort: rows=%d reserved=%d seq=%d", rows, reserved, seq);
  const size_t smem = (size_t)((seq + 31) / 32) * sizeof(unsigned);
  DSB_REQUIRE(smem <= 200 * 1024, "token_sort: seq=%d too long for the shared-memory bitmap", seq);
  static bool attr_set = false;
  if (!attr_set) {
    DSB_CHECK_CUDA(cudaFuncSetAttribute(ltd::token_sort_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    attr_set = true;
  }
  const int grid = rows < num_sms() * 4 ? rows : num_sms() * 4;
  ltd::token_sort_kernel<<<grid, ltd::SORT_THREADS, smem, stream>>>(static_cast<int*>(idx), rows, reserved, seq);
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

// part[B,R,row] (batch_first) or part[R,B,row]  <-  all[B,S,row] / all[S,B,row] at idx[B,R]
DSB_EXPORT int dsb200_token_gather(const void* all, void* part, const void* idx, int B, int S, int R, long long row_bytes,
                                   int batch_first, cudaStream_t stream) {
  DSB_REQUIRE(B > 0 && S > 0 && R > 0 && row_bytes > 0, "token_gather: bad shape");
  return dsb::ltd::launch_move<true>(const_cast<void*>(all), part, idx, B, S, R, row_bytes, batch_first, stream);
}

DSB_EXPORT int dsb200_token_scatter(void* all, const void* part, const void* idx, int B, int S, int R, long long row_bytes,
                                    int batch_first, cudaStream_t stream) {
  DSB_REQUIRE(B > 0 && S > 0 && R > 0 && row_bytes > 0, "token_scatter: bad shape");
  return dsb::ltd::launch_move<false>(all, const_cast<void*>(part), idx, B, S, R, row_bytes, batch_first, stream);
}

// Evoformer attention (AlphaFold / OpenFold MSA row / column and triangle attention): softmax(Q K^T * scale + bias1 + bias2) V with
//   Q, K, V, O : [B][N][L][H][D]      bias1 : [B][N][1][1][L] (per key)      bias2 : [B][1][H][L][L] (pair bias, shared over N)
// and gradients for all five inputs.  Reference counterpart: csrc/deepspeed4science/evoformer_attn (a CUTLASS fused-MHA fork).
//
// Head sizes are small here (D <= 64, typically 32) and L is a few hundred: the op is bound by the exponentials and the bias
// traffic, not by matrix throughput, so these are CUDA-core kernels organised for reuse rather than tensor-core kernels:
//   forward        block = (b, n, h, 16 query rows): 32-key tiles of K (transposed) and V staged in shared memory and shared by the 4
//                  warps; a warp owns 4 rows, lane == key for the scores, lane == output feature for P V; online softmax; LSE saved.
//   backward rows  same tiling: recomputes P from LSE, dP = dO V^T, dS = P (dP - delta); dQ in registers; dbias2 += dS by atomics
//                  (N-way sum), dbias1 through a per-block shared-memory partial; delta saved for the column pass.
//   backward cols  block = (b, n, h, 16 keys), lane == query row inside 32-row tiles of Q and dO: dK, dV in registers, written once.
// fp32 arithmetic throughout; inputs / outputs bf16 or fp16.
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include "../common/host.h"

namespace dsb {
namespace evo {

constexpr int TILE = 32;        // keys (or rows) per shared-memory tile == lanes
constexpr int WARPS = 4;
constexpr int PER_WARP = 4;     // rows (forward / backward rows) or keys (backward cols) per warp
constexpr int PER_BLOCK = WARPS * PER_WARP;
constexpr int DMAX = 64;

template <typename T> __device__ __forceinline__ float to_f(T v);
template <> __device__ __forceinline__ float to_f<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <> __device__ __forceinline__ float to_f<__half>(__half v) { return __half2float(v); }
template <typename T> __device__ __forceinline__ T from_f(float v);
template <> __device__ __forceinline__ __nv_bfloat16 from_f<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }
template <> __device__ __forceinline__ __half from_f<__half>(float v) { return __float2half_rn(v); }

__device__ __forceinline__ float wsum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float wmax(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

struct Shape {
  int B, N, L, H, D;
  float scale;
};
// element (b, n, l, h, d) of a [B][N][L][H][D] tensor
__device__ __forceinline__ size_t at(const Shape& s, int b, int n, int l, int h) {
  return (((static_cast<size_t>(b) * s.N + n) * s.L + l) * s.H + h) * s.D;
}

// stage a 32-row tile of X (rows r0.., head h) as transposed floats Xt[d][33] and / or row-major Xs[32][64]; rows >= L are zero
template <typename T>
__device__ __forceinline__ void load_tile(const T* __restrict__ x, const Shape& s, int b, int n, int h, int r0, float (*xt)[TILE + 1],
                                          float (*xs)[DMAX]) {
  for (int e = threadIdx.x; e < TILE * s.D; e += blockDim.x) {
    const int r = e / s.D, d = e - r * s.D;
    const float v = (r0 + r < s.L) ? to_f(x[at(s, b, n, r0 + r, h) + d]) : 0.f;
    if (xt) xt[d][r] = v;
    if (xs) xs[r][d] = v;
  }
}

// ------------------------------------------------------------------------------------------------ forward
template <typename T>
__global__ void __launch_bounds__(WARPS * 32)
evo_fwd_kernel(const T* __restrict__ q, const T* __restrict__ k, const T* __restrict__ v, const T* __restrict__ bias1,
               const T* __restrict__ bias2, T* __restrict__ o, float* __restrict__ lse, Shape s) {
  __shared__ float kt[DMAX][TILE + 1];
  __shared__ float vs[TILE][DMAX];
  __shared__ float qs[PER_BLOCK][DMAX];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int bnh = blockIdx.x;
  const int h = bnh % s.H; bnh /= s.H;
  const int n = bnh % s.N;
  const int b = bnh / s.N;
  const int row0 = blockIdx.y * PER_BLOCK;
  for (int e = threadIdx.x; e < PER_BLOCK * s.D; e += blockDim.x) {
    const int r = e / s.D, d = e - r * s.D;
    qs[r][d] = (row0 + r < s.L) ? to_f(q[at(s, b, n, row0 + r, h) + d]) * s.scale : 0.f;
  }
  float m[PER_WARP], l[PER_WARP], acc0[PER_WARP], acc1[PER_WARP];
#pragma unroll
  for (int r = 0; r < PER_WARP; ++r) { m[r] = -INFINITY; l[r] = 0.f; acc0[r] = 0.f; acc1[r] = 0.f; }
  for (int j0 = 0; j0 < s.L; j0 += TILE) {
    __syncthreads();  // previous tile consumed (and qs written, first round)
    load_tile(k, s, b, n, h, j0, kt, nullptr);
    load_tile(v, s, b, n, h, j0, nullptr, vs);
    __syncthreads();
    const int j = j0 + lane;
    const float b1 = (bias1 != nullptr && j < s.L) ? to_f(bias1[(static_cast<size_t>(b) * s.N + n) * s.L + j]) : 0.f;
#pragma unroll
    for (int r = 0; r < PER_WARP; ++r) {
      const int i = row0 + warp * PER_WARP + r;
      if (i >= s.L) continue;  // warp-uniform
      const float* qr = qs[warp * PER_WARP + r];
      float sc = 0.f;
      for (int d = 0; d < s.D; ++d) sc = fmaf(qr[d], kt[d][lane], sc);
      sc += b1;
      if (bias2 != nullptr && j < s.L) sc += to_f(bias2[((static_cast<size_t>(b) * s.H + h) * s.L + i) * s.L + j]);
      if (j >= s.L) sc = -INFINITY;
      const float m_new = fmaxf(m[r], wmax(sc));
      const float alpha = (m[r] == -INFINITY) ? 0.f : __expf(m[r] - m_new);
      const float p = (m_new == -INFINITY) ? 0.f : __expf(sc - m_new);
      l[r] = l[r] * alpha + wsum(p);
      m[r] = m_new;
      float a0 = acc0[r] * alpha, a1 = acc1[r] * alpha;
      for (int jj = 0; jj < TILE; ++jj) {
        const float pj = __shfl_sync(0xffffffffu, p, jj);
        a0 = fmaf(pj, vs[jj][lane], a0);
        if (s.D > 32) a1 = fmaf(pj, vs[jj][lane + 32], a1);
      }
      acc0[r] = a0;
      acc1[r] = a1;
    }
  }
#pragma unroll
  for (int r = 0; r < PER_WARP; ++r) {
    const int i = row0 + warp * PER_WARP + r;
    if (i >= s.L) continue;
    const float inv = l[r] > 0.f ? 1.f / l[r] : 0.f;
    T* orow = o + at(s, b, n, i, h);
    if (lane < s.D) orow[lane] = from_f<T>(acc0[r] * inv);
    if (lane + 32 < s.D) orow[lane + 32] = from_f<T>(acc1[r] * inv);
    if (lane == 0) lse[((static_cast<size_t>(b) * s.N + n) * s.H + h) * s.L + i] = m[r] + __logf(fmaxf(l[r], 1e-30f));
  }
}

// ------------------------------------------------------------------------------------------------ backward, row pass
template <typename T>
__global__ void __launch_bounds__(WARPS * 32)
evo_bwd_rows_kernel(const T* __restrict__ q, const T* __restrict__ k, const T* __restrict__ v, const T* __restrict__ bias1,
                    const T* __restrict__ bias2, const T* __restrict__ o, const T* __restrict__ d_o, const float* __restrict__ lse,
                    T* __restrict__ dq, float* __restrict__ delta, float* __restrict__ dbias1, float* __restrict__ dbias2, Shape s) {
  __shared__ float kt[DMAX][TILE + 1];
  __shared__ float ks[TILE][DMAX];
  __shared__ float vt[DMAX][TILE + 1];
  __shared__ float qs[PER_BLOCK][DMAX];
  __shared__ float dos[PER_BLOCK][DMAX];
  __shared__ float db1[TILE];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int bnh = blockIdx.x;
  const int h = bnh % s.H; bnh /= s.H;
  const int n = bnh % s.N;
  const int b = bnh / s.N;
  const int row0 = blockIdx.y * PER_BLOCK;
  for (int e = threadIdx.x; e < PER_BLOCK * s.D; e += blockDim.x) {
    const int r = e / s.D, d = e - r * s.D;
    const bool ok = row0 + r < s.L;
    qs[r][d] = ok ? to_f(q[at(s, b, n, row0 + r, h) + d]) * s.scale : 0.f;
    dos[r][d] = ok ? to_f(d_o[at(s, b, n, row0 + r, h) + d]) : 0.f;
  }
  __syncthreads();
  float lse_r[PER_WARP], del_r[PER_WARP], dq0[PER_WARP], dq1[PER_WARP];
#pragma unroll
  for (int r = 0; r < PER_WARP; ++r) {
    const int i = row0 + warp * PER_WARP + r;
    dq0[r] = dq1[r] = 0.f;
    lse_r[r] = del_r[r] = 0.f;
    if (i >= s.L) continue;
    const size_t li = ((static_cast<size_t>(b) * s.N + n) * s.H + h) * s.L + i;
    lse_r[r] = lse[li];
    const T* orow = o + at(s, b, n, i, h);
    float part = 0.f;
    if (lane < s.D) part = to_f(orow[lane]) * dos[warp * PER_WARP + r][lane];
    if (lane + 32 < s.D) part += to_f(orow[lane + 32]) * dos[warp * PER_WARP + r][lane + 32];
    del_r[r] = wsum(part);
    if (lane == 0) delta[li] = del_r[r];
  }
  for (int j0 = 0; j0 < s.L; j0 += TILE) {
    __syncthreads();
    load_tile(k, s, b, n, h, j0, kt, ks);
    load_tile(v, s, b, n, h, j0, vt, nullptr);
    if (threadIdx.x < TILE) db1[threadIdx.x] = 0.f;
    __syncthreads();
    const int j = j0 + lane;
    const float b1 = (bias1 != nullptr && j < s.L) ? to_f(bias1[(static_cast<size_t>(b) * s.N + n) * s.L + j]) : 0.f;
    float db1_lane = 0.f;
#pragma unroll
    for (int r = 0; r < PER_WARP; ++r) {
      const int i = row0 + warp * PER_WARP + r;
      if (i >= s.L) continue;
      const float* qr = qs[warp * PER_WARP + r];
      const float* dr = dos[warp * PER_WARP + r];
      float sc = 0.f, dp = 0.f;
      for (int d = 0; d < s.D; ++d) {
        sc = fmaf(qr[d], kt[d][lane], sc);
        dp = fmaf(dr[d], vt[d][lane], dp);
      }
      sc += b1;
      const size_t b2i = ((static_cast<size_t>(b) * s.H + h) * s.L + i) * s.L + j;
      if (bias2 != nullptr && j < s.L) sc += to_f(bias2[b2i]);
      const float p = (j < s.L) ? __expf(sc - lse_r[r]) : 0.f;
      const float ds = p * (dp - del_r[r]);
      if (j < s.L) {
        if (dbias2 != nullptr) atomicAdd(dbias2 + b2i, ds);
        db1_lane += ds;
      }
      float a0 = dq0[r], a1 = dq1[r];
      for (int jj = 0; jj < TILE; ++jj) {
        const float dsj = __shfl_sync(0xffffffffu, ds, jj);
        a0 = fmaf(dsj, ks[jj][lane], a0);
        if (s.D > 32) a1 = fmaf(dsj, ks[jj][lane + 32], a1);
      }
      dq0[r] = a0;
      dq1[r] = a1;
    }
    if (dbias1 != nullptr) {
      atomicAdd(&db1[lane], db1_lane);
      __syncthreads();
      if (threadIdx.x < TILE && j0 + threadIdx.x < s.L)
        atomicAdd(dbias1 + (static_cast<size_t>(b) * s.N + n) * s.L + j0 + threadIdx.x, db1[threadIdx.x]);
    }
  }
#pragma unroll
  for (int r = 0; r < PER_WARP; ++r) {
    const int i = row0 + warp * PER_WARP + r;
    if (i >= s.L) continue;
    T* drow = dq + at(s, b, n, i, h);
    if (lane < s.D) drow[lane] = from_f<T>(dq0[r] * s.scale);
    if (lane + 32 < s.D) drow[lane + 32] = from_f<T>(dq1[r] * s.scale);
  }
}

// ------------------------------------------------------------------------------------------------ backward, column pass
template <typename T>
__global__ void __launch_bounds__(WARPS * 32)
evo_bwd_cols_kernel(const T* __restrict__ q, const T* __restrict__ k, const T* __restrict__ v, const T* __restrict__ bias1,
                    const T* __restrict__ bias2, const T* __restrict__ d_o, const float* __restrict__ lse, const float* __restrict__ delta,
                    T* __restrict__ dk, T* __restrict__ dv, Shape s) {
  __shared__ float qt[DMAX][TILE + 1];
  __shared__ float qrow[TILE][DMAX];
  __shared__ float dot[DMAX][TILE + 1];
  __shared__ float dorow[TILE][DMAX];
  __shared__ float kk[PER_BLOCK][DMAX];
  __shared__ float vv[PER_BLOCK][DMAX];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int bnh = blockIdx.x;
  const int h = bnh % s.H; bnh /= s.H;
  const int n = bnh % s.N;
  const int b = bnh / s.N;
  const int key0 = blockIdx.y * PER_BLOCK;
  for (int e = threadIdx.x; e < PER_BLOCK * s.D; e += blockDim.x) {
    const int r = e / s.D, d = e - r * s.D;
    const bool ok = key0 + r < s.L;
    kk[r][d] = ok ? to_f(k[at(s, b, n, key0 + r, h) + d]) : 0.f;
    vv[r][d] = ok ? to_f(v[at(s, b, n, key0 + r, h) + d]) : 0.f;
  }
  float dk0[PER_WARP], dk1[PER_WARP], dv0[PER_WARP], dv1[PER_WARP], b1[PER_WARP];
#pragma unroll
  for (int r = 0; r < PER_WARP; ++r) {
    dk0[r] = dk1[r] = dv0[r] = dv1[r] = 0.f;
    const int j = key0 + warp * PER_WARP + r;
    b1[r] = (bias1 != nullptr && j < s.L) ? to_f(bias1[(static_cast<size_t>(b) * s.N + n) * s.L + j]) : 0.f;
  }
  for (int i0 = 0; i0 < s.L; i0 += TILE) {
    __syncthreads();
    load_tile(q, s, b, n, h, i0, qt, qrow);
    load_tile(d_o, s, b, n, h, i0, dot, dorow);
    __syncthreads();
    const int i = i0 + lane;  // lane == query row of the tile
    const size_t li = ((static_cast<size_t>(b) * s.N + n) * s.H + h) * s.L + min(i, s.L - 1);
    const float lse_i = lse[li], del_i = delta[li];
#pragma unroll
    for (int r = 0; r < PER_WARP; ++r) {
      const int j = key0 + warp * PER_WARP + r;
      if (j >= s.L) continue;
      const float* kr = kk[warp * PER_WARP + r];
      const float* vr = vv[warp * PER_WARP + r];
      float sc = 0.f, dp = 0.f;
      for (int d = 0; d < s.D; ++d) {
        sc = fmaf(kr[d], qt[d][lane], sc);
        dp = fmaf(vr[d], dot[d][lane], dp);
      }
      sc = sc * s.scale + b1[r];
      if (bias2 != nullptr && i < s.L) sc += to_f(bias2[((static_cast<size_t>(b) * s.H + h) * s.L + i) * s.L + j]);
      const float p = (i < s.L) ? __expf(sc - lse_i) : 0.f;
      const float ds = p * (dp - del_i);
      float a0 = dk0[r], a1 = dk1[r], c0 = dv0[r], c1 = dv1[r];
      for (int ii = 0; ii < TILE; ++ii) {
        const float pi = __shfl_sync(0xffffffffu, p, ii);
        const float dsi = __shfl_sync(0xffffffffu, ds, ii);
        a0 = fmaf(dsi, qrow[ii][lane], a0);
        c0 = fmaf(pi, dorow[ii][lane], c0);
        if (s.D > 32) {
          a1 = fmaf(dsi, qrow[ii][lane + 32], a1);
          c1 = fmaf(pi, dorow[ii][lane + 32], c1);
        }
      }
      dk0[r] = a0; dk1[r] = a1; dv0[r] = c0; dv1[r] = c1;
    }
  }
#pragma unroll
  for (int r = 0; r < PER_WARP; ++r) {
    const int j = key0 + warp * PER_WARP + r;
    if (j >= s.L) continue;
    T* dkr = dk + at(s, b, n, j, h);
    T* dvr = dv + at(s, b, n, j, h);
    if (lane < s.D) { dkr[lane] = from_f<T>(dk0[r] * s.scale); dvr[lane] = from_f<T>(dv0[r]); }
    if (lane + 32 < s.D) { dkr[lane + 32] = from_f<T>(dk1[r] * s.scale); dvr[lane + 32] = from_f<T>(dv1[r]); }
  }
}

template <typename T>
static int run_fwd(const void* q, const void* k, const void* v, const void* b1, const void* b2, void* o, float* lse, const Shape& s,
                   cudaStream_t st) {
  dim3 grid(static_cast<unsigned>(s.B * s.N * s.H), static_cast<unsigned>((s.L + PER_BLOCK - 1) / PER_BLOCK));
  evo_fwd_kernel<T><<<grid, WARPS * 32, 0, st>>>(static_cast<const T*>(q), static_cast<const T*>(k), static_cast<const T*>(v),
                                                 static_cast<const T*>(b1), static_cast<const T*>(b2), static_cast<T*>(o), lse, s);
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

template <typename T>
static int run_bwd(const void* q, const void* k, const void* v, const void* b1, const void* b2, const void* o, const void* d_o, const float* lse,
                   void* dq, void* dk, void* dv, float* delta, float* db1, float* db2, const Shape& s, cudaStream_t st) {
  dim3 grid(static_cast<unsigned>(s.B * s.N * s.H), static_cast<unsigned>((s.L + PER_BLOCK - 1) / PER_BLOCK));
  evo_bwd_rows_kernel<T><<<grid, WARPS * 32, 0, st>>>(static_cast<const T*>(q), static_cast<const T*>(k), static_cast<const T*>(v),
                                                      static_cast<const T*>(b1), static_cast<const T*>(b2), static_cast<const T*>(o),
                                                      static_cast<const T*>(d_o), lse, static_cast<T*>(dq), delta, db1, db2, s);
  DSB_CHECK_CUDA(cudaGetLastError());
  evo_bwd_cols_kernel<T><<<grid, WARPS * 32, 0, st>>>(static_cast<const T*>(q), static_cast<const T*>(k), static_cast<const T*>(v),
                                                      static_cast<const T*>(b1), static_cast<const T*>(b2), static_cast<const T*>(d_o), lse, delta,
                                                      static_cast<T*>(dk), static_cast<T*>(dv), s);
  DSB_CHECK_CUDA(cudaGetLastError());
  return DSB_OK;
}

}  // namespace evo
}  // namespace dsb

// dtype: 0 bf16, 1 fp16.  q, k, v, o: [B][N][L][H][D] contiguous, D <= 64; bias1 [B][N][L] or NULL; bias2 [B][H][L][L] or NULL (same dtype);
// lse: [B][N][H][L] fp32 (written).
DSB_EXPORT int dsb200_evoformer_attn_fwd(const void* q, const void* k, const void* v, const void* bias1, const void* bias2, void* o, float* lse,
                                         int B, int N, int L, int H, int D, float scale, int dtype, cudaStream_t stream) {
  using namespace dsb;
  DSB_REQUIRE(D >= 1 && D <= evo::DMAX && B > 0 && N > 0 && L > 0 && H > 0, "evoformer_attn: head size %d unsupported (<= 64)", D);
  DSB_REQUIRE(static_cast<long long>(B) * N * H <= 2147483647LL, "evoformer_attn: B * N * H too large for one grid");
  evo::Shape s{B, N, L, H, D, scale};
  if (dtype == 0) return evo::run_fwd<__nv_bfloat16>(q, k, v, bias1, bias2, o, lse, s, stream);
  if (dtype == 1) return evo::run_fwd<__half>(q, k, v, bias1, bias2, o, lse, s, stream);
  set_last_error("evoformer_attn: unsupported dtype %d", dtype);
  return DSB_ERR_UNSUPPORTED;
}

// Gradients: dq, dk, dv as q; dbias1 [B][N][L] / dbias2 [B][H][L][L] fp32, ZERO-INITIALISED by the caller (NULL = not wanted);
// delta: [B][N][H][L] fp32 scratch.
DSB_EXPORT int dsb200_evoformer_attn_bwd(const void* q, const void* k, const void* v, const void* bias1, const void* bias2, const void* o,
                                         const void* d_o, const float* lse, void* dq, void* dk, void* dv, float* delta, float* dbias1,
                                         float* dbias2, int B, int N, int L, int H, int D, float scale, int dtype, cudaStream_t stream) {
  using namespace dsb;
  DSB_REQUIRE(D >= 1 && D <= evo::DMAX && B > 0 && N > 0 && L > 0 && H > 0, "evoformer_attn: head size %d unsupported (<= 64)", D);
  evo::Shape s{B, N, L, H, D, scale};
  if (dtype == 0) return evo::run_bwd<__nv_bfloat16>(q, k, v, bias1, bias2, o, d_o, lse, dq, dk, dv, delta, dbias1, dbias2, s, stream);
  if (dtype == 1) return evo::run_bwd<__half>(q, k, v, bias1, bias2, o, d_o, lse, dq, dk, dv, delta, dbias1, dbias2, s, stream);
  set_last_error("evoformer_attn: unsupported dtype %d", dtype);
  return DSB_ERR_UNSUPPORTED;
}

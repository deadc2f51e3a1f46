// Host optimizers for ZeRO-Offload: Adam/AdamW, Adagrad, Lion on fp32 master weights in (pinned) host memory,
// AVX-512 / AVX2 / scalar selected at compile time (-march=native), OpenMP over tiles, with an optional
// low-precision (bf16/fp16) copy of the updated parameters written in the same pass so the H2D transfer of the
// 16-bit weights can start tile by tile on a side stream.
// Reference counterparts: csrc/adam/cpu_adam_impl.cpp:24-112,226-304 + csrc/includes/cpu_adam.h, simd.h;
// csrc/adagrad/cpu_adagrad.cpp; csrc/lion/cpu_lion_impl.cpp.
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>

#include <immintrin.h>
// Built for x86-64-v3 (AVX2/FMA/F16C); the AVX-512 inner loop is a target-attributed function chosen at run time,
// so one binary runs on the build box and on the GPU host whatever their CPUs.
#define DSBH_AVX512 __attribute__((target("avx512f,avx512bw,avx512vl,avx512dq")))
static inline bool has_avx512() {
  static const bool ok = __builtin_cpu_supports("avx512f") && __builtin_cpu_supports("avx512bw") &&
                         __builtin_cpu_supports("avx512vl");
  return ok;
}

#define DSBH_EXPORT extern "C" __attribute__((visibility("default")))

static inline uint16_t f32_to_bf16(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  if ((u & 0x7fffffffu) > 0x7f800000u) return 0x7fc0;  // NaN
  u += 0x7fffu + ((u >> 16) & 1u);                      // round to nearest even
  return static_cast<uint16_t>(u >> 16);
}
static inline float bf16_to_f32(uint16_t h) {
  uint32_t u = static_cast<uint32_t>(h) << 16;
  float f;
  memcpy(&f, &u, 4);
  return f;
}
static inline uint16_t f32_to_f16(float f) {
#if defined(__F16C__)
  return static_cast<uint16_t>(_cvtss_sh(f, _MM_FROUND_TO_NEAREST_INT));
#else
  uint32_t x;
  memcpy(&x, &f, 4);
  uint32_t sign = (x >> 16) & 0x8000u;
  int32_t e = static_cast<int32_t>((x >> 23) & 0xff) - 127 + 15;
  uint32_t m = x & 0x7fffffu;
  if (e <= 0) return static_cast<uint16_t>(sign);
  if (e >= 31) return static_cast<uint16_t>(sign | 0x7c00u);
  uint32_t r = sign | (static_cast<uint32_t>(e) << 10) | (m >> 13);
  if ((m & 0x1000u) && ((m & 0x2fffu) != 0)) r += 1;
  return static_cast<uint16_t>(r);
#endif
}
static inline float f16_to_f32(uint16_t h) {
#if defined(__F16C__)
  return _cvtsh_ss(h);
#else
  uint32_t sign = (h & 0x8000u) << 16, e = (h >> 10) & 0x1f, m = h & 0x3ffu, u;
  if (e == 0) u = sign; else if (e == 31) u = sign | 0x7f800000u | (m << 13); else u = sign | ((e + 112) << 23) | (m << 13);
  float f;
  memcpy(&f, &u, 4);
  return f;
#endif
}

// grad_type: 0 fp32, 1 bf16, 2 fp16 ; lp_type: 0 none, 1 bf16, 2 fp16
static inline float load_grad(const void* g, int grad_type, int64_t i) {
  if (grad_type == 0) return static_cast<const float*>(g)[i];
  if (grad_type == 1) return bf16_to_f32(static_cast<const uint16_t*>(g)[i]);
  return f16_to_f32(static_cast<const uint16_t*>(g)[i]);
}
static inline void store_lp(void* lp, int lp_type, int64_t i, float v) {
  if (lp_type == 1) static_cast<uint16_t*>(lp)[i] = f32_to_bf16(v);
  else if (lp_type == 2) static_cast<uint16_t*>(lp)[i] = f32_to_f16(v);
}

static const int64_t TILE = 1 << 16;

DSBH_AVX512 static int64_t adam_tile_avx512(float* p, float* m, float* v, const float* gf, void* lp, int lp_type, int64_t lo,
                                            int64_t hi, float beta1, float beta2, float eps, float step_size, float bc2_rsqrt,
                                            float weight_decay, float grad_scale, float lr, int adamw_mode) {
  const __m512 vb1 = _mm512_set1_ps(beta1), vb2 = _mm512_set1_ps(beta2), v1mb1 = _mm512_set1_ps(1.f - beta1),
               v1mb2 = _mm512_set1_ps(1.f - beta2), veps = _mm512_set1_ps(eps), vss = _mm512_set1_ps(step_size),
               vbc2 = _mm512_set1_ps(bc2_rsqrt), vwd = _mm512_set1_ps(weight_decay), vgs = _mm512_set1_ps(grad_scale),
               vlrwd = _mm512_set1_ps(lr * weight_decay);
  int64_t i = lo;
  for (; i + 16 <= hi; i += 16) {
    __m512 gp = _mm512_mul_ps(_mm512_loadu_ps(gf + i), vgs);
    __m512 pp = _mm512_loadu_ps(p + i), mm = _mm512_loadu_ps(m + i), vv = _mm512_loadu_ps(v + i);
    if (!adamw_mode) gp = _mm512_fmadd_ps(vwd, pp, gp);
    mm = _mm512_fmadd_ps(vb1, mm, _mm512_mul_ps(v1mb1, gp));
    vv = _mm512_fmadd_ps(vb2, vv, _mm512_mul_ps(v1mb2, _mm512_mul_ps(gp, gp)));
    __m512 denom = _mm512_add_ps(_mm512_mul_ps(_mm512_sqrt_ps(vv), vbc2), veps);
    __m512 upd = _mm512_div_ps(mm, denom);
    if (adamw_mode) pp = _mm512_fnmadd_ps(vlrwd, pp, pp);
    pp = _mm512_fnmadd_ps(vss, upd, pp);
    _mm512_storeu_ps(p + i, pp);
    _mm512_storeu_ps(m + i, mm);
    _mm512_storeu_ps(v + i, vv);
    if (lp_type == 1) {  // bf16, round to nearest even
      __m512i u = _mm512_castps_si512(pp);
      __m512i r = _mm512_add_epi32(u, _mm512_add_epi32(_mm512_set1_epi32(0x7fff),
                                                       _mm512_and_si512(_mm512_srli_epi32(u, 16), _mm512_set1_epi32(1))));
      __m256i h = _mm512_cvtepi32_epi16(_mm512_srli_epi32(r, 16));
      _mm256_storeu_si256(reinterpret_cast<__m256i*>(static_cast<uint16_t*>(lp) + i), h);
    } else if (lp_type == 2) {
      __m256i h = _mm512_cvtps_ph(pp, _MM_FROUND_TO_NEAREST_INT);
      _mm256_storeu_si256(reinterpret_cast<__m256i*>(static_cast<uint16_t*>(lp) + i), h);
    }
  }
  return i;
}


DSBH_EXPORT int dsbh_simd_width() { return has_avx512() ? 16 : 8; }

DSBH_EXPORT int dsbh_adam_step(float* p, float* m, float* v, const void* g, int grad_type, void* lp, int lp_type, int64_t n,
                               float lr, float beta1, float beta2, float eps, float weight_decay, int adamw_mode,
                               int bias_correction, int step, float grad_scale) {
  const float bc1 = bias_correction ? 1.f - powf(beta1, (float)step) : 1.f;
  const float bc2 = bias_correction ? 1.f - powf(beta2, (float)step) : 1.f;
  const float step_size = lr / bc1;
  const float bc2_rsqrt = 1.f / sqrtf(bc2);
  const float one_m_b1 = 1.f - beta1, one_m_b2 = 1.f - beta2;
  const int64_t ntiles = (n + TILE - 1) / TILE;
#pragma omp parallel for schedule(static)
  for (int64_t t = 0; t < ntiles; ++t) {
    const int64_t lo = t * TILE, hi = std::min(n, lo + TILE);
    int64_t i = lo;
    if (grad_type == 0 && has_avx512())
      i = adam_tile_avx512(p, m, v, static_cast<const float*>(g), lp, lp_type, lo, hi, beta1, beta2, eps, step_size, bc2_rsqrt,
                           weight_decay, grad_scale, lr, adamw_mode);
    for (; i < hi; ++i) {
      float gp = load_grad(g, grad_type, i) * grad_scale;
      float pp = p[i];
      if (!adamw_mode) gp += weight_decay * pp;
      float mm = beta1 * m[i] + one_m_b1 * gp;
      float vv = beta2 * v[i] + one_m_b2 * gp * gp;
      float denom = sqrtf(vv) * bc2_rsqrt + eps;
      if (adamw_mode) pp -= lr * weight_decay * pp;
      pp -= step_size * (mm / denom);
      p[i] = pp; m[i] = mm; v[i] = vv;
      store_lp(lp, lp_type, i, pp);
    }
  }
  return 0;
}

DSBH_EXPORT int dsbh_adagrad_step(float* p, float* h, const void* g, int grad_type, void* lp, int lp_type, int64_t n, float lr,
                                  float eps, float weight_decay, float grad_scale) {
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < n; ++i) {
    float gp = load_grad(g, grad_type, i) * grad_scale;
    float pp = p[i];
    gp += weight_decay * pp;
    float hh = h[i] + gp * gp;
    pp -= lr * gp / (sqrtf(hh) + eps);
    h[i] = hh; p[i] = pp;
    store_lp(lp, lp_type, i, pp);
  }
  return 0;
}

DSBH_EXPORT int dsbh_lion_step(float* p, float* m, const void* g, int grad_type, void* lp, int lp_type, int64_t n, float lr,
                               float beta1, float beta2, float weight_decay, float grad_scale) {
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < n; ++i) {
    const float gp = load_grad(g, grad_type, i) * grad_scale;
    float pp = p[i], mm = m[i];
    const float c = beta1 * mm + (1.f - beta1) * gp;
    const float s = c > 0.f ? 1.f : (c < 0.f ? -1.f : 0.f);
    pp = pp * (1.f - lr * weight_decay) - lr * s;
    mm = beta2 * mm + (1.f - beta2) * gp;
    p[i] = pp; m[i] = mm;
    store_lp(lp, lp_type, i, pp);
  }
  return 0;
}

// fp32 -> bf16/fp16 cast of a host buffer (the reference does this on the GPU after H2D: custom_cuda_kernel.cu:8-43;
// casting on the host halves the bytes that cross PCIe).
DSBH_EXPORT int dsbh_cast_f32_lp(const float* src, void* dst, int lp_type, int64_t n) {
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < n; ++i) store_lp(dst, lp_type, i, src[i]);
  return 0;
}

DSBH_EXPORT double dsbh_sumsq(const void* g, int grad_type, int64_t n) {
  double acc = 0.0;
#pragma omp parallel for reduction(+ : acc) schedule(static)
  for (int64_t i = 0; i < n; ++i) {
    const double v = load_grad(g, grad_type, i);
    acc += v * v;
  }
  return acc;
}

/*
 * crt_detmath.h -- deterministic f32 math shared by the CUDA kernels and the CPU oracle.
 *
 * WHY THIS EXISTS.  The reference shader (src/shaders/ComputeShader.wgsl) calls the WGSL
 * builtins sin/cos (CS:756-757), exp (CS:177, CS:438) and pow (CS:449,457,465).  Their
 * precision is implementation-defined in WGSL (it belongs to the browser's shader compiler,
 * which the reference does not pin), and CUDA libm and glibc differ from each other by ulps.
 * A path tracer is chaotic: a 1-ulp difference in one bounce direction flips a silhouette hit
 * and the whole path diverges, so CPU/GPU parity needs *one* definition of these functions,
 * made only of operations that are exactly rounded on both sides (+, -, *, /, fmaf, rintf,
 * int<->float conversion, bit casts).  This header is that definition.  It is a contract
 * header (like crt.h), compiled by nvcc into the product and by gcc into the oracle; nothing
 * else is shared between the two.
 *
 * Accuracy: sin/cos/exp about 1-2 ulp on the ranges the path uses, pow about 1e-6 relative
 * (WGSL itself defines pow's accuracy as that of exp2(y*log2(x))).
 *
 * Build rules that make this deterministic: gcc -ffp-contract=off -mfma (fmaf -> one vfmadd),
 * nvcc -fmad=false -prec-div=true -prec-sqrt=true -ftz=false (fmaf -> one FFMA, nothing else fused).
 */
#ifndef CRT_DETMATH_H
#define CRT_DETMATH_H

#include <stdint.h>
#include <math.h>
#include <string.h>

#ifdef __CUDACC__
#define CRT_HD __host__ __device__ __forceinline__
#else
#define CRT_HD static inline
#endif

CRT_HD uint32_t crt_f2u(float f) {
#ifdef __CUDA_ARCH__
  return __float_as_uint(f);
#else
  uint32_t u; memcpy(&u, &f, 4); return u;
#endif
}
CRT_HD float crt_u2f(uint32_t u) {
#ifdef __CUDA_ARCH__
  return __uint_as_float(u);
#else
  float f; memcpy(&f, &u, 4); return f;
#endif
}

/* sin and cos of x (radians), |x| up to a few thousand.  Cody-Waite reduction to
 * [-pi/4, pi/4] with a three-part pi/2, then the classic minimax polynomials. */
CRT_HD void crt_sincosf(float x, float* s_out, float* c_out) {
  float kf = rintf(x * 0.636619772f);                 /* x * 2/pi, to nearest even */
  int   k  = (int)kf;
  float r  = fmaf(kf, -1.5703125f, x);                /* pi/2 = 1.5703125 + 4.8375e-4 + 7.5498e-8 */
  r = fmaf(kf, -4.837512969970703125e-4f, r);
  r = fmaf(kf, -7.54978995489188216e-8f, r);
  float r2 = r * r;
  float sp = fmaf(r2, -1.9515295891e-4f, 8.3321608736e-3f);
  sp = fmaf(sp, r2, -1.6666654611e-1f);
  float sn = fmaf(sp * r2, r, r);
  float cp = fmaf(r2, 2.443315711809948e-5f, -1.388731625493765e-3f);
  cp = fmaf(cp, r2, 4.166664568298827e-2f);
  cp = fmaf(cp, r2, -0.5f);
  float cs = fmaf(cp, r2, 1.0f);
  float s = (k & 1) ? cs : sn;
  float c = (k & 1) ? sn : cs;
  if (k & 2) s = -s;
  if ((k + 1) & 2) c = -c;
  *s_out = s;
  *c_out = c;
}

/* 2^k * y for integer k in [-252, 254], split so neither factor over/underflows alone. */
CRT_HD float crt_scale2(float y, int k) {
  int k1 = k >> 1;
  int k2 = k - k1;
  float s1 = crt_u2f((uint32_t)(k1 + 127) << 23);
  float s2 = crt_u2f((uint32_t)(k2 + 127) << 23);
  return (y * s1) * s2;
}

/* e^r for |r| <= ~0.35 (polynomial core shared by exp and exp2) */
CRT_HD float crt_exp_core(float r) {
  float p = 1.9875691500e-4f;
  p = fmaf(p, r, 1.3981999507e-3f);
  p = fmaf(p, r, 8.3334519073e-3f);
  p = fmaf(p, r, 4.1665795894e-2f);
  p = fmaf(p, r, 1.6666665459e-1f);
  p = fmaf(p, r, 5.0000001201e-1f);
  return fmaf(p, r * r, r) + 1.0f;
}

/* e^x.  NaN -> NaN, x > 88.72 -> +inf, x < -87.33 -> 0 (results below FLT_MIN are flushed;
 * this is part of the pin, not an accident). */
CRT_HD float crt_expf(float x) {
  if (x != x) return x;
  if (x > 88.7228394f) return crt_u2f(0x7f800000u);
  if (x < -87.3365479f) return 0.0f;
  float kf = rintf(x * 1.44269504f);
  float r  = fmaf(kf, -0.693359375f, x);              /* ln2 = 0.693359375 - 2.12194440e-4 */
  r = fmaf(kf, 2.12194440e-4f, r);
  return crt_scale2(crt_exp_core(r), (int)kf);
}

/* 2^t with the same flush/overflow rules as crt_expf */
CRT_HD float crt_exp2f(float t) {
  if (t != t) return t;
  if (t >= 128.0f) return crt_u2f(0x7f800000u);
  if (t < -126.0f) return 0.0f;
  float kf = rintf(t);
  float r  = (t - kf) * 0.693147182f;
  return crt_scale2(crt_exp_core(r), (int)kf);
}

/* log2(x) for finite x > 0 (normal numbers; the path only calls it on [0.003, 1]) */
CRT_HD float crt_log2f(float x) {
  uint32_t ix = crt_f2u(x);
  int32_t  e  = ((int32_t)(ix - 0x3f3504f3u)) >> 23;  /* mantissa into [sqrt(1/2), sqrt(2)) */
  float m = crt_u2f(ix - ((uint32_t)e << 23));
  float f = m - 1.0f;
  float z = f * f;
  float p = 7.0376836292e-2f;
  p = fmaf(p, f, -1.1514610310e-1f);
  p = fmaf(p, f, 1.1676998740e-1f);
  p = fmaf(p, f, -1.2420140846e-1f);
  p = fmaf(p, f, 1.4249322787e-1f);
  p = fmaf(p, f, -1.6668057665e-1f);
  p = fmaf(p, f, 2.0000714765e-1f);
  p = fmaf(p, f, -2.4999993993e-1f);
  p = fmaf(p, f, 3.3333331174e-1f);
  float y = (f * z) * p;
  y = fmaf(-0.5f, z, y);
  float ln = f + y;
  return fmaf(ln, 1.44269504f, (float)e);
}

/* x^y for the gamma curve (CS:449): NaN -> NaN, x <= 0 -> 0, else exp2(y*log2(x)) */
CRT_HD float crt_powf(float x, float y) {
  if (x != x) return x;
  if (x <= 0.0f) return 0.0f;
  return crt_exp2f(y * crt_log2f(x));
}

#endif /* CRT_DETMATH_H */

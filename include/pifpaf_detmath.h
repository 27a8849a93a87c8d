/*
 * pifpaf_detmath.h -- deterministic single-precision exp() shared by the CUDA
 * assembly kernel and (as the checker's "det" mode) the CPU oracle.
 *
 * Why: the reference scores CAF candidates with numpy's float32 `np.exp`
 * (/root/reference/openpifpaf/decoder/generator/cifcaf.py:139), a SIMD kernel
 * that is not correctly rounded and differs between CPUs; CUDA's expf differs
 * again.  Every operation below is a single IEEE-754 binary32 operation with
 * round-to-nearest-even (explicit fmaf, no contraction), so the host and the
 * device produce bit-identical values and the GPU path can be compared with the
 * oracle bit for bit.  Accuracy: <= 1 ulp on [-104, 0] (tests/test_detmath.py),
 * i.e. inside the spread between numpy's and libm's own exp.
 *
 * Plain C99 / CUDA C++.  No dependencies.
 */
#ifndef PIFPAF_DETMATH_H_
#define PIFPAF_DETMATH_H_

#include <math.h>
#include <stdint.h>
#include <string.h>

#if defined(__CUDACC__)
#define PP_HD __host__ __device__ __forceinline__
#else
#define PP_HD static inline
#endif

PP_HD float pp_fma(float a, float b, float c) {
#if defined(__CUDA_ARCH__)
  return __fmaf_rn(a, b, c);
#else
  return fmaf(a, b, c);
#endif
}

PP_HD float pp_mul(float a, float b) {
#if defined(__CUDA_ARCH__)
  return __fmul_rn(a, b);
#else
  volatile float r = a * b; /* volatile: no contraction with a following add */
  return r;
#endif
}

PP_HD float pp_add(float a, float b) {
#if defined(__CUDA_ARCH__)
  return __fadd_rn(a, b);
#else
  volatile float r = a + b;
  return r;
#endif
}

/* exp(x) for x <= 0 (the only range the decoder uses: x = -0.5 d^2 / sigma^2).
 * Cody-Waite reduction x = n ln2 + r, |r| <= ln2/2, degree-7 polynomial,
 * scaling by 2^n through the exponent field.  Positive / NaN / -inf inputs are
 * handled so that the function is total: NaN -> NaN, x > 0 uses the same code
 * (valid up to 88), x < -104 -> 0. */
PP_HD float pp_expf(float x) {
  if (x != x) return x;
  if (x < -104.0f) return 0.0f;
  if (x > 88.0f) return INFINITY;
  const float kLog2e = 1.4426950408889634f;
  const float kLn2Hi = 0.693145751953125f;        /* 0x3f317200 */
  const float kLn2Lo = 1.42860682030941723212e-6f; /* ln2 - hi  */
  float t = pp_mul(x, kLog2e);
  float n = rintf(t);
  float r = pp_fma(-n, kLn2Hi, x);
  r = pp_fma(-n, kLn2Lo, r);
  /* e^r ~= 1 + r + r^2 P(r), P = degree-5 minimax (Cephes expf constants) */
  float p = 1.9875691500e-4f;
  p = pp_fma(p, r, 1.3981999507e-3f);
  p = pp_fma(p, r, 8.3334519073e-3f);
  p = pp_fma(p, r, 4.1665795894e-2f);
  p = pp_fma(p, r, 1.6666665459e-1f);
  p = pp_fma(p, r, 5.0000001201e-1f);
  float r2 = pp_mul(r, r);
  float e = pp_fma(p, r2, r);
  e = pp_add(e, 1.0f);
  /* scale by 2^n in two steps so that subnormal results round once */
  int ni = (int)n;
  int n1 = ni < -126 ? ni + 64 : ni;
  uint32_t bits = (uint32_t)(n1 + 127) << 23;
  float s;
#if defined(__CUDA_ARCH__)
  s = __uint_as_float(bits);
#else
  memcpy(&s, &bits, sizeof s);
#endif
  e = pp_mul(e, s);
  if (ni < -126) e = pp_mul(e, 5.42101086242752217e-20f); /* 2^-64 */
  return e;
}

#endif /* PIFPAF_DETMATH_H_ */

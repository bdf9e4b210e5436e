// airc_math.cuh -- scalar FP32 building blocks of the aircraft kernels (sm_100a).
//
// Every function is AIRC_HD so that oracle/host_emul.cpp can compile the very same FP32
// formulas with g++ and check them on a machine without a GPU (test infrastructure only;
// the shipped library contains the device versions alone).
#pragma once
#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define AIRC_HD __host__ __device__ __forceinline__
#define AIRC_D __device__ __forceinline__
#else
#define AIRC_HD inline
#define AIRC_D inline
#endif

namespace airc {

// ---------------------------------------------------------------- MUFU wrappers
AIRC_HD float fast_ex2(float x) {
#if defined(__CUDA_ARCH__)
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
#else
  return exp2f(x);
#endif
}

AIRC_HD float fast_rcp(float x) {
#if defined(__CUDA_ARCH__)
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
#else
  return 1.0f / x;
#endif
}

AIRC_HD float fast_rsqrt(float x) {
#if defined(__CUDA_ARCH__)
  float y;
  asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
#else
  return 1.0f / sqrtf(x);
#endif
}

// 1/x to <= 1 ulp: MUFU.RCP seed + one Newton step on the FMA pipe.
AIRC_HD float rcp_nr(float x) {
  float r = fast_rcp(x);
  float e = fmaf(-x, r, 1.0f);
  return fmaf(r, e, r);
}

// sqrt(x) and 1/sqrt(x) for x > 0 (one MUFU.RSQ + Newton); both ~1 ulp.
AIRC_HD void sqrt_rsqrt(float x, float& s, float& rs) {
  float r = fast_rsqrt(x);
  // one Newton step for rsqrt: r <- r*(1.5 - 0.5*x*r*r)
  float xr = x * r;                       // ~ sqrt(x)
  float h = 0.5f * r;
  float e = fmaf(-xr, h, 0.5f);           // 0.5 - 0.5*x*r*r
  s = fmaf(xr, e, xr);
  rs = fmaf(r, e, r);
}

// ---------------------------------------------------------------- logistic in base 2
// The surrogate's tanh is evaluated as tanh(a) = 1 - 2 r, r = 1/(1 + 2^h), h = 2*log2(e)*a.
// The factor 2*log2(e) is folded into the first-layer weights and the affine map t -> r into
// the output layer (see fuse_network in airc_api.cu), so the kernels only ever need r.
// h is clamped at +40 (tanh saturated to < 1 ulp beyond |a| > 9.1; clamp at a = 13.9) so that
// the product of two denominators cannot overflow in the shared-reciprocal form below.
#define AIRC_H_CLAMP 40.0f

AIRC_HD float sigm2(float h) {
  float e = fast_ex2(fminf(h, AIRC_H_CLAMP));
  return fast_rcp(1.0f + e);
}

// two logistic values with ONE reciprocal (Montgomery's trick): 3 MUFU instead of 4.
AIRC_HD void sigm2_pair(float h0, float h1, float& r0, float& r1) {
  float d0 = 1.0f + fast_ex2(fminf(h0, AIRC_H_CLAMP));
  float d1 = 1.0f + fast_ex2(fminf(h1, AIRC_H_CLAMP));
  float inv = fast_rcp(d0 * d1);
  r0 = d1 * inv;
  r1 = d0 * inv;
}

// ---------------------------------------------------------------- inverse trig
// Round-1 versions call the CUDA math library (accurate to 2 ulp); they are 8 calls per
// RK4 step against 128 logistic evaluations.
AIRC_HD float atan2_acc(float y, float x) { return atan2f(y, x); }
AIRC_HD float asin_acc(float x) { return asinf(x); }

AIRC_HD float sign_casadi(float x) {  // casadi sign(): -1, 0, +1 ; derivative 0
  return (x > 0.0f) ? 1.0f : ((x < 0.0f) ? -1.0f : (x == 0.0f ? 0.0f : x));
}

}  // namespace airc

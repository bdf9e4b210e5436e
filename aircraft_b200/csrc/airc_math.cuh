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

// min / max that PROPAGATE NaN (fminf / fmaxf return the other operand): a NaN command or state must reach the
// results, as it does through CasADi's tanh / fmin in the reference (main/dynamics/dynamics.py:108 checks for it).
AIRC_HD float min_nan(float a, float b) {
#if defined(__CUDA_ARCH__)
  float r;
  asm("min.NaN.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
  return r;
#else
  return (a != a) ? a : ((b != b) ? b : (a < b ? a : b));
#endif
}
AIRC_HD float max_nan(float a, float b) {
#if defined(__CUDA_ARCH__)
  float r;
  asm("max.NaN.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
  return r;
#else
  return (a != a) ? a : ((b != b) ? b : (a > b ? a : b));
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

// ---------------------------------------------------------------- packed FP32x2 (Blackwell FFMA2)
// sm_100 adds fma/add/mul.rn.f32x2 on 64-bit register pairs (SASS FFMA2/FADD2/FMUL2): the same
// FMA-lane throughput as two FFMA but ONE issue slot, and the multiplier may be a pair of uniform
// registers (LDCU.128 feeds two FFMA2 = four weights).  Measured on B200 (profiles/
// r1_pipe_probe.txt): FFMA2 with a uniform operand 123 FMA-lanes/clk/SM vs 115 for FFMA, so the
// gain is purely in issue slots -- which is what bounds the rollout kernel.
struct F2 {
  float lo, hi;
};

#if defined(__CUDA_ARCH__)
AIRC_D unsigned long long f2_pack(const F2& a) {
  unsigned long long r;
  asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(a.lo), "f"(a.hi));
  return r;
}
AIRC_D F2 f2_unpack(unsigned long long v) {
  F2 a;
  asm("mov.b64 {%0,%1}, %2;" : "=f"(a.lo), "=f"(a.hi) : "l"(v));
  return a;
}
#endif

// Two spellings of the same three instructions.  fma2 / add2 / mul2 are inline PTX: the form the packed rollout
// kernels (airc_packed.cuh) were tuned with -- ptxas then keeps the weights in uniform registers with the fewest
// moves.  fma2b / add2b / mul2b are the CUDA 12.8+ builtins: identical arithmetic, but the front end handles
// thousands of them per function in seconds where the same number of asm statements takes minutes, so the
// surrogate-with-Jacobian (airc_dynamics.cuh, airc_coop.cuh: 16 unrolled unit pairs) uses these.  The tangent / jet
// vocabulary below stays on the PTX spelling: with the builtins the jet kernels (airc_jet.cuh) spill six times more.
AIRC_HD F2 fma2(const F2& a, const F2& b, const F2& c) {
#if defined(__CUDA_ARCH__)
  unsigned long long d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(f2_pack(a)), "l"(f2_pack(b)), "l"(f2_pack(c)));
  return f2_unpack(d);
#else
  return F2{fmaf(a.lo, b.lo, c.lo), fmaf(a.hi, b.hi, c.hi)};
#endif
}
AIRC_HD F2 add2(const F2& a, const F2& b) {
#if defined(__CUDA_ARCH__)
  unsigned long long d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(f2_pack(a)), "l"(f2_pack(b)));
  return f2_unpack(d);
#else
  return F2{a.lo + b.lo, a.hi + b.hi};
#endif
}
AIRC_HD F2 mul2(const F2& a, const F2& b) {
#if defined(__CUDA_ARCH__)
  unsigned long long d;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(f2_pack(a)), "l"(f2_pack(b)));
  return f2_unpack(d);
#else
  return F2{a.lo * b.lo, a.hi * b.hi};
#endif
}
AIRC_HD F2 fma2b(const F2& a, const F2& b, const F2& c) {
#if defined(__CUDA_ARCH__)
  const float2 d = __ffma2_rn(make_float2(a.lo, a.hi), make_float2(b.lo, b.hi), make_float2(c.lo, c.hi));
  return F2{d.x, d.y};
#else
  return F2{fmaf(a.lo, b.lo, c.lo), fmaf(a.hi, b.hi, c.hi)};
#endif
}
AIRC_HD F2 add2b(const F2& a, const F2& b) {
#if defined(__CUDA_ARCH__)
  const float2 d = __fadd2_rn(make_float2(a.lo, a.hi), make_float2(b.lo, b.hi));
  return F2{d.x, d.y};
#else
  return F2{a.lo + b.lo, a.hi + b.hi};
#endif
}
AIRC_HD F2 mul2b(const F2& a, const F2& b) {
#if defined(__CUDA_ARCH__)
  const float2 d = __fmul2_rn(make_float2(a.lo, a.hi), make_float2(b.lo, b.hi));
  return F2{d.x, d.y};
#else
  return F2{a.lo * b.lo, a.hi * b.hi};
#endif
}
AIRC_HD F2 bc2(float a) { return F2{a, a}; }

// ---- one vocabulary for scalar (float) and packed (F2 = two tangent columns) arithmetic, so that the
// forward-mode tangent code is written once.  Primal quantities stay scalar and enter packed
// instructions as broadcast operands (free in SASS: "R.F32" operand form of FFMA2).
AIRC_HD float tfma(float a, float b, float c) { return fmaf(a, b, c); }
AIRC_HD F2 tfma(float a, const F2& b, const F2& c) { return fma2(bc2(a), b, c); }
AIRC_HD F2 tfma(const F2& a, float b, const F2& c) { return fma2(a, bc2(b), c); }
AIRC_HD F2 tfma(const F2& a, const F2& b, const F2& c) { return fma2(a, b, c); }
AIRC_HD float tmul(float a, float b) { return a * b; }
AIRC_HD F2 tmul(float a, const F2& b) { return mul2(bc2(a), b); }
AIRC_HD F2 tmul(const F2& a, float b) { return mul2(a, bc2(b)); }
AIRC_HD F2 tmul(const F2& a, const F2& b) { return mul2(a, b); }
AIRC_HD float tadd(float a, float b) { return a + b; }
AIRC_HD F2 tadd(const F2& a, const F2& b) { return add2(a, b); }
AIRC_HD float tsub(float a, float b) { return a - b; }
AIRC_HD F2 tsub(const F2& a, const F2& b) { return fma2(bc2(-1.0f), b, a); }
template <class T> AIRC_HD T tzero();
template <> AIRC_HD float tzero<float>() { return 0.0f; }
template <> AIRC_HD F2 tzero<F2>() { return F2{0.0f, 0.0f}; }
// a*b - c*d and a*b + c*d with scalar (primal) a, c
template <class T> AIRC_HD T tdiff(float a, const T& b, float c, const T& d) { return tfma(a, b, tmul(-c, d)); }
template <class T> AIRC_HD T tsum(float a, const T& b, float c, const T& d) { return tfma(a, b, tmul(c, d)); }

// ---------------------------------------------------------------- logistic in base 2
// The surrogate's tanh is evaluated as tanh(a) = 1 - 2 r, r = 1/(1 + 2^h), h = 2*log2(e)*a.
// The factor 2*log2(e) is folded into the first-layer weights and the affine map t -> r into
// the output layer (see fuse_network in airc_api.cu), so the kernels only ever need r.
// h is clamped at +40 (tanh saturated to < 1 ulp beyond |a| > 9.1; clamp at a = 13.9) so that
// the product of two denominators cannot overflow in the shared-reciprocal form below.  The clamp
// is NaN-preserving (min_nan): a NaN input makes every pre-activation NaN and must come out as NaN.
#define AIRC_H_CLAMP 40.0f

AIRC_HD float sigm2(float h) {
  float e = fast_ex2(min_nan(h, AIRC_H_CLAMP));
  return fast_rcp(1.0f + e);
}

// Logistic pair without a clamp: 2^h -> +inf gives 1+inf = inf and rcp(inf) = 0 (tanh = +1);
// 2^h -> 0 (ftz) gives r = 1 (tanh = -1); NaN propagates.  2 issue slots + 2 MUFU per unit.
// MUFU runs at 4 lanes/clk/SMSP (8 cycles per warp instruction, measured), i.e. 16 XU-cycles per
// hidden unit here against 12 for the shared-reciprocal form below, which in turn costs 2 more
// issue slots per unit (clamp + 1.5 multiplies); the kernel is issue-bound, so this form wins.
// A/B on B200 (profiles/r1_variants.txt): mode 1 (shared reciprocal) 2.72 ms vs mode 0 (no clamp,
// two reciprocals) 2.78 ms per 25M state-steps: the XU pipe (47 % busy in mode 1) matters more than
// the two extra issue slots.
#ifndef AIRC_SIGM_MODE
#define AIRC_SIGM_MODE 1
#endif
AIRC_HD F2 sigm2_x2(const F2& h) {
#if AIRC_SIGM_MODE == 0
  const F2 d = add2b(F2{fast_ex2(h.lo), fast_ex2(h.hi)}, F2{1.0f, 1.0f});
  return F2{fast_rcp(d.lo), fast_rcp(d.hi)};
#else  // shared reciprocal (3 MUFU per pair), needs the clamp so that d0*d1 stays finite
  const F2 d = add2b(F2{fast_ex2(min_nan(h.lo, 40.0f)), fast_ex2(min_nan(h.hi, 40.0f))}, F2{1.0f, 1.0f});
  const float inv = fast_rcp(d.lo * d.hi);
  return F2{d.hi * inv, d.lo * inv};
#endif
}

// two logistic values with ONE reciprocal (Montgomery's trick): 3 MUFU instead of 4.
AIRC_HD void sigm2_pair(float h0, float h1, float& r0, float& r1) {
  float d0 = 1.0f + fast_ex2(min_nan(h0, AIRC_H_CLAMP));
  float d1 = 1.0f + fast_ex2(min_nan(h1, AIRC_H_CLAMP));
  float inv = fast_rcp(d0 * d1);
  r0 = d1 * inv;
  r1 = d0 * inv;
}

// ---------------------------------------------------------------- inverse trig
AIRC_HD F2 neg2(const F2& a) { return F2{-a.lo, -a.hi}; }
AIRC_HD float sel(bool c, float a, float b) { return c ? a : b; }

// atan2(y, x), all quadrants, branch-free (an infinite argument yields NaN: a diverged
// trajectory is reported as NaN one step earlier than libm would).  atan(t) = t + t s Q(s), s = t^2, t in [0,1]:
// degree-8 near-minimax Q (max error 6.3e-8 abs, 8.5e-8 rel incl. FP32 evaluation).
AIRC_HD F2 atan2_2(const F2& y, const F2& x) {
  const F2 ax{fabsf(x.lo), fabsf(x.hi)}, ay{fabsf(y.lo), fabsf(y.hi)};
  // selects instead of min/max so that a NaN argument reaches t (FMNMX would drop it)
  const bool glo = ay.lo > ax.lo, ghi = ay.hi > ax.hi;
  const F2 mx{sel(glo, ay.lo, ax.lo), sel(ghi, ay.hi, ax.hi)};
  const F2 mn{sel(glo, ax.lo, ay.lo), sel(ghi, ax.hi, ay.hi)};
  // t = mn/mx: MUFU.RCP (1 ulp) + a residual correction of the quotient (~1 ulp); 0/0 -> t = 0
  const F2 r{fast_rcp(mx.lo == 0.0f ? 1.0f : mx.lo), fast_rcp(mx.hi == 0.0f ? 1.0f : mx.hi)};
  F2 t = mul2(mn, r);
  t = fma2(fma2(neg2(t), mx, mn), r, t);      // one residual step on the quotient: no Newton step on r needed
  const F2 s = mul2(t, t);
  F2 q = bc2(-2.447031215e-03f);
  q = fma2(q, s, bc2(1.375028427e-02f));
  q = fma2(q, s, bc2(-3.627017439e-02f));
  q = fma2(q, s, bc2(6.284360929e-02f));
  q = fma2(q, s, bc2(-8.673170879e-02f));
  q = fma2(q, s, bc2(1.103799429e-01f));
  q = fma2(q, s, bc2(-1.427911092e-01f));
  q = fma2(q, s, bc2(1.999976600e-01f));
  q = fma2(q, s, bc2(-3.333333195e-01f));
  F2 p = fma2(mul2(t, s), q, t);
  // |y| > |x| : pi/2 - p ;  x < 0 : pi - p ;  sign of y
  const float hp = 1.57079632679f, pi = 3.14159265359f;
  p.lo = sel(glo, hp - p.lo, p.lo);
  p.hi = sel(ghi, hp - p.hi, p.hi);
  p.lo = sel(x.lo < 0.0f, pi - p.lo, p.lo);
  p.hi = sel(x.hi < 0.0f, pi - p.hi, p.hi);
  p.lo = copysignf(p.lo, y.lo);
  p.hi = copysignf(p.hi, y.hi);
  return p;
}

// alpha = atan2(vb_z, vb_x + eps) and beta = asin(vb_y / V) = atan2(vb_y, sqrt(vb_x^2 + vb_z^2 + eps)) in ONE packed
// evaluation of the polynomial (base.py:167,176,234): ~35 instructions, no branches, no libm call.
AIRC_HD void alpha_beta(float vbz, float xa, float vby, float rxz, float& alpha, float& beta) {
  const F2 r = atan2_2(F2{vbz, vby}, F2{xa, rxz});
  alpha = r.lo;
  beta = r.hi;
}

// The CUDA math library versions (2 ulp) remain for the registry's table models and the jets (airc_alt.cuh,
// airc_jet.cuh), which are not latency-critical.
AIRC_HD float atan2_acc(float y, float x) { return atan2f(y, x); }
AIRC_HD float asin_acc(float x) { return asinf(x); }

AIRC_HD float sign_casadi(float x) {  // casadi sign(): -1, 0, +1 ; derivative 0
  return (x > 0.0f) ? 1.0f : ((x < 0.0f) ? -1.0f : (x == 0.0f ? 0.0f : x));
}

}  // namespace airc

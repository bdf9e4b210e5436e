// airc_packed.cuh -- the primal rollout arithmetic on packed FP32x2 registers: every thread carries
// TWO trajectories (lo, hi halves of 64-bit register pairs) through f, RK4 and the loss.
//
// Why (B200): sm_100's FFMA2/FMUL2/FADD2 do two FP32 lanes per issue slot, and a scalar multiplier
// may be a broadcast uniform register ("UR.F32"), so one LDCU.128 of network weights feeds FOUR
// packed FMAs.  The one-trajectory kernel (airc_step.cuh) was issue-bound (78 % issue-active, FMA
// pipe 68 %); here all non-MUFU arithmetic of two trajectories shares its issue slots.
//
// Same mathematics as airc_dynamics.cuh (reference lines cited there), with two differences that
// keep everything branch-free and packable:
//   * alpha = atan2(vb_z, vb_x+eps) and beta = asin(vb_y/V) use one polynomial atan2:
//     asin(y/sqrt(y^2+r^2)) == atan2(y, r) with r = sqrt(vb_x^2+vb_z^2+eps)   (base.py:167,176,234)
//   * sign(), the quadrant fix-ups and clamps are selects.
// All arithmetic goes through explicit PTX (fma/mul/add.rn.f32x2), so the compiler cannot contract
// or reassociate differently in different kernels: the sweep, the winner re-roll and the rollout
// kernels produce bit-identical trajectories.
#pragma once
#include "airc_dynamics.cuh"

namespace airc {

// network weights in unit-major order for broadcast operands: 12 floats = 3 x LDCU.128 per unit
struct alignas(16) UnitW {
  float a0, Au0, Au1, Ax0;   // bias, aileron, elevator, qbar
  float Ax1, Ax2, wo0, wo1;  // alpha, beta, output rows 0,1
  float wo2, wo3, wo4, wo5;
};
struct alignas(16) UnitNet {
  UnitW u[NH];
  float c0[6];
  float pad_[2];
};

AIRC_HD void make_unit_net(const FusedNet& f, UnitNet& n) {
  for (int j = 0; j < NH; ++j) {
    const HidPair& w = f.hid[j / 2];
    const int h = j & 1;
    UnitW& o = n.u[j];
    o.a0 = w.a0[h]; o.Au0 = w.Au[0][h]; o.Au1 = w.Au[1][h];
    o.Ax0 = w.Ax[0][h]; o.Ax1 = w.Ax[1][h]; o.Ax2 = w.Ax[2][h];
    o.wo0 = f.Wo[0][j]; o.wo1 = f.Wo[1][j]; o.wo2 = f.Wo[2][j];
    o.wo3 = f.Wo[3][j]; o.wo4 = f.Wo[4][j]; o.wo5 = f.Wo[5][j];
  }
  for (int k = 0; k < 6; ++k) n.c0[k] = f.c0[k];
  n.pad_[0] = n.pad_[1] = 0.0f;
}

// ---------------------------------------------------------------- packed helpers
AIRC_HD F2 sub2(const F2& a, const F2& b) { return fma2(bc2(-1.0f), b, a); }   // a - b, one FFMA2
AIRC_HD F2 fmas(float a, const F2& b, const F2& c) { return fma2(bc2(a), b, c); }     // scalar * pair + pair
AIRC_HD F2 fmass(float a, const F2& b, float c) { return fma2(bc2(a), b, bc2(c)); }   // scalar * pair + scalar
AIRC_HD F2 muls(float a, const F2& b) { return mul2(bc2(a), b); }
AIRC_HD F2 adds(const F2& a, float b) { return add2(a, bc2(b)); }

AIRC_HD F2 rcp2_nr(const F2& x) {            // 1/x to ~1 ulp: 2 MUFU.RCP + packed Newton step
  const F2 r{fast_rcp(x.lo), fast_rcp(x.hi)};
  const F2 e = fma2(neg2(x), r, bc2(1.0f));
  return fma2(r, e, r);
}

AIRC_HD void sqrt_rsqrt2(const F2& x, F2& s, F2& rs) {   // sqrt and 1/sqrt, x > 0
  const F2 r{fast_rsqrt(x.lo), fast_rsqrt(x.hi)};
  const F2 xr = mul2(x, r);
  const F2 h = muls(0.5f, r);
  const F2 e = fma2(neg2(xr), h, bc2(0.5f));
  s = fma2(xr, e, xr);
  rs = fma2(r, e, r);
}

// ---------------------------------------------------------------- state, controls
struct State2 {
  F2 p[3], v[3], q[4], w[3];
};
struct Ctl2 {
  F2 da, de, dr, df;
};

AIRC_HD void rot_matrix2(const F2 q[4], F2 M[9], F2& n2) {
  const F2 x = q[0], y = q[1], z = q[2], w = q[3];
  const F2 xx = mul2(x, x), yy = mul2(y, y), zz = mul2(z, z), ww = mul2(w, w);
  n2 = add2(add2(xx, yy), add2(zz, ww));
  M[0] = sub2(add2(ww, xx), add2(yy, zz));
  M[4] = sub2(add2(ww, yy), add2(xx, zz));
  M[8] = sub2(add2(ww, zz), add2(xx, yy));
  const F2 x2 = add2(x, x), y2 = add2(y, y), w2 = add2(w, w);
  const F2 xy = mul2(x2, y), xz = mul2(x2, z), yz = mul2(y2, z);
  const F2 wx = mul2(w2, x), wy = mul2(w2, y), wz = mul2(w2, z);
  M[1] = sub2(xy, wz);
  M[2] = add2(xz, wy);
  M[3] = add2(xy, wz);
  M[5] = sub2(yz, wx);
  M[6] = sub2(xz, wy);
  M[7] = add2(yz, wx);
}

// Logistic of both halves with INDEPENDENT reciprocals: the halves are different trajectories, and
// the shared-reciprocal trick of airc_math.cuh (r = d_other / (d_lo d_hi)) would make a trajectory's
// rounding depend on its lane partner, i.e. on how the batch is sharded.  2^h -> inf gives r = 0,
// 2^h -> 0 gives r = 1, NaN propagates; no clamp needed.
AIRC_HD F2 sigm2_indep(const F2& h) {
  const F2 d = add2(F2{fast_ex2(h.lo), fast_ex2(h.hi)}, bc2(1.0f));
  return F2{fast_rcp(d.lo), fast_rcp(d.hi)};
}

// casadi sign(): -1, 0, +1; NaN stays NaN
AIRC_HD float sign1(float x) { return (x > 0.0f) ? 1.0f : ((x < 0.0f) ? -1.0f : x); }

// Partial unrolling of the unit loop bounds the register footprint (full unrolling: 228 registers,
// 2 blocks/SM; by 8: 162 registers, 3 blocks/SM, no spills), weights then come from
// uniform-register-indexed constant loads.
#ifndef AIRC_NN_UNROLL
#define AIRC_NN_UNROLL 16
#endif
constexpr int kNNUnroll = AIRC_NN_UNROLL;
// (Tried and rejected, profiles/r1_variants_packed.txt: parking the control part of the hidden
// pre-activations in shared memory once per step -- 10 % fewer FMAs but 20 % slower, LDS competes with
// MUFU for the MIO queue; sharing one reciprocal between two units of a trajectory -- 2 % slower, the FMA
// pipe, not the XU pipe, is the limiter.)
// the surrogate for two trajectories: one hidden unit per 11 packed FMAs, weights broadcast
AIRC_HD void surrogate2(const UnitNet& net, const F2& qbar, const F2& alpha, const F2& beta,
                        const Ctl2& c, F2 C[6]) {
#pragma unroll
  for (int k = 0; k < 6; ++k) C[k] = bc2(net.c0[k]);
#pragma unroll kNNUnroll
  for (int j = 0; j < NH; ++j) {
    const UnitW& w = net.u[j];
    F2 h = fmass(w.Au0, c.da, w.a0);
    h = fmas(w.Au1, c.de, h);
    h = fmas(w.Ax0, qbar, h);
    h = fmas(w.Ax1, alpha, h);
    h = fmas(w.Ax2, beta, h);
    const F2 r = sigm2_indep(h);
    C[0] = fmas(w.wo0, r, C[0]);
    C[1] = fmas(w.wo1, r, C[1]);
    C[2] = fmas(w.wo2, r, C[2]);
    C[3] = fmas(w.wo3, r, C[3]);
    C[4] = fmas(w.wo4, r, C[4]);
    C[5] = fmas(w.wo5, r, C[5]);
  }
}

// How f gets its coefficients: on the FMA pipe from the unit-major weights (here), or on the tensor cores
// (SurTc, airc_tc.cuh).  Everything around the surrogate is the same code for both.
// begin() gets the network's inputs, end() delivers the coefficients; f puts the work that needs neither in between.
struct SurFfma {
  const UnitNet& net;
  F2 Cs[6];
  AIRC_HD void begin(const F2& qbar, const F2& alpha, const F2& beta, const Ctl2& c) {
    surrogate2(net, qbar, alpha, beta, c, Cs);
  }
  AIRC_HD void end(F2 C[6]) {
#pragma unroll
    for (int k = 0; k < 6; ++k) C[k] = Cs[k];
  }
};

// f(x,u) for two trajectories (p_dot = v is handled by the caller).  base.py:385-406.
template <class Sur>
AIRC_HD void f_eval2(const DevConst& P, Sur& sur, const F2 v[3], const F2 q[4],
                     const F2 w[3], const Ctl2& c, F2 vdot[3], F2 qdot[4], F2 wdot[3], F2& vb2_out) {
  F2 M[9], n2;
  rot_matrix2(q, M, n2);
  const F2 inv_n2 = rcp2_nr(n2);
  // v_b = M^T v / n2 + eps                                   base.py:156-159
  F2 vb[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    F2 a = mul2(M[i], v[0]);
    a = fma2(M[3 + i], v[1], a);
    a = fma2(M[6 + i], v[2], a);
    vb[i] = fma2(a, inv_n2, bc2(P.eps));
  }
  const F2 xz2 = fma2(vb[2], vb[2], mul2(vb[0], vb[0]));
  const F2 vb2 = fma2(vb[1], vb[1], xz2);
  vb2_out = vb2;
  const F2 qbar = muls(0.6125f, vb2);                         // base.py:240
  const F2 xa = adds(vb[0], P.eps);
  const F2 alpha = atan2_2(vb[2], xa);                        // base.py:176
  F2 rxz, rs_unused;
  sqrt_rsqrt2(adds(xz2, P.eps), rxz, rs_unused);              // sqrt(V^2 - vb_y^2), V^2 = vb2+eps
  const F2 beta = atan2_2(vb[1], rxz);                        // == asin(vb_y/V), base.py:167,234

  sur.begin(qbar, alpha, beta, c);
  // ---- the part of f that does not need the coefficients
  // q_dot = 0.5 q (x) (w,0)                                  base.py:292-295
  const F2 hw0 = muls(0.5f, w[0]), hw1 = muls(0.5f, w[1]), hw2 = muls(0.5f, w[2]);
  qdot[0] = fma2(q[3], hw0, fma2(q[1], hw2, mul2(neg2(q[2]), hw1)));
  qdot[1] = fma2(q[3], hw1, fma2(q[2], hw0, mul2(neg2(q[0]), hw2)));
  qdot[2] = fma2(q[3], hw2, fma2(q[0], hw1, mul2(neg2(q[1]), hw0)));
  qdot[3] = neg2(fma2(q[0], hw0, fma2(q[1], hw1, mul2(q[2], hw2))));
  // w x (I w)                                                base.py:376-377
  F2 Iw[3], gy[3];
#pragma unroll
  for (int i = 0; i < 3; ++i)
    Iw[i] = fmas(P.I[3 * i + 2], w[2], fmas(P.I[3 * i + 1], w[1], muls(P.I[3 * i], w[0])));
  gy[0] = fma2(w[1], Iw[2], mul2(neg2(w[2]), Iw[1]));
  gy[1] = fma2(w[2], Iw[0], mul2(neg2(w[0]), Iw[2]));
  gy[2] = fma2(w[0], Iw[1], mul2(neg2(w[1]), Iw[0]));
  F2 C[6];
  sur.end(C);
  C[5] = fmas(P.k_rudder, c.dr, C[5]);                        // coefficient_models.py:102-103
  if (P.stall_scaling) {                                      // aircraft.py:280-294 (default off)
    const float k = 10.0f, a30 = 0.52359877559f;
    const F2 sa{1.0f / (1.0f + expf(k * (fabsf(alpha.lo) - a30))), 1.0f / (1.0f + expf(k * (fabsf(alpha.hi) - a30)))};
    const F2 sb{1.0f / (1.0f + expf(k * (fabsf(beta.lo) - a30))), 1.0f / (1.0f + expf(k * (fabsf(beta.hi) - a30)))};
    C[2] = mul2(mul2(C[2], sa), sb);
    C[4] = mul2(C[4], sa);
  }
  C[0] = fmas(P.k_flap_x, c.df, C[0]);                        // aircraft.py:297-299
  C[2] = fmas(P.k_flap_z, c.df, C[2]);

  // forces and moments, body frame                           aircraft.py:313-328 ; base.py:255,275
  const F2 qS = muls(P.S, qbar);
  const F2 sgn{sign1(vb[0].lo), sign1(vb[0].hi)};
  F2 F[3];
  F[0] = mul2(mul2(C[0], qS), sgn);
  F[1] = mul2(C[1], qS);
  F[2] = mul2(C[2], qS);
  const F2 qSb = muls(P.Sb, qbar), qSc = muls(P.Sc, qbar);
  F2 Mo[3];
  Mo[0] = fma2(C[3], qSb, fmas(P.com[1], F[2], muls(-P.com[2], F[1])));
  Mo[1] = fma2(C[4], qSc, fmas(P.com[2], F[0], muls(-P.com[0], F[2])));
  Mo[2] = fma2(C[5], qSb, fmas(P.com[0], F[1], muls(-P.com[1], F[0])));

  // v_dot = M F /(n2 m) + g                                  base.py:284-286,359
  const F2 sc = muls(P.inv_mass, inv_n2);
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    F2 a = mul2(M[3 * i], F[0]);
    a = fma2(M[3 * i + 1], F[1], a);
    a = fma2(M[3 * i + 2], F[2], a);
    vdot[i] = fma2(a, sc, bc2(P.g[i]));
  }
  // w_dot = I^-1 (M - w x (I w))                             base.py:376-377
  F2 t[3];
  t[0] = sub2(Mo[0], gy[0]);
  t[1] = sub2(Mo[1], gy[1]);
  t[2] = sub2(Mo[2], gy[2]);
#pragma unroll
  for (int i = 0; i < 3; ++i)
    wdot[i] = fmas(P.Iinv[3 * i + 2], t[2], fmas(P.Iinv[3 * i + 1], t[1], muls(P.Iinv[3 * i], t[0])));
}

// ---------------------------------------------------------------- RK4, normalise, loss pieces
// state_step for two trajectories, base.py:408-446 (all 13 components additively).
// speed2 receives |v_b(x)|^2 of the step's FIRST stage, i.e. of the node itself: the loss's speed term
// (control.py:63-64, aircraft.v_frd_rel on every node) costs nothing extra.
template <class Sur>
AIRC_HD void rk4_step2(const DevConst& P, Sur& sur, State2& x, const Ctl2& c, float h, F2& speed2) {
  const float hh = 0.5f * h, h6 = h * (1.0f / 6.0f);
  F2 sv[3], sq[4], sw[3], ap[3], av[3], aq[4], aw[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) { sv[i] = x.v[i]; sw[i] = x.w[i]; ap[i] = av[i] = aw[i] = bc2(0.0f); }
#pragma unroll
  for (int i = 0; i < 4; ++i) { sq[i] = x.q[i]; aq[i] = bc2(0.0f); }
#pragma unroll 1
  for (int stage = 0; stage < 4; ++stage) {
    const float wgt = (stage == 0 || stage == 3) ? 1.0f : 2.0f;
    const float cn = (stage == 2) ? h : hh;
    F2 kv[3], kq[4], kw[3], vb2;
    f_eval2(P, sur, sv, sq, sw, c, kv, kq, kw, vb2);
    if (stage == 0) speed2 = vb2;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      ap[i] = fmas(wgt, sv[i], ap[i]);
      av[i] = fmas(wgt, kv[i], av[i]);
      aw[i] = fmas(wgt, kw[i], aw[i]);
      sv[i] = fmas(cn, kv[i], x.v[i]);
      sw[i] = fmas(cn, kw[i], x.w[i]);
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      aq[i] = fmas(wgt, kq[i], aq[i]);
      sq[i] = fmas(cn, kq[i], x.q[i]);
    }
  }
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    x.p[i] = fmas(h6, ap[i], x.p[i]);
    x.v[i] = fmas(h6, av[i], x.v[i]);
    x.w[i] = fmas(h6, aw[i], x.w[i]);
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) x.q[i] = fmas(h6, aq[i], x.q[i]);
}

AIRC_HD void normalise_q2(State2& x) {          // base.py:443-444, 473-474
  const F2 n2 = add2(fma2(x.q[1], x.q[1], mul2(x.q[0], x.q[0])), fma2(x.q[3], x.q[3], mul2(x.q[2], x.q[2])));
  F2 n, rn;
  sqrt_rsqrt2(n2, n, rn);
#pragma unroll
  for (int i = 0; i < 4; ++i) x.q[i] = mul2(x.q[i], rn);
}

// Phi(x,u,dt), base.py:450-480
// speed2: |v_b|^2 at the node x (before the update), a by-product of the first RK4 stage
template <class Sur>
AIRC_HD void state_update2_with(const DevConst& P, Sur& sur, State2& x, const Ctl2& c, float dt,
                                int substeps, bool normalise, F2& speed2) {
  const float h = (substeps == 1) ? dt : dt / (float)substeps;
  F2 sp2;
  rk4_step2(P, sur, x, c, h, speed2);
  for (int s = 1; s < substeps; ++s) rk4_step2(P, sur, x, c, h, sp2);
  if (normalise) normalise_q2(x);
}
AIRC_HD void state_update2(const DevConst& P, const UnitNet& net, State2& x, const Ctl2& c, float dt,
                           int substeps, bool normalise, F2& speed2) {
  SurFfma sur{net};
  state_update2_with(P, sur, x, c, dt, substeps, normalise, speed2);
}

// |v_b|^2 of the node for the speed term of the loss (control.py:63-64)
AIRC_HD F2 body_speed2_2(const DevConst& P, const State2& s) {
  F2 M[9], n2;
  rot_matrix2(s.q, M, n2);
  const F2 inv = rcp2_nr(n2);
  F2 acc = bc2(0.0f);
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    F2 a = mul2(M[i], s.v[0]);
    a = fma2(M[3 + i], s.v[1], a);
    a = fma2(M[6 + i], s.v[2], a);
    a = fma2(a, inv, bc2(P.eps));
    acc = fma2(a, a, acc);
  }
  return acc;
}

// smooth-L0 penalty of one control increment pair: 1 - exp(-d^2/eps)
AIRC_HD F2 smooth_l0_2(const F2& un, const F2& up, float neg_log2e_over_eps) {
  const F2 d = sub2(un, up);
  const F2 a = muls(neg_log2e_over_eps, mul2(d, d));
  return F2{1.0f - fast_ex2(a.lo), 1.0f - fast_ex2(a.hi)};
}

}  // namespace airc

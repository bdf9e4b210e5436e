// airc_jet.cuh -- second-order forward-mode arithmetic for the Lagrangian Hessian kernel (K4, jets formulation).
//
// What it replaces: IPOPT's nlp_hess_l with hessian_approximation 'exact'
// (/root/reference/src/aircraft/config.py:29), i.e. CasADi's second-order AD of the defect
// constraints x[k+1] - Phi(x[k],u[k],dt) (control/base.py:275-286) contracted with the multipliers:
//   H_k = sum_r lambda_k[r] d2 Phi_r / d z2 ,  z = (x, u, dt).
//
// A Jet carries, for ONE "outer" direction a and TWO packed "inner" directions (b0,b1):
//   v   value                     o   d/da
//   g   (d/db0, d/db1)            og  (d2/da db0, d2/da db1)
// so one evaluation of lambda^T Phi on jets yields two entries of H (and one of the gradient).
// The pair arithmetic is FFMA2 (fma.rn.f32x2), value and outer parts are scalar.
#pragma once
#include "airc_dynamics.cuh"
#include "airc_tangent.cuh"

namespace airc {

struct Jet {
  float v, o;
  F2 g, og;
  static AIRC_HD Jet cst(float c) { return Jet{c, 0.0f, F2{0.0f, 0.0f}, F2{0.0f, 0.0f}}; }
};

AIRC_HD Jet jconst(float c) { return Jet{c, 0.0f, F2{0.0f, 0.0f}, F2{0.0f, 0.0f}}; }

AIRC_HD Jet operator+(const Jet& a, const Jet& b) {
  return Jet{a.v + b.v, a.o + b.o, add2(a.g, b.g), add2(a.og, b.og)};
}
AIRC_HD Jet operator-(const Jet& a, const Jet& b) {
  return Jet{a.v - b.v, a.o - b.o, tsub(a.g, b.g), tsub(a.og, b.og)};
}
AIRC_HD Jet operator-(const Jet& a) { return Jet{-a.v, -a.o, tmul(-1.0f, a.g), tmul(-1.0f, a.og)}; }
AIRC_HD Jet operator+(const Jet& a, float c) { return Jet{a.v + c, a.o, a.g, a.og}; }
AIRC_HD Jet operator*(float s, const Jet& a) { return Jet{s * a.v, s * a.o, tmul(s, a.g), tmul(s, a.og)}; }
// s*a + b with a scalar (primal constant) s
AIRC_HD Jet jfma(float s, const Jet& a, const Jet& b) {
  return Jet{fmaf(s, a.v, b.v), fmaf(s, a.o, b.o), tfma(s, a.g, b.g), tfma(s, a.og, b.og)};
}
AIRC_HD Jet operator*(const Jet& a, const Jet& b) {
  Jet r;
  r.v = a.v * b.v;
  r.o = fmaf(a.o, b.v, a.v * b.o);
  r.g = tfma(a.v, b.g, tmul(b.v, a.g));
  r.og = tfma(a.o, b.g, tfma(b.o, a.g, tfma(a.v, b.og, tmul(b.v, a.og))));
  return r;
}
// a*b + c
AIRC_HD Jet jmad(const Jet& a, const Jet& b, const Jet& c) {
  Jet r;
  r.v = fmaf(a.v, b.v, c.v);
  r.o = fmaf(a.o, b.v, fmaf(a.v, b.o, c.o));
  r.g = tfma(a.v, b.g, tfma(b.v, a.g, c.g));
  r.og = tfma(a.o, b.g, tfma(b.o, a.g, tfma(a.v, b.og, tfma(b.v, a.og, c.og))));
  return r;
}
// phi(a) from (phi, phi', phi'') at a.v
AIRC_HD Jet jchain(const Jet& a, float f0, float f1, float f2) {
  return Jet{f0, f1 * a.o, tmul(f1, a.g), tfma(f2 * a.o, a.g, tmul(f1, a.og))};
}
// phi(y,x) from its value, gradient and Hessian at (y.v, x.v)
AIRC_HD Jet jchain2(const Jet& y, const Jet& x, float f0, float fy, float fx, float fyy, float fxy, float fxx) {
  Jet r;
  r.v = f0;
  r.o = fmaf(fy, y.o, fx * x.o);
  r.g = tfma(fy, y.g, tmul(fx, x.g));
  const float cy = fmaf(fyy, y.o, fxy * x.o);   // d(fy)/da
  const float cx = fmaf(fxy, y.o, fxx * x.o);   // d(fx)/da
  r.og = tfma(cy, y.g, tfma(cx, x.g, tfma(fy, y.og, tmul(fx, x.og))));
  return r;
}

AIRC_HD Jet jrcp(const Jet& a) {
  const float r = rcp_nr(a.v), r2 = r * r;
  return jchain(a, r, -r2, 2.0f * r2 * r);
}
AIRC_HD Jet jrsqrt(const Jet& a) {   // a^-1/2, a > 0
  float s, rs;
  sqrt_rsqrt(a.v, s, rs);
  const float r2 = rs * rs;
  return jchain(a, rs, -0.5f * rs * r2, 0.75f * rs * r2 * r2);
}
AIRC_HD Jet jasin(const Jet& a) {
  const float ic = fast_rsqrt(fmaf(-a.v, a.v, 1.0f));
  return jchain(a, asin_acc(a.v), ic, a.v * ic * ic * ic);
}
AIRC_HD Jet jatan2(const Jet& y, const Jet& x) {
  const float ir2 = rcp_nr(fmaf(x.v, x.v, y.v * y.v));
  const float fy = x.v * ir2, fx = -y.v * ir2;
  const float fyy = 2.0f * fx * fy;                       // -2xy/r^4
  const float fxy = (y.v * y.v - x.v * x.v) * ir2 * ir2;
  return jchain2(y, x, atan2_acc(y.v, x.v), fy, fx, fyy, fxy, -fyy);
}
// r = 1/(1+2^h)
AIRC_HD Jet jsigm2(const Jet& h) {
  const float r = sigm2(h.v);
  const float r1 = -0.69314718056f * r * (1.0f - r);
  const float r2 = -0.69314718056f * (1.0f - 2.0f * r) * r1;
  return jchain(h, r, r1, r2);
}
// 1/(1+exp(k(|a|-a0)))  (stall scaling, aircraft.py:280-294); |.| has derivative sign(a), second derivative 0
AIRC_HD Jet jstall(const Jet& a, float k, float a0) {
  const float s = 1.0f / (1.0f + expf(k * (fabsf(a.v) - a0)));
  const float sg = a.v < 0.0f ? -1.0f : 1.0f;
  const float s1 = -k * s * (1.0f - s);
  return jchain(a, s, s1 * sg, -k * (1.0f - 2.0f * s) * s1);
}

using CtlJ = CtlT<Jet>;   // CtlT: airc_tangent.cuh

// ---- the registry's other coefficient models on jets (the float versions are in airc_alt.cuh)
// outputs K0..K1-1 of the degree-3 polynomial table at x (sklearn monomial order), without storing the monomials
template <int K0, int K1>
AIRC_HD void poly_eval_jet(const float (*W)[NPOLY + 1], const Jet x[4], Jet* out) {
#pragma unroll
  for (int k = K0; k < K1; ++k) {
    Jet a = jconst(W[k][NPOLY]);
#pragma unroll
    for (int i = 0; i < 4; ++i) a = jfma(W[k][i], x[i], a);
    out[k - K0] = a;
  }
  int t2 = 4, t3 = 14;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
#pragma unroll
    for (int j = i; j < 4; ++j) {
      const Jet pij = x[i] * x[j];
#pragma unroll
      for (int k = K0; k < K1; ++k) out[k - K0] = jfma(W[k][t2], pij, out[k - K0]);
      ++t2;
#pragma unroll
      for (int l = j; l < 4; ++l) {
        const Jet pijl = pij * x[l];
#pragma unroll
        for (int k = K0; k < K1; ++k) out[k - K0] = jfma(W[k][t3], pijl, out[k - K0]);
        ++t3;
      }
    }
  }
}
AIRC_HD Jet poly_alpha_only_jet(const float w[NPOLY + 1], const Jet& a) {
  return jmad(a, jmad(a, jfma(w[14], a, jconst(w[4])), jconst(w[0])), jconst(w[NPOLY]));
}

// NetAlt::adjust on jets (coefficient_models.py:41-135); C excludes the rudder term f_jet adds afterwards
AIRC_HD void alt_coefficients_jet(const DevConst& P, const Jet vb[3], const Jet& qbar, const Jet& alpha,
                                  const Jet& beta, const Jet w[3], const CtlJ& c, Jet C[6]) {
  const AltParams& A = alt_params(P);
  if (P.model == AIRC_MODEL_POLY_) {
    const Jet xu = vb[0] + P.eps;
    const Jet a_el = jatan2(jfma(A.arm, w[1], vb[2]), xu);
    const Jet a_lw = jatan2(jfma(-0.25f * A.span, w[0], vb[2]), xu);
    const Jet a_rw = jatan2(jfma(0.25f * A.span, w[0], vb[2]), xu);
    const Jet vr = jfma(-A.arm, w[2], vb[1]);
    const Jet b_rud = jasin(vr * jrsqrt(jmad(vb[0], vb[0], jmad(vr, vr, vb[2] * vb[2])) + P.eps));
    const Jet x0[4] = {alpha, beta, c.da, c.de};
    poly_eval_jet<0, 4>(A.poly, x0, C);
    C[3] = jfma(0.125f * A.span, poly_alpha_only_jet(A.poly[2], a_rw) - poly_alpha_only_jet(A.poly[2], a_lw), C[3]);
    const Jet x1[4] = {a_el, beta, c.da, c.de};
    poly_eval_jet<4, 5>(A.poly, x1, C + 4);
    const Jet x2[4] = {alpha, b_rud, c.da, c.de};
    poly_eval_jet<5, 6>(A.poly, x2, C + 5);
  } else if (P.model == AIRC_MODEL_LINEAR_) {
#pragma unroll
    for (int k = 0; k < 6; ++k) {
      const float* L = A.lin[k];
      C[k] = jfma(L[0], qbar, jfma(L[1], alpha, jfma(L[2], beta, jfma(L[3], c.da, jfma(L[4], c.de, jconst(L[5]))))));
    }
  } else {
    const float d2r = 0.01745329252f;
    C[0] = -(0.3f * (alpha * alpha) + 0.02f);
    C[1] = -0.98f * beta;
    C[2] = -5.0f * alpha;
    C[3] = jfma(0.08f * 4.0f * d2r, c.da, -0.05f * w[0]);
    C[4] = jfma(-1.2f * 5.0f * d2r, c.de, -0.5f * w[1]);
    C[5] = -0.05f * w[2];
  }
  C[5] = jfma(A.k_rudder_model - P.k_rudder, c.dr, C[5]);
}

// f(x,u) on jets, the same rotation-matrix form as f_eval_net (base.py:385-406); p_dot = v is left
// to the caller.  MODEL 0: the NN surrogate; 1: the table models of airc_alt.cuh.
// J is the jet type: Jet (second order) or Dual (value + two first-order directions, airc_hess2.cuh); the
// table models exist on Jet only.
// NET evaluates the network model's coefficients: NetInlineJet = every thread the whole network on its jets;
// NetSharedJet (airc_hess.cuh) = the 64 threads of a node share the value-level sums.
struct NetInlineJet {
  template <class J>
  AIRC_HD void operator()(const DevConst& P, const J& da, const J& de, const J& qbar, const J& alpha, const J& beta,
                          J C[6]) const {
#pragma unroll
    for (int k = 0; k < 6; ++k) C[k] = J::cst(P.net.c0[k]);
#pragma unroll 2
    for (int j = 0; j < NH; ++j) {
      const HidPair& hp = P.net.hid[j >> 1];
      const int e = j & 1;
      J h = J::cst(hp.a0[e]);
      h = jfma(hp.Au[0][e], da, h);
      h = jfma(hp.Au[1][e], de, h);
      h = jfma(hp.Ax[0][e], qbar, h);
      h = jfma(hp.Ax[1][e], alpha, h);
      h = jfma(hp.Ax[2][e], beta, h);
      const J r = jsigm2(h);
#pragma unroll
      for (int k = 0; k < 6; ++k) C[k] = jfma(P.net.Wo[k][j], r, C[k]);
    }
  }
};

template <int MODEL = 0, class J = Jet, class NET = NetInlineJet>
AIRC_HD void f_jet(const DevConst& P, const J v[3], const J q[4], const J w[3], const CtlT<J>& c,
                   J vdot[3], J qdot[4], J wdot[3], const NET& net = NET()) {
  const J &x = q[0], &y = q[1], &z = q[2], &s = q[3];
  const J xx = x * x, yy = y * y, zz = z * z, ww = s * s;
  const J n2 = (xx + yy) + (zz + ww);
  J M[9];
  M[0] = (ww + xx) - (yy + zz);
  M[4] = (ww + yy) - (xx + zz);
  M[8] = (ww + zz) - (xx + yy);
  {
    const J xy = x * y, xz = x * z, yz = y * z, wx = s * x, wy = s * y, wz = s * z;
    M[1] = 2.0f * (xy - wz);
    M[2] = 2.0f * (xz + wy);
    M[3] = 2.0f * (xy + wz);
    M[5] = 2.0f * (yz - wx);
    M[6] = 2.0f * (xz - wy);
    M[7] = 2.0f * (yz + wx);
  }
  const J inv_n2 = jrcp(n2);
  J vb[3];
#pragma unroll
  for (int i = 0; i < 3; ++i)
    vb[i] = jmad(M[6 + i], v[2], jmad(M[3 + i], v[1], M[i] * v[0])) * inv_n2 + P.eps;   // base.py:156-159
  const J vb2 = jmad(vb[2], vb[2], jmad(vb[1], vb[1], vb[0] * vb[0]));
  const J qbar = 0.6125f * vb2;                                                       // base.py:240
  const J invV = jrsqrt(vb2 + P.eps);                                                 // base.py:167
  const J alpha = jatan2(vb[2], vb[0] + P.eps);                                       // base.py:176
  const J beta = jasin(vb[1] * invV);                                                 // base.py:234

  J C[6];
  if constexpr (MODEL == 0) {
    net(P, c.da, c.de, qbar, alpha, beta, C);
  } else {
    alt_coefficients_jet(P, vb, qbar, alpha, beta, w, c, C);
  }
  C[5] = jfma(P.k_rudder, c.dr, C[5]);                       // coefficient_models.py:102-103
  if (P.stall_scaling) {                                     // aircraft.py:280-294
    const J sa = jstall(alpha, 10.0f, 0.52359877559f), sbt = jstall(beta, 10.0f, 0.52359877559f);
    C[2] = C[2] * sa * sbt;
    C[4] = C[4] * sa;
  }
  C[0] = jfma(P.k_flap_x, c.df, C[0]);                       // aircraft.py:297-299
  C[2] = jfma(P.k_flap_z, c.df, C[2]);

  const J qS = P.S * qbar;
  J F[3];
  F[0] = sign_casadi(vb[0].v) * (C[0] * qS);                 // sign() has zero derivative, as in CasADi
  F[1] = C[1] * qS;
  F[2] = C[2] * qS;
  J Mo[3];
  Mo[0] = C[3] * (P.Sb * qbar) + (P.com[1] * F[2] - P.com[2] * F[1]);
  Mo[1] = C[4] * (P.Sc * qbar) + (P.com[2] * F[0] - P.com[0] * F[2]);
  Mo[2] = C[5] * (P.Sb * qbar) + (P.com[0] * F[1] - P.com[1] * F[0]);

  const J sc = P.inv_mass * inv_n2;                        // base.py:284-286,359
#pragma unroll
  for (int i = 0; i < 3; ++i)
    vdot[i] = jmad(M[3 * i + 2], F[2], jmad(M[3 * i + 1], F[1], M[3 * i] * F[0])) * sc + P.g[i];
  // q_dot = 0.5 q (x) (w,0)                                 base.py:292-295
  qdot[0] = 0.5f * (s * w[0] + (y * w[2] - z * w[1]));
  qdot[1] = 0.5f * (s * w[1] + (z * w[0] - x * w[2]));
  qdot[2] = 0.5f * (s * w[2] + (x * w[1] - y * w[0]));
  qdot[3] = -0.5f * (x * w[0] + y * w[1] + z * w[2]);
  // w_dot = I^-1 (M - w x (I w))                            base.py:376-377
  J Iw[3], t[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) Iw[i] = jfma(P.I[3 * i + 2], w[2], jfma(P.I[3 * i + 1], w[1], P.I[3 * i] * w[0]));
  t[0] = Mo[0] - (w[1] * Iw[2] - w[2] * Iw[1]);
  t[1] = Mo[1] - (w[2] * Iw[0] - w[0] * Iw[2]);
  t[2] = Mo[2] - (w[0] * Iw[1] - w[1] * Iw[0]);
#pragma unroll
  for (int i = 0; i < 3; ++i)
    wdot[i] = jfma(P.Iinv[3 * i + 2], t[2], jfma(P.Iinv[3 * i + 1], t[1], P.Iinv[3 * i] * t[0]));
}

}  // namespace airc

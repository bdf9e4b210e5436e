// airc_hess2.cuh -- K4: Hessian of the multiplier-weighted step by FORWARD-OVER-REVERSE differentiation.
//
// Replaces CasADi's second-order AD behind IPOPT's exact nlp_hess_l
// (/root/reference/src/aircraft/config.py:29; constraints control/base.py:275-286; the dt row and column
// belong to time='progress', control/base.py:276,306-307), like airc_hess.cuh, with a tenth of its arithmetic:
//
//   L(z) = lambda^T Phi(x, u, dt),  z = (x, u, dt);  g(z) = dL/dz by ONE hand-written reverse sweep through the
//   RK4 step (f_dual<true> below is f with its adjoint); H = dg/dz by running that sweep on DUAL numbers: every
//   quantity carries its value and its derivatives along TWO input directions, packed into the lanes of FFMA2.
//   15 directions are active (v q w | aileron elevator rudder flaps | dt), so EIGHT threads ("lanes") per node give
//   the whole 15 x 15 block (airc_hess.cuh needs 64 threads per node, each on second-order jets).  The network's
//   value-level sums are the same in the eight lanes, which share them through shuffles (net_linearise).
//
// The sweep keeps no tape: the forward pass stores the three stage derivatives k1..k3 of a substep, the reverse
// pass rebuilds each stage point from them and f_dual<true> recomputes f's interior.  Only the NN surrogate
// (model 0) has a hand-written adjoint; the table models stay on airc_hess.cuh, and so do short horizons (there
// the 64 threads of a node share the network, NetSharedJet).  Measured variants: profiles/r2_variants_hessian.txt.
#pragma once
#include <type_traits>
#include "airc_hess.cuh"

#ifndef AIRC_AB_H2_MINB            // build-variant A/B only
#define AIRC_AB_H2_MINB 1
#endif
#ifndef AIRC_AB_H2_BS              // 0: 256 threads per block for long batches, 128 below (shipped); else fixed
#define AIRC_AB_H2_BS 0
#endif

#ifndef AIRC_AB_H2_DIRS            // directions per lane: 2 (eight lanes per node, shipped) or 4 (four lanes)
#define AIRC_AB_H2_DIRS 2
#endif

namespace airc {

constexpr int H2_DIRS = AIRC_AB_H2_DIRS;
constexpr int HESS2_LANES = (15 + H2_DIRS - 1) / H2_DIRS;   // threads per node
static_assert(H2_DIRS == 2 || H2_DIRS == 4, "two or four directions per lane");

// two packed pairs, for the four-direction variant
struct F4 {
  F2 a, b;
};
AIRC_HD F4 add2(const F4& x, const F4& y) { return F4{add2(x.a, y.a), add2(x.b, y.b)}; }
AIRC_HD F4 tsub(const F4& x, const F4& y) { return F4{tsub(x.a, y.a), tsub(x.b, y.b)}; }
AIRC_HD F4 neg2(const F4& x) { return F4{neg2(x.a), neg2(x.b)}; }
AIRC_HD F4 tmul(float s, const F4& x) { return F4{tmul(s, x.a), tmul(s, x.b)}; }
AIRC_HD F4 tfma(float s, const F4& x, const F4& y) { return F4{tfma(s, x.a, y.a), tfma(s, x.b, y.b)}; }
template <class T> struct TanOps;
template <> struct TanOps<F2> {
  static AIRC_HD F2 zero() { return F2{0.0f, 0.0f}; }
  static AIRC_HD F2 unit(int slot, int first) { return F2{slot == first ? 1.0f : 0.0f, slot == first + 1 ? 1.0f : 0.0f}; }
  static AIRC_HD float get(const F2& t, int i) { return i ? t.hi : t.lo; }
};
template <> struct TanOps<F4> {
  static AIRC_HD F4 zero() { return F4{F2{0.0f, 0.0f}, F2{0.0f, 0.0f}}; }
  static AIRC_HD F4 unit(int slot, int first) { return F4{TanOps<F2>::unit(slot, first), TanOps<F2>::unit(slot, first + 2)}; }
  static AIRC_HD float get(const F4& t, int i) { return i < 2 ? TanOps<F2>::get(t.a, i) : TanOps<F2>::get(t.b, i - 2); }
};
using Tan = std::conditional<H2_DIRS == 2, F2, F4>::type;

// value + derivatives along H2_DIRS directions
struct Dual {
  float v;
  Tan t;
  static AIRC_HD Dual cst(float c) { return Dual{c, TanOps<Tan>::zero()}; }
};

AIRC_HD Dual operator+(const Dual& a, const Dual& b) { return Dual{a.v + b.v, add2(a.t, b.t)}; }
AIRC_HD Dual operator-(const Dual& a, const Dual& b) { return Dual{a.v - b.v, tsub(a.t, b.t)}; }
AIRC_HD Dual operator-(const Dual& a) { return Dual{-a.v, neg2(a.t)}; }
AIRC_HD Dual operator+(const Dual& a, float c) { return Dual{a.v + c, a.t}; }
AIRC_HD Dual operator*(float s, const Dual& a) { return Dual{s * a.v, tmul(s, a.t)}; }
AIRC_HD Dual operator*(const Dual& a, const Dual& b) {
  return Dual{a.v * b.v, tfma(a.v, b.t, tmul(b.v, a.t))};
}
// s*a + b with a plain number s
AIRC_HD Dual jfma(float s, const Dual& a, const Dual& b) { return Dual{fmaf(s, a.v, b.v), tfma(s, a.t, b.t)}; }
// a*b + c
AIRC_HD Dual jmad(const Dual& a, const Dual& b, const Dual& c) {
  return Dual{fmaf(a.v, b.v, c.v), tfma(a.v, b.t, tfma(b.v, a.t, c.t))};
}
// c - a*b
AIRC_HD Dual jnmad(const Dual& a, const Dual& b, const Dual& c) {
  return Dual{fmaf(-a.v, b.v, c.v), tfma(-a.v, b.t, tfma(-b.v, a.t, c.t))};
}
// phi(a) from (phi, phi') at a.v
AIRC_HD Dual dchain(const Dual& a, float f0, float f1) { return Dual{f0, tmul(f1, a.t)}; }

AIRC_HD Dual jrcp(const Dual& a) {
  const float r = rcp_nr(a.v);
  return dchain(a, r, -r * r);
}
AIRC_HD Dual jrsqrt(const Dual& a) {   // a^-1/2, a > 0
  float s, rs;
  sqrt_rsqrt(a.v, s, rs);
  return dchain(a, rs, -0.5f * rs * rs * rs);
}
AIRC_HD Dual jasin(const Dual& a) { return dchain(a, asin_acc(a.v), fast_rsqrt(fmaf(-a.v, a.v, 1.0f))); }
AIRC_HD Dual jatan2(const Dual& y, const Dual& x) {
  const float ir2 = rcp_nr(fmaf(x.v, x.v, y.v * y.v));
  return Dual{atan2_acc(y.v, x.v), tfma(x.v * ir2, y.t, tmul(-y.v * ir2, x.t))};
}
// r = 1/(1+2^h)
AIRC_HD Dual jsigm2(const Dual& h) {
  const float r = sigm2(h.v);
  return dchain(h, r, -0.69314718056f * r * (1.0f - r));
}
// 1/(1+exp(k(|a|-a0)))  (stall scaling, aircraft.py:280-294); |.| has derivative sign(a)
AIRC_HD Dual jstall(const Dual& a, float k, float a0) {
  const float s = 1.0f / (1.0f + expf(k * (fabsf(a.v) - a0)));
  const float sg = a.v < 0.0f ? -1.0f : 1.0f;
  return dchain(a, s, -k * s * (1.0f - s) * sg);
}
// d/da of jstall as a dual number, from s = jstall(a)
AIRC_HD Dual dstall_slope(const Dual& a, const Dual& s, float k) {
  const float sg = a.v < 0.0f ? -1.0f : 1.0f;
  return (-k * sg) * (s * (-s + 1.0f));
}

struct DState {   // the ten state entries f reads: v (0..2), q (3..6), w (7..9); p never enters f
  Dual e[10];
};
using DCtl = CtlT<Dual>;

AIRC_HD void dcross(const Dual a[3], const Dual b[3], Dual o[3]) {
  o[0] = jnmad(a[2], b[1], a[1] * b[2]);
  o[1] = jnmad(a[0], b[2], a[2] * b[0]);
  o[2] = jnmad(a[1], b[0], a[0] * b[1]);
}

// ---- the network, shared by the eight lanes of a node -----------------------------------------------------
// The network's VALUE-level quantities are the same in all eight lanes (the lanes differ in their directions
// only), so each lane visits FOUR of the 32 hidden units and the lanes add their shares:
//   C = c0 + Wo r(h),   J = dC/dy = Wo diag(r') A   (6 x 5),   G = sum_j (Wo^T Cb)_j r''_j a_j a_j^T   (5 x 5, symmetric)
// for the inputs y = (aileron, elevator, qbar, alpha, beta), h = a0 + A y.  With them the dual parts are small
// products:  C.t = J y.t,   yb = J^T Cb,   yb.t = J^T Cb.t + G y.t   (G is the second derivative of Cb^T C).
constexpr int NET_UNITS = NH / HESS2_LANES;  // hidden units per lane
constexpr int NET_TAB_FLOATS = NH * 12;      // [unit of the lane][lane][a0, a(5), Wo(6)]
constexpr int NET_NC = 6, NET_NJ = 30, NET_NG = 15;

// the weight image a lane reads its units from (shared memory on the device: conflict-free LDS.128)
AIRC_HD void net_tab_entry(const DevConst& P, int idx, float* tab) {   // idx in [0, 32*12)
  const int f = idx % 12, sub = (idx / 12) % HESS2_LANES, loc = idx / (12 * HESS2_LANES);
  const int j = NET_UNITS * sub + loc;
  const HidPair& hp = P.net.hid[j >> 1];
  const int e = j & 1;
  float v;
  if (f == 0) v = hp.a0[e];
  else if (f < 3) v = hp.Au[f - 1][e];
  else if (f < 6) v = hp.Ax[f - 3][e];
  else v = P.net.Wo[f - 6][j];
  tab[idx] = v;
}

// the share of lane `sub`: acc[0..5] C, acc[6 + 5 k + i] J[k][i], acc[36..50] G (upper triangle, row-major)
template <bool ADJ>
AIRC_HD void net_share(const float* tab, int sub, const float y[5], const float Cbv[6], float* acc) {
  constexpr int NACC = NET_NC + NET_NJ + (ADJ ? NET_NG : 0);
#pragma unroll
  for (int i = 0; i < NACC; ++i) acc[i] = 0.0f;
#pragma unroll
  for (int loc = 0; loc < NET_UNITS; ++loc) {
    struct alignas(16) Q4 { float x, y, z, w; };   // three 16-byte loads
    const Q4* wp = reinterpret_cast<const Q4*>(tab + (loc * HESS2_LANES + sub) * 12);
    const Q4 w0 = wp[0], w1 = wp[1], w2 = wp[2];
    const float a[5] = {w0.y, w0.z, w0.w, w1.x, w1.y};
    const float wo[6] = {w1.z, w1.w, w2.x, w2.y, w2.z, w2.w};
    float h = w0.x;
#pragma unroll
    for (int i = 0; i < 5; ++i) h = fmaf(a[i], y[i], h);
    const float r = sigm2(h);
    const float r1 = -0.69314718056f * r * (1.0f - r);
    float m[5];
#pragma unroll
    for (int i = 0; i < 5; ++i) m[i] = r1 * a[i];
#pragma unroll
    for (int k = 0; k < 6; ++k) {
      acc[k] = fmaf(wo[k], r, acc[k]);
#pragma unroll
      for (int i = 0; i < 5; ++i) acc[NET_NC + 5 * k + i] = fmaf(wo[k], m[i], acc[NET_NC + 5 * k + i]);
    }
    if constexpr (ADJ) {
      float rb = wo[0] * Cbv[0];
#pragma unroll
      for (int k = 1; k < 6; ++k) rb = fmaf(wo[k], Cbv[k], rb);
      const float g = rb * (-0.69314718056f * (1.0f - 2.0f * r));   // (Wo^T Cb)_j r''_j / r'_j
      int e = NET_NC + NET_NJ;
#pragma unroll
      for (int i = 0; i < 5; ++i) {
        const float t = g * m[i];
#pragma unroll
        for (int l = i; l < 5; ++l, ++e) acc[e] = fmaf(t, a[l], acc[e]);
      }
    }
  }
}

// C, J, G of the whole network in every lane.  Device: butterfly over the eight lanes of the node (the sums are
// the same bits in all of them: a + b == b + a).  Host: the eight shares one after the other, added in the
// butterfly's order.
template <bool ADJ>
AIRC_HD void net_linearise(const DevConst& P, const float* tab, int sub, const float y[5], const float Cbv[6],
                           float* acc) {
  constexpr int NACC = NET_NC + NET_NJ + (ADJ ? NET_NG : 0);
#if defined(__CUDA_ARCH__)
  net_share<ADJ>(tab, sub, y, Cbv, acc);
#pragma unroll
  for (int d = 1; d < HESS2_LANES; d <<= 1)
#pragma unroll
    for (int i = 0; i < NACC; ++i) acc[i] += __shfl_xor_sync(0xffffffffu, acc[i], d);
#else
  (void)sub;
  float part[HESS2_LANES][NACC];
  for (int l = 0; l < HESS2_LANES; ++l) net_share<ADJ>(tab, l, y, Cbv, part[l]);
  for (int d = 1; d < HESS2_LANES; d <<= 1)      // the butterfly: after level d, part[l] holds the sum of its group of 2d
    for (int l = 0; l < HESS2_LANES; ++l)
      if (!(l & d))
        for (int i = 0; i < NACC; ++i) part[l][i] = part[l | d][i] = part[l][i] + part[l | d][i];
  for (int i = 0; i < NACC; ++i) acc[i] = part[0][i];
#endif
#pragma unroll
  for (int k = 0; k < 6; ++k) acc[k] += P.net.c0[k];
}

// f on dual numbers, k = f(X, c) (p_dot = v left to the caller; base.py:385-406, rotation-matrix form as
// f_jet), and with ADJ its adjoint for the weights kb on k:
//   Xb = (df/dX)^T kb  (overwritten),   cb += (df/dc)^T kb.
// The reverse half walks the forward half backwards.  The coefficient adjoints Cb do not depend on the
// coefficients themselves, so the network is visited once, between the two halves.
template <bool ADJ>
AIRC_HD void f_dual(const DevConst& P, const float* tab, int sub, const DState& X, const DCtl& c, const DState& kb,
                    DState& k, DState& Xb, DCtl& cb) {
  const Dual &x = X.e[3], &y = X.e[4], &z = X.e[5], &s = X.e[6];
  const Dual *v = X.e, *w = X.e + 7;
  const Dual *kbv = kb.e, *kbq = kb.e + 3, *kbw = kb.e + 7;
  Dual *kv = k.e, *kq = k.e + 3, *kw = k.e + 7;
  Dual *Xbv = Xb.e, *Xbq = Xb.e + 3, *Xbw = Xb.e + 7;
  // ---- forward, up to the network inputs
  const Dual xx = x * x, yy = y * y, zz = z * z, ww = s * s;
  const Dual n2 = (xx + yy) + (zz + ww);
  Dual M[9];
  M[0] = (ww + xx) - (yy + zz);
  M[4] = (ww + yy) - (xx + zz);
  M[8] = (ww + zz) - (xx + yy);
  {
    const Dual xy = x * y, xz = x * z, yz = y * z, wx = s * x, wy = s * y, wz = s * z;
    M[1] = 2.0f * (xy - wz);
    M[2] = 2.0f * (xz + wy);
    M[3] = 2.0f * (xy + wz);
    M[5] = 2.0f * (yz - wx);
    M[6] = 2.0f * (xz - wy);
    M[7] = 2.0f * (yz + wx);
  }
  const Dual inv_n2 = jrcp(n2);
  Dual uu[3], vb[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    uu[i] = jmad(M[6 + i], v[2], jmad(M[3 + i], v[1], M[i] * v[0]));
    vb[i] = uu[i] * inv_n2 + P.eps;
  }
  const Dual vb2 = jmad(vb[2], vb[2], jmad(vb[1], vb[1], vb[0] * vb[0]));
  const Dual qbar = 0.6125f * vb2;
  const Dual invV = jrsqrt(vb2 + P.eps);
  const Dual xa = vb[0] + P.eps;
  const Dual alpha = jatan2(vb[2], xa);
  const Dual sbeta = vb[1] * invV;
  const Dual beta = jasin(sbeta);
  const float sgn = sign_casadi(vb[0].v);
  const Dual qS = P.S * qbar, qSb = P.Sb * qbar, qSc = P.Sc * qbar;
  const Dual sc = P.inv_mass * inv_n2;
  Dual sa = Dual::cst(1.0f), sbt = Dual::cst(1.0f);
  if (P.stall_scaling) {
    sa = jstall(alpha, 10.0f, 0.52359877559f);
    sbt = jstall(beta, 10.0f, 0.52359877559f);
  }

  // ---- adjoints that need no coefficient: w_dot, q_dot, and the way from v_dot / the moments to Cb
  Dual tb[3], Iw[3];   // tb = adjoint of t = Mo - w x (I w), also the adjoint of Mo
#pragma unroll
  for (int j = 0; j < 3; ++j) Iw[j] = jfma(P.I[3 * j + 2], w[2], jfma(P.I[3 * j + 1], w[1], P.I[3 * j] * w[0]));
  Dual qb[4], Gb[3], Fb[3], Cb[6], Cb2f, Cb4s;
  float Cbv[6] = {0.0f, 0.0f, 0.0f, 0.0f, 0.0f, 0.0f};
  if constexpr (ADJ) {
#pragma unroll
    for (int j = 0; j < 3; ++j)
      tb[j] = jfma(P.Iinv[6 + j], kbw[2], jfma(P.Iinv[3 + j], kbw[1], P.Iinv[j] * kbw[0]));
    {
      // d(w x Iw).(-tb) = dw.(tb x Iw) + dw.I^T (w x tb)
      Dual e[3], f[3];
      dcross(tb, Iw, e);
      dcross(w, tb, f);
#pragma unroll
      for (int i = 0; i < 3; ++i) Xbw[i] = e[i] + jfma(P.I[6 + i], f[2], jfma(P.I[3 + i], f[1], P.I[i] * f[0]));
    }
    {
      const Dual a0 = 0.5f * kbq[0], a1 = 0.5f * kbq[1], a2 = 0.5f * kbq[2], a3 = 0.5f * kbq[3];
      qb[0] = jnmad(a3, w[0], jnmad(a1, w[2], a2 * w[1]));
      qb[1] = jnmad(a3, w[1], jnmad(a2, w[0], a0 * w[2]));
      qb[2] = jnmad(a3, w[2], jnmad(a0, w[1], a1 * w[0]));
      qb[3] = jmad(a2, w[2], jmad(a1, w[1], a0 * w[0]));
      Xbw[0] = Xbw[0] + jnmad(a3, x, jnmad(a2, y, jmad(a1, z, a0 * s)));
      Xbw[1] = Xbw[1] + jnmad(a3, y, jnmad(a0, z, jmad(a2, x, a1 * s)));
      Xbw[2] = Xbw[2] + jnmad(a3, z, jnmad(a1, x, jmad(a0, y, a2 * s)));
    }
#pragma unroll
    for (int i = 0; i < 3; ++i) Gb[i] = kbv[i] * sc;
#pragma unroll
    for (int j = 0; j < 3; ++j) Fb[j] = jmad(M[6 + j], Gb[2], jmad(M[3 + j], Gb[1], M[j] * Gb[0]));
    Fb[0] = Fb[0] + (P.com[2] * tb[1] - P.com[1] * tb[2]);
    Fb[1] = Fb[1] + (P.com[0] * tb[2] - P.com[2] * tb[0]);
    Fb[2] = Fb[2] + (P.com[1] * tb[0] - P.com[0] * tb[1]);
    Cb[0] = sgn * (Fb[0] * qS);
    Cb[1] = Fb[1] * qS;
    Cb2f = Fb[2] * qS;       // adjoint of C2 after stall scaling (and of the flap term)
    Cb[3] = tb[0] * qSb;
    Cb4s = tb[1] * qSc;      // adjoint of C4 after stall scaling
    Cb[5] = tb[2] * qSb;
    Cb[2] = P.stall_scaling ? Cb2f * (sa * sbt) : Cb2f;
    Cb[4] = P.stall_scaling ? Cb4s * sa : Cb4s;
#pragma unroll
    for (int kk = 0; kk < 6; ++kk) Cbv[kk] = Cb[kk].v;
  }

  // ---- the network: value-level C, J, G from the eight lanes, dual parts by small products
  Dual C[6], yb[5];
  {
    const Dual* yin[5] = {&c.da, &c.de, &qbar, &alpha, &beta};
    const float yv[5] = {c.da.v, c.de.v, qbar.v, alpha.v, beta.v};
    float acc[NET_NC + NET_NJ + (ADJ ? NET_NG : 0)];
    net_linearise<ADJ>(P, tab, sub, yv, Cbv, acc);
    const float* J = acc + NET_NC;
#pragma unroll
    for (int kk = 0; kk < 6; ++kk) {
      Tan t = tmul(J[5 * kk], yin[0]->t);
#pragma unroll
      for (int i = 1; i < 5; ++i) t = tfma(J[5 * kk + i], yin[i]->t, t);
      C[kk] = Dual{acc[kk], t};
    }
    if constexpr (ADJ) {
      const float* G = acc + NET_NC + NET_NJ;
#pragma unroll
      for (int i = 0; i < 5; ++i) {
        Dual a = J[i] * Cb[0];
#pragma unroll
        for (int kk = 1; kk < 6; ++kk) a = jfma(J[5 * kk + i], Cb[kk], a);
#pragma unroll
        for (int l = 0; l < 5; ++l) {
          const int lo = i < l ? i : l, hi = i < l ? l : i;
          a.t = tfma(G[lo * 5 - (lo * (lo - 1)) / 2 + (hi - lo)], yin[l]->t, a.t);
        }
        yb[i] = a;
      }
    }
  }
  Dual qbarb, alphab, betab;
  if constexpr (ADJ) {
    cb.da = cb.da + yb[0];
    cb.de = cb.de + yb[1];
    qbarb = yb[2]; alphab = yb[3]; betab = yb[4];
  }

  // ---- forward from the coefficients to k, with the adjoints that need the coefficients
  C[5] = jfma(P.k_rudder, c.dr, C[5]);                       // coefficient_models.py:102-103
  if constexpr (ADJ) cb.dr = jfma(P.k_rudder, Cb[5], cb.dr);
  if (P.stall_scaling) {                                     // aircraft.py:280-294
    if constexpr (ADJ) {
      const Dual sab = jmad(Cb2f * C[2], sbt, Cb4s * C[4]);
      const Dual sbtb = (Cb2f * C[2]) * sa;
      alphab = jmad(sab, dstall_slope(alpha, sa, 10.0f), alphab);
      betab = jmad(sbtb, dstall_slope(beta, sbt, 10.0f), betab);
    }
    C[2] = C[2] * (sa * sbt);
    C[4] = C[4] * sa;
  }
  C[0] = jfma(P.k_flap_x, c.df, C[0]);                       // aircraft.py:297-299
  C[2] = jfma(P.k_flap_z, c.df, C[2]);
  if constexpr (ADJ) cb.df = jfma(P.k_flap_x, Cb[0], jfma(P.k_flap_z, Cb2f, cb.df));
  Dual F[3];
  F[0] = sgn * (C[0] * qS);
  F[1] = C[1] * qS;
  F[2] = C[2] * qS;
  if constexpr (ADJ) {   // q_bar enters through qS (forces) and through the three moment products
    qbarb = jfma(P.S, jmad(Fb[2], C[2], jmad(Fb[1], C[1], sgn * (Fb[0] * C[0]))), qbarb);
    qbarb = jfma(P.Sb, jmad(tb[2], C[5], tb[0] * C[3]), jfma(P.Sc, tb[1] * C[4], qbarb));
  }
  Dual Mo[3];
  Mo[0] = C[3] * qSb + (P.com[1] * F[2] - P.com[2] * F[1]);
  Mo[1] = C[4] * qSc + (P.com[2] * F[0] - P.com[0] * F[2]);
  Mo[2] = C[5] * qSb + (P.com[0] * F[1] - P.com[1] * F[0]);
  Dual Mb[9];   // adjoint of the rotation matrix
  Dual scb = Dual::cst(0.0f);
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    const Dual G = jmad(M[3 * i + 2], F[2], jmad(M[3 * i + 1], F[1], M[3 * i] * F[0]));
    kv[i] = G * sc + P.g[i];
    if constexpr (ADJ) {
      scb = jmad(kbv[i], G, scb);
#pragma unroll
      for (int j = 0; j < 3; ++j) Mb[3 * i + j] = Gb[i] * F[j];
    }
  }
  kq[0] = 0.5f * (s * w[0] + (y * w[2] - z * w[1]));
  kq[1] = 0.5f * (s * w[1] + (z * w[0] - x * w[2]));
  kq[2] = 0.5f * (s * w[2] + (x * w[1] - y * w[0]));
  kq[3] = -0.5f * (x * w[0] + y * w[1] + z * w[2]);
  {
    Dual cr[3], t[3];
    dcross(w, Iw, cr);
#pragma unroll
    for (int i = 0; i < 3; ++i) t[i] = Mo[i] - cr[i];
#pragma unroll
    for (int i = 0; i < 3; ++i)
      kw[i] = jfma(P.Iinv[3 * i + 2], t[2], jfma(P.Iinv[3 * i + 1], t[1], P.Iinv[3 * i] * t[0]));
  }
  if constexpr (!ADJ) return;

  // ---- backward from the network inputs to v and q
  Dual vbb[3];
  {
    const Dual sbb = betab * jrsqrt(jnmad(sbeta, sbeta, Dual::cst(1.0f)));   // beta = asin(sbeta)
    const Dual invVb = sbb * vb[1];
    const Dual vb2b = jfma(0.6125f, qbarb, invVb * (-0.5f * (invV * (invV * invV))));
    const Dual ir2 = jrcp(jmad(xa, xa, vb[2] * vb[2]));
    const Dual ai = alphab * ir2;
    const Dual two_n = 2.0f * vb2b;
    vbb[0] = jnmad(ai, vb[2], two_n * vb[0]);
    vbb[1] = jmad(sbb, invV, two_n * vb[1]);
    vbb[2] = jmad(ai, xa, two_n * vb[2]);
  }
  Dual inv_n2b = P.inv_mass * scb;
  Dual ub[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    ub[i] = vbb[i] * inv_n2;
    inv_n2b = jmad(vbb[i], uu[i], inv_n2b);
  }
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    Xbv[j] = jmad(M[3 * j + 2], ub[2], jmad(M[3 * j + 1], ub[1], M[3 * j] * ub[0]));
#pragma unroll
    for (int i = 0; i < 3; ++i) Mb[3 * j + i] = jmad(ub[i], v[j], Mb[3 * j + i]);
  }
  const Dual n2b = -(inv_n2b * (inv_n2 * inv_n2));
  {
    const Dual A = Mb[1] + Mb[3], Bm = Mb[3] - Mb[1], Cc = Mb[2] + Mb[6], D = Mb[2] - Mb[6], E = Mb[5] + Mb[7],
               Fm = Mb[7] - Mb[5];
    const Dual dx = (Mb[0] - Mb[4]) - Mb[8] + n2b, dy = (Mb[4] - Mb[0]) - Mb[8] + n2b,
               dz = (Mb[8] - Mb[0]) - Mb[4] + n2b, ds = (Mb[0] + Mb[4]) + Mb[8] + n2b;
    Xbq[0] = qb[0] + 2.0f * jmad(x, dx, jmad(y, A, jmad(z, Cc, s * Fm)));
    Xbq[1] = qb[1] + 2.0f * jmad(y, dy, jmad(x, A, jmad(z, E, s * D)));
    Xbq[2] = qb[2] + 2.0f * jmad(z, dz, jmad(x, Cc, jmad(y, E, s * Bm)));
    Xbq[3] = qb[3] + 2.0f * jmad(s, ds, jmad(z, Bm, jmad(y, D, x * Fm)));
  }
}

constexpr int HESS2_MAXSUB = 8;    // substeps whose start states the reverse sweep keeps; beyond: airc_hess.cuh

// One RK4 substep forward on duals (base.py:408-480): k1..k3 into ks, and with `advance` also k4 and x <- x+.
AIRC_HD void rk4_forward_dual(const DevConst& P, const float* tab, int sub, DState& x, const DCtl& c, const Dual& h,
                              DState ks[3], bool advance) {
  const Dual hh = 0.5f * h;
  DState X = x, acc;
  const int stages = advance ? 4 : 3;
#pragma unroll 1
  for (int stage = 0; stage < stages; ++stage) {
    DState k, none;
    DCtl nonec;
    f_dual<false>(P, tab, sub, X, c, none, k, none, nonec);
    const float wgt = (stage == 0 || stage == 3) ? 1.0f : 2.0f;
    const Dual cn = (stage == 2) ? h : hh;
#pragma unroll
    for (int i = 0; i < 10; ++i) {
      acc.e[i] = stage ? jfma(wgt, k.e[i], acc.e[i]) : k.e[i];
      X.e[i] = jmad(cn, k.e[i], x.e[i]);
    }
    if (stage < 3) ks[stage] = k;
  }
  if (advance) {
    const Dual h6 = (1.0f / 6.0f) * h;
#pragma unroll
    for (int i = 0; i < 10; ++i) x.e[i] = jmad(h6, acc.e[i], x.e[i]);
  }
}

// The same substep backwards.  In: x (its start state), ks (k1..k3), xb = weights on x+, lp = lambda_p (the p
// rows: p+ = p + h/6 (v1 + 2 v2 + 2 v3 + v4), f never reads p).  Out: xb = weights on x, cb / hb accumulated,
// ksum = k1 + 2 k2 + 2 k3 + k4 and vsum = v1 + 2 v2 + 2 v3 + v4 for the value of the step.
AIRC_HD void rk4_reverse_dual(const DevConst& P, const float* tab, int sub, const DState& x, const DCtl& c,
                              const Dual& h, const DState ks[3], const float lp[3], DState& xb, DCtl& cb, Dual& hb,
                              DState& ksum, Dual vsum[3]) {
  const Dual hh = 0.5f * h, h6 = (1.0f / 6.0f) * h;
  DState Xb, xacc = xb;
#pragma unroll 1
  for (int stage = 3; stage >= 0; --stage) {
    const float wgt = (stage == 0 || stage == 3) ? 1.0f : 2.0f;
    DState X = x;
    if (stage > 0) {
      const Dual cn = (stage == 3) ? h : hh;
#pragma unroll
      for (int i = 0; i < 10; ++i) X.e[i] = jmad(cn, ks[stage - 1].e[i], x.e[i]);
    }
    // weights on k_stage: through x+ and through the next stage point X_{stage+1} = x + cnx k_stage
    const Dual hw = wgt * h6;
    const Dual cnx = (stage == 2) ? h : hh;
    DState kb;
#pragma unroll
    for (int i = 0; i < 10; ++i) {
      kb.e[i] = hw * xb.e[i];
      if (stage < 3) kb.e[i] = jmad(cnx, Xb.e[i], kb.e[i]);
    }
    DState k;
    f_dual<true>(P, tab, sub, X, c, kb, k, Xb, cb);
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      Xb.e[i] = jfma(lp[i], hw, Xb.e[i]);
      vsum[i] = (stage == 3) ? X.e[i] : jfma(wgt, X.e[i], vsum[i]);
    }
    if (stage > 0) {   // dX_stage / dh = (1/2 or 1) k_{stage-1}
      Dual d = Xb.e[0] * ks[stage - 1].e[0];
#pragma unroll
      for (int i = 1; i < 10; ++i) d = jmad(Xb.e[i], ks[stage - 1].e[i], d);
      hb = jfma(stage == 3 ? 1.0f : 0.5f, d, hb);
    }
#pragma unroll
    for (int i = 0; i < 10; ++i) {
      ksum.e[i] = (stage == 3) ? k.e[i] : jfma(wgt, k.e[i], ksum.e[i]);
      xacc.e[i] = xacc.e[i] + Xb.e[i];
    }
  }
  {   // dx+ / dh = ksum / 6,  dp+ / dh = vsum / 6
    Dual d = lp[0] * vsum[0];
    d = jfma(lp[1], vsum[1], d);
    d = jfma(lp[2], vsum[2], d);
#pragma unroll
    for (int i = 0; i < 10; ++i) d = jmad(xb.e[i], ksum.e[i], d);
    hb = jfma(1.0f / 6.0f, d, hb);
  }
  xb = xacc;
}

// One of the eight lanes of a node.  g[i].v = dL/dz_i and g[i].t = (H[i][j0], H[i][j0 + 1]) over the 15 active
// directions i, j0 = 2 lane (direction 15 on the last lane: unused); value = lambda^T Phi.
// tab: the weight image of net_tab_entry.  On the device the eight lanes of a node run in lockstep (the network
// sums cross them); the host runs them one after the other.
AIRC_HD void hessian2_lane(const DevConst& P, const float* tab, const State& x0, const Ctl& c0, float dt, int substeps,
                           bool normalise, const float lam[13], int lane, Dual g[NHACT], float& value) {
  const int j0 = H2_DIRS * lane;
  auto seed = [&](float val, int slot) { return Dual{val, TanOps<Tan>::unit(slot, j0)}; };
  DState x;
#pragma unroll
  for (int i = 0; i < 3; ++i) { x.e[i] = seed(x0.v[i], i); x.e[7 + i] = seed(x0.w[i], 7 + i); }
#pragma unroll
  for (int i = 0; i < 4; ++i) x.e[3 + i] = seed(x0.q[i], 3 + i);
  const DCtl c{seed(c0.da, 10), seed(c0.de, 11), seed(c0.dr, 12), seed(c0.df, 13)};
  const float inv_s = 1.0f / (float)substeps;
  const Dual h = inv_s * seed(dt, 14);

  // forward: the start state of every substep; the last substep only when its end point is needed
  DState xs[HESS2_MAXSUB], ks[3];
  for (int sub = 0; sub < substeps; ++sub) {
    if (substeps > 1) xs[sub] = x;
    if (sub + 1 < substeps || normalise) rk4_forward_dual(P, tab, lane, x, c, h, ks, true);
  }
  DState xb;
#pragma unroll
  for (int i = 0; i < 10; ++i) xb.e[i] = Dual::cst(lam[3 + i]);
  float rn = 1.0f;
  if (normalise) {   // q+ = q / |q| (base.py:443-444, 473-474)
    const Dual* q = x.e + 3;
    const Dual rnd = jrsqrt(jmad(q[0], q[0], q[1] * q[1]) + jmad(q[2], q[2], q[3] * q[3]));
    rn = rnd.v;
    Dual qn[4], d = Dual::cst(0.0f);
#pragma unroll
    for (int i = 0; i < 4; ++i) { qn[i] = q[i] * rnd; d = jfma(lam[6 + i], qn[i], d); }
#pragma unroll
    for (int i = 0; i < 4; ++i) xb.e[3 + i] = rnd * jnmad(d, qn[i], Dual::cst(lam[6 + i]));
  }

  DCtl cb{Dual::cst(0.0f), Dual::cst(0.0f), Dual::cst(0.0f), Dual::cst(0.0f)};
  Dual hb = Dual::cst(0.0f);
  const float lp[3] = {lam[0], lam[1], lam[2]};
  value = fmaf(lp[2], x0.p[2], fmaf(lp[1], x0.p[1], lp[0] * x0.p[0]));
  for (int sub = substeps - 1; sub >= 0; --sub) {
    if (substeps > 1) x = xs[sub];
    else if (normalise) {   // the forward pass moved x to the end point
#pragma unroll
      for (int i = 0; i < 3; ++i) { x.e[i] = seed(x0.v[i], i); x.e[7 + i] = seed(x0.w[i], 7 + i); }
#pragma unroll
      for (int i = 0; i < 4; ++i) x.e[3 + i] = seed(x0.q[i], 3 + i);
    }
    if (!(normalise && sub == substeps - 1)) rk4_forward_dual(P, tab, lane, x, c, h, ks, false);
    DState ksum;
    Dual vsum[3];
    const bool last = sub == substeps - 1;
    float lx[10];
#pragma unroll
    for (int i = 0; i < 10; ++i) lx[i] = lam[3 + i];
    rk4_reverse_dual(P, tab, lane, x, c, h, ks, lp, xb, cb, hb, ksum, vsum);
    const float h6 = h.v * (1.0f / 6.0f);
    value = fmaf(h6, fmaf(lp[2], vsum[2].v, fmaf(lp[1], vsum[1].v, lp[0] * vsum[0].v)), value);
    if (last) {
#pragma unroll
      for (int i = 0; i < 10; ++i) {
        const float xp = fmaf(h6, ksum.e[i].v, x.e[i].v);
        value = fmaf(lx[i], (i >= 3 && i < 7) ? xp * rn : xp, value);
      }
    }
  }
#pragma unroll
  for (int i = 0; i < 10; ++i) g[i] = xb.e[i];
  g[10] = cb.da; g[11] = cb.de; g[12] = cb.dr; g[13] = cb.df;
  g[14] = inv_s * hb;
}

#if defined(__CUDACC__)
// BS: 256 (one block per SM at 255 registers) keeps the eight warps of an SM close together in the long
// straight-line code, whose fetch is the limiter otherwise (+17 % over two blocks of 128); 128 spreads shorter
// batches over more SMs and wastes less in the last wave.
template <bool SOA, int BS>
__global__ void __launch_bounds__(BS, AIRC_AB_H2_MINB)
hessian2_kernel(const __grid_constant__ DevConst P, const HessArgs a) {
  __shared__ __align__(16) float tab[NET_TAB_FLOATS];
  for (int i = threadIdx.x; i < NET_TAB_FLOATS; i += BS) net_tab_entry(P, i, tab);
  __syncthreads();
  const long long gt = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const bool live = gt / HESS2_LANES < a.n;
  const long long node = live ? gt / HESS2_LANES : a.n - 1;   // idle lanes of the last warp keep step (shuffles)
  const int lane = (int)(gt % HESS2_LANES);
  State s;
  load_state(a.x, node, s);
  const Ctl c = load_ctl(a.u.base + node * a.u.sb, a.u.sc);
  float lam[13];
#pragma unroll
  for (int r = 0; r < 13; ++r) lam[r] = __ldg(a.lam.base + node * a.lam.sb + r * a.lam.sc);
  const float dt = a.dt_per_node ? __ldg(a.dt_per_node + node) : a.dt;
  Dual g[NHACT];
  float value;
  hessian2_lane(P, tab, s, c, dt, a.substeps, a.normalise != 0, lam, lane, g, value);
  if (!live) return;
  const long long es = SOA ? a.n : 1;
  float* Hn = a.H + (SOA ? node : node * (NZ * NZ));
  // column j of the active block comes from one lane: entries (i <= j) are written with their mirror image,
  // so H is symmetric to the bit
#pragma unroll
  for (int half = 0; half < H2_DIRS; ++half) {
    const int j = H2_DIRS * lane + half;
    if (j < NHACT) {
      const int zj = hess_zindex(j);
#pragma unroll
      for (int i = 0; i < NHACT; ++i)
        if (i <= j) {
          const int zi = hess_zindex(i);
          const float val = TanOps<Tan>::get(g[i].t, half);
          Hn[(zi * NZ + zj) * es] = val;
          Hn[(zj * NZ + zi) * es] = val;
        }
    }
  }
  if (lane == 0) {
    if (a.grad) {
      float* gr = a.grad + (SOA ? node : node * NZ);
#pragma unroll
      for (int i = 0; i < NHACT; ++i) gr[hess_zindex(i) * es] = g[i].v;
#pragma unroll
      for (int r = 0; r < 3; ++r) { gr[r * es] = lam[r]; gr[(16 + r) * es] = 0.0f; }   // d/dp = lambda_p, thrust inert
    }
    if (a.value) a.value[node] = value;
  }
}
#endif

}  // namespace airc

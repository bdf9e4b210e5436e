// airc_tangent.cuh -- forward-mode directional derivative of f, written once for scalar tangents
// (T = float: one column) and packed tangents (T = F2: two columns per FFMA2 instruction).
// Primal quantities (Stage, J) are scalar and enter packed instructions as broadcast operands.
// Same mathematics as f_tangent in airc_dynamics.cuh (base.py:385-406 differentiated; sign() has
// zero derivative as in CasADi).
#pragma once
#include "airc_dynamics.cuh"

namespace airc {

template <class T>
struct CtlT {
  T da, de, dr, df;
};

template <class T>
AIRC_HD void rot_matrix_tangent_t(const float q[4], const T dq[4], T dM[9], T& dn2) {
  const float x = q[0], y = q[1], z = q[2], w = q[3];
  const T xdx = tmul(x, dq[0]), ydy = tmul(y, dq[1]), zdz = tmul(z, dq[2]), wdw = tmul(w, dq[3]);
  dn2 = tmul(2.0f, tadd(tadd(xdx, ydy), tadd(zdz, wdw)));
  dM[0] = tmul(2.0f, tsub(tadd(wdw, xdx), tadd(ydy, zdz)));
  dM[4] = tmul(2.0f, tsub(tadd(wdw, ydy), tadd(xdx, zdz)));
  dM[8] = tmul(2.0f, tsub(tadd(wdw, zdz), tadd(xdx, ydy)));
  const T dxy = tsum(x, dq[1], y, dq[0]), dxz = tsum(x, dq[2], z, dq[0]), dyz = tsum(y, dq[2], z, dq[1]);
  const T dwx = tsum(w, dq[0], x, dq[3]), dwy = tsum(w, dq[1], y, dq[3]), dwz = tsum(w, dq[2], z, dq[3]);
  dM[1] = tmul(2.0f, tsub(dxy, dwz));
  dM[2] = tmul(2.0f, tadd(dxz, dwy));
  dM[3] = tmul(2.0f, tadd(dxy, dwz));
  dM[5] = tmul(2.0f, tsub(dyz, dwx));
  dM[6] = tmul(2.0f, tsub(dxz, dwy));
  dM[7] = tmul(2.0f, tadd(dyz, dwx));
}

// (dv_dot, dq_dot, dw_dot) along (dv, dq, dw; dc) at the point described by (v, q, w, s, J)
template <class T>
AIRC_HD void f_tangent_t(const DevConst& P, const float v[3], const float q[4], const float w[3],
                         const Stage& s, const float J[6][5], const T dv[3], const T dq[4],
                         const T dw[3], const CtlT<T>& dc, T dvdot[3], T dqdot[4], T dwdot[3]) {
  T dM[9], dn2;
  rot_matrix_tangent_t<T>(q, dq, dM, dn2);
  const T rn = tmul(dn2, s.inv_n2);  // d(n2)/n2
  T dvb[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    T a = tmul(v[0], dM[i]);
    a = tfma(v[1], dM[3 + i], a);
    a = tfma(v[2], dM[6 + i], a);
    a = tfma(s.M[i], dv[0], a);
    a = tfma(s.M[3 + i], dv[1], a);
    a = tfma(s.M[6 + i], dv[2], a);
    dvb[i] = tfma(a, s.inv_n2, tmul(-s.vbr[i], rn));
  }
  const T dvb2 = tmul(2.0f, tfma(s.vb[2], dvb[2], tfma(s.vb[1], dvb[1], tmul(s.vb[0], dvb[0]))));
  const T dqbar = tmul(0.6125f, dvb2);
  const T dalpha = tmul(tdiff<T>(s.xa, dvb[2], s.vb[2], dvb[0]), s.inv_den_a);
  const T dV = tmul(0.5f * s.invV, dvb2);
  const T dsb = tmul(tfma(-s.sb, dV, dvb[1]), s.invV);
  const T dbeta = tmul(dsb, s.inv_cosb);

  T dC[6];
#pragma unroll
  for (int k = 0; k < 6; ++k) {
    T a = tmul(J[k][0], dqbar);
    a = tfma(J[k][1], dalpha, a);
    a = tfma(J[k][2], dbeta, a);
    a = tfma(J[k][3], dc.da, a);
    a = tfma(J[k][4], dc.de, a);
    dC[k] = a;
  }
  dC[5] = tfma(P.k_rudder, dc.dr, dC[5]);
  if (P.stall_scaling) {
    const float k = 10.0f;
    const T dsa = tmul(-k * s.sa * (1.0f - s.sa) * (s.alpha < 0.0f ? -1.0f : 1.0f), dalpha);
    const T dsb_s = tmul(-k * s.sbeta_s * (1.0f - s.sbeta_s) * (s.beta < 0.0f ? -1.0f : 1.0f), dbeta);
    dC[2] = tfma(s.Craw2, tsum<T>(s.sbeta_s, dsa, s.sa, dsb_s), tmul(s.sa * s.sbeta_s, dC[2]));
    dC[4] = tfma(s.Craw4, dsa, tmul(s.sa, dC[4]));
  }
  dC[0] = tfma(P.k_flap_x, dc.df, dC[0]);
  dC[2] = tfma(P.k_flap_z, dc.df, dC[2]);

  const T dqS = tmul(P.S, dqbar);
  T dF[3];
  dF[0] = tmul(s.sgn, tsum<T>(s.C[0], dqS, s.qS, dC[0]));
  dF[1] = tsum<T>(s.C[1], dqS, s.qS, dC[1]);
  dF[2] = tsum<T>(s.C[2], dqS, s.qS, dC[2]);
  T dMo[3];
  dMo[0] = tadd(tmul(P.Sb, tsum<T>(s.C[3], dqbar, s.qbar, dC[3])), tdiff<T>(P.com[1], dF[2], P.com[2], dF[1]));
  dMo[1] = tadd(tmul(P.Sc, tsum<T>(s.C[4], dqbar, s.qbar, dC[4])), tdiff<T>(P.com[2], dF[0], P.com[0], dF[2]));
  dMo[2] = tadd(tmul(P.Sb, tsum<T>(s.C[5], dqbar, s.qbar, dC[5])), tdiff<T>(P.com[0], dF[1], P.com[1], dF[0]));

  const float sc = s.inv_n2 * P.inv_mass;
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    T a = tmul(s.F[0], dM[3 * i]);
    a = tfma(s.F[1], dM[3 * i + 1], a);
    a = tfma(s.F[2], dM[3 * i + 2], a);
    a = tfma(s.M[3 * i], dF[0], a);
    a = tfma(s.M[3 * i + 1], dF[1], a);
    a = tfma(s.M[3 * i + 2], dF[2], a);
    dvdot[i] = tfma(a, sc, tmul(-s.acc[i], rn));
  }
  // d(0.5 q (x) (w,0)) = 0.5 (dq (x) (w,0) + q (x) (dw,0))
  dqdot[0] = tmul(0.5f, tadd(tfma(w[0], dq[3], tdiff<T>(w[2], dq[1], w[1], dq[2])),
                             tfma(q[3], dw[0], tdiff<T>(q[1], dw[2], q[2], dw[1]))));
  dqdot[1] = tmul(0.5f, tadd(tfma(w[1], dq[3], tdiff<T>(w[0], dq[2], w[2], dq[0])),
                             tfma(q[3], dw[1], tdiff<T>(q[2], dw[0], q[0], dw[2]))));
  dqdot[2] = tmul(0.5f, tadd(tfma(w[2], dq[3], tdiff<T>(w[1], dq[0], w[0], dq[1])),
                             tfma(q[3], dw[2], tdiff<T>(q[0], dw[1], q[1], dw[0]))));
  dqdot[3] = tmul(-0.5f, tadd(tfma(w[2], dq[2], tsum<T>(w[0], dq[0], w[1], dq[1])),
                              tfma(q[2], dw[2], tsum<T>(q[0], dw[0], q[1], dw[1]))));
  T dIw[3];
#pragma unroll
  for (int i = 0; i < 3; ++i)
    dIw[i] = tfma(P.I[3 * i + 2], dw[2], tsum<T>(P.I[3 * i], dw[0], P.I[3 * i + 1], dw[1]));
  // d(w x Iw) = dw x Iw + w x dIw
  T t[3];
  t[0] = tsub(tsub(dMo[0], tdiff<T>(s.Iw[2], dw[1], s.Iw[1], dw[2])), tdiff<T>(w[1], dIw[2], w[2], dIw[1]));
  t[1] = tsub(tsub(dMo[1], tdiff<T>(s.Iw[0], dw[2], s.Iw[2], dw[0])), tdiff<T>(w[2], dIw[0], w[0], dIw[2]));
  t[2] = tsub(tsub(dMo[2], tdiff<T>(s.Iw[1], dw[0], s.Iw[0], dw[1])), tdiff<T>(w[0], dIw[1], w[1], dIw[0]));
#pragma unroll
  for (int i = 0; i < 3; ++i)
    dwdot[i] = tfma(P.Iinv[3 * i + 2], t[2], tsum<T>(P.Iinv[3 * i], t[0], P.Iinv[3 * i + 1], t[1]));
}

}  // namespace airc

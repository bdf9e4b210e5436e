// airc_dynamics.cuh -- FP32 6-DOF dynamics with the fused surrogate, primal and tangent.
//
// Mathematics restated from (never copied; the reference is a CasADi graph in Python):
//   /root/reference/src/aircraft/dynamics/base.py:147-177,229-241   v_frd_rel, alpha, beta, qbar
//   /root/reference/src/aircraft/dynamics/aircraft.py:255-330        coefficients, forces, moments
//   /root/reference/src/aircraft/dynamics/base.py:280-297,344-406    forces_ned, q_dot, v_dot, omega_dot, f
//   /root/reference/src/aircraft/surrogates/models.py:143-155        ScaledModel.forward
//   /root/reference/src/aircraft/dynamics/coefficient_models.py:99-104  NeuralModel (+ rudder term)
//
// The quaternion sandwiches q^-1 (x) (v,0) (x) q and q (x) (F,0) (x) q^-1 (liecasadi, xyzw,
// inverse = conj/|q|^2) are evaluated through the unnormalised rotation matrix
//   M(q) = (w^2-|qv|^2) I + 2 qv qv^T + 2 w [qv]x ,   v_b = M^T v / |q|^2 ,  F_n = M F / |q|^2
// which is algebraically identical and shares M between both uses.
#pragma once
#include "airc_math.cuh"

namespace airc {

constexpr int NH = 32;  // hidden units after folding Linear(5,16).Linear(16,32)

// Fused surrogate.  With z_in = (in - mu_in)/sigma_in and W21 = W2.W1, b21 = W2.b1 + b2:
//   h_j  = kappa*(W21 z_in + b21)_j,  kappa = 2 log2(e)          (so tanh(a_j) = 1 - 2/(1+2^h_j))
//        = a0_j + Ax[j].(qbar,alpha,beta) + Au[j].(aileron,elevator)
//   C_k  = c0_k + sum_j Wo[k][j] r_j,   r_j = 1/(1+2^h_j)
//   Wo   = -2 sigma_out W3,  c0 = sigma_out (b3 + W3.1) + mu_out
//   dC_k/din_i = sum_j Wo[k][j] g_j A[j][i],  g_j = -ln2 r_j (1-r_j)
// Hidden units are stored in PAIRS (2p, 2p+1) with the two weights of every input adjacent, so
// that one LDCU.128 delivers two packed FFMA2 multiplicands.
struct alignas(16) HidPair {
  float a0[2];
  float Au[2][2];   // [aileron|elevator][unit in pair]
  float Ax[3][2];   // [qbar|alpha|beta][unit in pair]
};
struct alignas(16) FusedNet {
  HidPair hid[NH / 2];
  float Wo[6][NH];  // (Wo[k][2p], Wo[k][2p+1]) is the packed multiplicand of the output layer
  float c0[6];
  float pad_[2];
};

// Everything a kernel needs, passed BY VALUE as a __grid_constant__ kernel parameter so that
// weights are constant-bank operands of FFMA (no load instructions, no global state, one
// block per handle).
struct DevConst {
  FusedNet net;
  float S;            // reference area
  float Sb, Sc;       // S*span, S*chord
  float inv_mass;
  float com[3];
  float I[9];
  float Iinv[9];
  float g[3];
  float eps;
  float k_rudder;     // -0.1*6*pi/180   (coefficient_models.py:102-103)
  float k_flap_x;     // -0.1            (aircraft.py:297-299)
  float k_flap_z;     // -0.6
  int stall_scaling;
  int model;           // 0 nn | AIRC_MODEL_* (airc_alt.cuh): then `net` holds AltParams instead of the network
};

// control inputs that the NN path actually reads (thrust is inert, aircraft.py:320)
struct Ctl {
  float da, de, dr, df;
};

// primal intermediates of one evaluation of f, kept only when tangents follow
struct Stage {
  float M[9];
  float inv_n2;
  float vbr[3];      // M^T v / n2   (v_b without eps)
  float vb[3];       // v_b = vbr + eps
  float xa;          // vb.x + eps
  float inv_den_a;   // 1/(xa^2 + vb.z^2)
  float invV;        // 1/V
  float sb;          // vb.y / V
  float inv_cosb;    // 1/sqrt(1 - sb^2)
  float qbar, alpha, beta;
  float C[6];        // after rudder/flaps(/stall)
  float qS;
  float sgn;
  float F[3];
  float acc[3];      // vdot - g = M F inv_n2 / m
  float Iw[3];
  float sa, sbeta_s; // stall scalings (1 when disabled)
  float Craw2, Craw4;// NN outputs before stall scaling (for the tangent of the scaling)
};

AIRC_HD void rot_matrix(const float q[4], float M[9], float& n2) {
  const float x = q[0], y = q[1], z = q[2], w = q[3];
  const float xx = x * x, yy = y * y, zz = z * z, ww = w * w;
  n2 = (xx + yy) + (zz + ww);
  M[0] = (ww + xx) - (yy + zz);
  M[4] = (ww + yy) - (xx + zz);
  M[8] = (ww + zz) - (xx + yy);
  const float xy = x * y, xz = x * z, yz = y * z, wx = w * x, wy = w * y, wz = w * z;
  M[1] = 2.0f * (xy - wz);
  M[2] = 2.0f * (xz + wy);
  M[3] = 2.0f * (xy + wz);
  M[5] = 2.0f * (yz - wx);
  M[6] = 2.0f * (xz - wy);
  M[7] = 2.0f * (yz + wx);
}

// directional derivative of M along dq
AIRC_HD void rot_matrix_tangent(const float q[4], const float dq[4], float dM[9], float& dn2) {
  const float x = q[0], y = q[1], z = q[2], w = q[3];
  const float dx = dq[0], dy = dq[1], dz = dq[2], dw = dq[3];
  const float xdx = x * dx, ydy = y * dy, zdz = z * dz, wdw = w * dw;
  dn2 = 2.0f * ((xdx + ydy) + (zdz + wdw));
  dM[0] = 2.0f * ((wdw + xdx) - (ydy + zdz));
  dM[4] = 2.0f * ((wdw + ydy) - (xdx + zdz));
  dM[8] = 2.0f * ((wdw + zdz) - (xdx + ydy));
  const float dxy = x * dy + y * dx, dxz = x * dz + z * dx, dyz = y * dz + z * dy;
  const float dwx = w * dx + x * dw, dwy = w * dy + y * dw, dwz = w * dz + z * dw;
  dM[1] = 2.0f * (dxy - dwz);
  dM[2] = 2.0f * (dxz + dwy);
  dM[3] = 2.0f * (dxy + dwz);
  dM[5] = 2.0f * (dyz - dwx);
  dM[6] = 2.0f * (dxz - dwy);
  dM[7] = 2.0f * (dyz + dwx);
}

// hidden layer + output layer, two hidden units per packed instruction.
template <bool WITH_JAC>
AIRC_HD void surrogate(const FusedNet& net, float qbar, float alpha, float beta, const Ctl& c,
                       float C[6], float J[6][5]) {
  F2 acc[6];
  F2 Jacc[WITH_JAC ? 6 : 1][WITH_JAC ? 5 : 1];
#pragma unroll
  for (int k = 0; k < 6; ++k) {
    acc[k] = F2{net.c0[k], 0.0f};
    if (WITH_JAC) {
#pragma unroll
      for (int i = 0; i < 5; ++i) Jacc[k][i] = F2{0.0f, 0.0f};
    }
  }
  const F2 da = bc2(c.da), de = bc2(c.de), qb = bc2(qbar), al = bc2(alpha), be = bc2(beta);
#pragma unroll
  for (int p = 0; p < NH / 2; ++p) {
    const HidPair& w = net.hid[p];
    F2 h = F2{w.a0[0], w.a0[1]};
    h = fma2b(F2{w.Au[0][0], w.Au[0][1]}, da, h);
    h = fma2b(F2{w.Au[1][0], w.Au[1][1]}, de, h);
    h = fma2b(F2{w.Ax[0][0], w.Ax[0][1]}, qb, h);
    h = fma2b(F2{w.Ax[1][0], w.Ax[1][1]}, al, h);
    h = fma2b(F2{w.Ax[2][0], w.Ax[2][1]}, be, h);
    const F2 r = sigm2_x2(h);   // AIRC_SIGM_MODE selects the variant (airc_math.cuh)
#pragma unroll
    for (int k = 0; k < 6; ++k) acc[k] = fma2b(F2{net.Wo[k][2 * p], net.Wo[k][2 * p + 1]}, r, acc[k]);
    if (WITH_JAC) {
      // g_j = d r_j / d h_j = -ln2 r (1 - r)
      const F2 g = mul2b(mul2b(r, bc2(-0.69314718056f)), F2{1.0f - r.lo, 1.0f - r.hi});
#pragma unroll
      for (int i = 0; i < 5; ++i) {
        const F2 wi = (i < 3) ? F2{w.Ax[i][0], w.Ax[i][1]} : F2{w.Au[i - 3][0], w.Au[i - 3][1]};
        const F2 gw = mul2b(g, wi);
#pragma unroll
        for (int k = 0; k < 6; ++k)
          Jacc[k][i] = fma2b(F2{net.Wo[k][2 * p], net.Wo[k][2 * p + 1]}, gw, Jacc[k][i]);
      }
    }
  }
#pragma unroll
  for (int k = 0; k < 6; ++k) {
    C[k] = acc[k].lo + acc[k].hi;
    if (WITH_JAC) {
#pragma unroll
      for (int i = 0; i < 5; ++i) J[k][i] = Jacc[k][i].lo + Jacc[k][i].hi;
    }
  }
}

// How the network is evaluated: by this thread alone (NetFull) or shared among the lanes that
// cooperate on one node (NetCoop in airc_coop.cuh).
struct NetFull {
  static constexpr bool kPolyTrig = true;   // alpha, beta by the packed polynomial atan2 (airc_math.cuh)
  const FusedNet* net;
  template <bool WITH_JAC>
  AIRC_HD void eval(float qbar, float alpha, float beta, const Ctl& c, float C[6], float J[6][5]) const {
    surrogate<WITH_JAC>(*net, qbar, alpha, beta, c, C, J);
  }
  // hook for coefficient models that also read the body rates (airc_alt.cuh); the NN model does not
  AIRC_HD void adjust(const DevConst&, Stage&, const float*, const Ctl&) const {}
};

// f(x,u) without the trivial p_dot = v rows.  base.py:385-406.
template <bool WITH_JAC, class Net>
AIRC_HD void f_eval_net(const DevConst& P, const Net& net, const float v[3], const float q[4],
                        const float w[3], const Ctl& c, float vdot[3], float qdot[4], float wdot[3],
                        Stage& s, float J[6][5]) {
  float n2;
  rot_matrix(q, s.M, n2);
  s.inv_n2 = rcp_nr(n2);
  // v_b = M^T v / n2 + eps                                   base.py:156-159
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    float a = s.M[i] * v[0];
    a = fmaf(s.M[3 + i], v[1], a);
    a = fmaf(s.M[6 + i], v[2], a);
    s.vbr[i] = a * s.inv_n2;
    s.vb[i] = s.vbr[i] + P.eps;
  }
  const float xz2 = fmaf(s.vb[2], s.vb[2], s.vb[0] * s.vb[0]);
  const float vb2 = fmaf(s.vb[1], s.vb[1], xz2);
  s.qbar = 0.6125f * vb2;                                   // 0.5*1.225*|v_b|^2   base.py:240
  s.xa = s.vb[0] + P.eps;
  if constexpr (Net::kPolyTrig) {
    // NN path: alpha and beta from one packed evaluation of the branch-free polynomial atan2;
    // asin(vb_y / V) == atan2(vb_y, sqrt(vb_x^2 + vb_z^2 + eps))            base.py:167,176,234
    float rxz, inv_rxz;
    sqrt_rsqrt(xz2 + P.eps, rxz, inv_rxz);                  // V cos(beta)
    alpha_beta(s.vb[2], s.xa, s.vb[1], rxz, s.alpha, s.beta);
    if (WITH_JAC) {
      float V;
      sqrt_rsqrt(vb2 + P.eps, V, s.invV);                   // base.py:167
      s.sb = s.vb[1] * s.invV;
      s.inv_den_a = rcp_nr(fmaf(s.xa, s.xa, s.vb[2] * s.vb[2]));
      s.inv_cosb = inv_rxz * V;                             // 1/cos(beta) = V / sqrt(vb_x^2 + vb_z^2 + eps)
    }
  } else {
    float V;
    sqrt_rsqrt(vb2 + P.eps, V, s.invV);                     // base.py:167
    s.alpha = atan2_acc(s.vb[2], s.xa);                     // base.py:176
    s.sb = s.vb[1] * s.invV;
    s.beta = asin_acc(s.sb);                                // base.py:234
    if (WITH_JAC) {
      s.inv_den_a = rcp_nr(fmaf(s.xa, s.xa, s.vb[2] * s.vb[2]));
      s.inv_cosb = fast_rsqrt(fmaf(-s.sb, s.sb, 1.0f));
    }
  }

  net.template eval<WITH_JAC>(s.qbar, s.alpha, s.beta, c, s.C, J);
  net.adjust(P, s, w, c);
  s.C[5] = fmaf(P.k_rudder, c.dr, s.C[5]);                  // coefficient_models.py:102-103
  s.sa = 1.0f;
  s.sbeta_s = 1.0f;
  s.Craw2 = s.C[2];
  s.Craw4 = s.C[4];
  if (P.stall_scaling) {                                    // aircraft.py:280-294
    const float k = 10.0f, a30 = 0.52359877559f;
    s.sa = 1.0f / (1.0f + expf(k * (fabsf(s.alpha) - a30)));
    s.sbeta_s = 1.0f / (1.0f + expf(k * (fabsf(s.beta) - a30)));
    s.C[2] = s.C[2] * s.sa * s.sbeta_s;
    s.C[4] = s.C[4] * s.sa;
  }
  s.C[0] = fmaf(P.k_flap_x, c.df, s.C[0]);                  // aircraft.py:297-299
  s.C[2] = fmaf(P.k_flap_z, c.df, s.C[2]);

  // forces and moments, body frame                          aircraft.py:313-328 ; base.py:255,275
  s.qS = s.qbar * P.S;
  s.sgn = sign_casadi(s.vb[0]);
  s.F[0] = s.C[0] * s.qS * s.sgn;
  s.F[1] = s.C[1] * s.qS;
  s.F[2] = s.C[2] * s.qS;
  float Mo[3];
  Mo[0] = s.C[3] * (s.qbar * P.Sb) + (P.com[1] * s.F[2] - P.com[2] * s.F[1]);
  Mo[1] = s.C[4] * (s.qbar * P.Sc) + (P.com[2] * s.F[0] - P.com[0] * s.F[2]);
  Mo[2] = s.C[5] * (s.qbar * P.Sb) + (P.com[0] * s.F[1] - P.com[1] * s.F[0]);

  // v_dot = M F /(n2 m) + g                                  base.py:284-286,359
  const float sc = s.inv_n2 * P.inv_mass;
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    float a = s.M[3 * i] * s.F[0];
    a = fmaf(s.M[3 * i + 1], s.F[1], a);
    a = fmaf(s.M[3 * i + 2], s.F[2], a);
    s.acc[i] = a * sc;
    vdot[i] = s.acc[i] + P.g[i];
  }
  // q_dot = 0.5 q (x) (w,0)                                  base.py:292-295
  qdot[0] = 0.5f * (q[3] * w[0] + (q[1] * w[2] - q[2] * w[1]));
  qdot[1] = 0.5f * (q[3] * w[1] + (q[2] * w[0] - q[0] * w[2]));
  qdot[2] = 0.5f * (q[3] * w[2] + (q[0] * w[1] - q[1] * w[0]));
  qdot[3] = -0.5f * (q[0] * w[0] + q[1] * w[1] + q[2] * w[2]);
  // w_dot = I^-1 (M - w x (I w))                             base.py:376-377
#pragma unroll
  for (int i = 0; i < 3; ++i)
    s.Iw[i] = fmaf(P.I[3 * i + 2], w[2], fmaf(P.I[3 * i + 1], w[1], P.I[3 * i] * w[0]));
  float t[3];
  t[0] = Mo[0] - (w[1] * s.Iw[2] - w[2] * s.Iw[1]);
  t[1] = Mo[1] - (w[2] * s.Iw[0] - w[0] * s.Iw[2]);
  t[2] = Mo[2] - (w[0] * s.Iw[1] - w[1] * s.Iw[0]);
#pragma unroll
  for (int i = 0; i < 3; ++i)
    wdot[i] = fmaf(P.Iinv[3 * i + 2], t[2], fmaf(P.Iinv[3 * i + 1], t[1], P.Iinv[3 * i] * t[0]));
}

}  // namespace airc
#include "airc_alt.cuh"
namespace airc {

// MODEL 0: the NN surrogate (every tuned kernel); MODEL 1: the registry's other coefficient models
// (airc_alt.cuh), primal only.
template <bool WITH_JAC, int MODEL = 0>
AIRC_HD void f_eval(const DevConst& P, const float v[3], const float q[4], const float w[3],
                    const Ctl& c, float vdot[3], float qdot[4], float wdot[3], Stage& s,
                    float J[6][5]) {
  if constexpr (MODEL == 0) f_eval_net<WITH_JAC>(P, NetFull{&P.net}, v, q, w, c, vdot, qdot, wdot, s, J);
  else f_eval_net<false>(P, NetAlt{&alt_params(P)}, v, q, w, c, vdot, qdot, wdot, s, J);
}

// Directional derivative of f at the point described by (v,q,w,c,s,J) along
// (dv,dq,dw ; dda,dde,ddr,ddf).  Forward mode; sign() has zero derivative as in CasADi.
AIRC_HD void f_tangent(const DevConst& P, const float v[3], const float q[4], const float w[3],
                       const Stage& s, const float J[6][5], const float dv[3], const float dq[4],
                       const float dw[3], const Ctl& dc, float dvdot[3], float dqdot[4],
                       float dwdot[3]) {
  float dM[9], dn2;
  rot_matrix_tangent(q, dq, dM, dn2);
  const float rn = dn2 * s.inv_n2;  // d(n2)/n2
  float dvb[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    float a = dM[i] * v[0];
    a = fmaf(dM[3 + i], v[1], a);
    a = fmaf(dM[6 + i], v[2], a);
    a = fmaf(s.M[i], dv[0], a);
    a = fmaf(s.M[3 + i], dv[1], a);
    a = fmaf(s.M[6 + i], dv[2], a);
    dvb[i] = fmaf(a, s.inv_n2, -s.vbr[i] * rn);
  }
  const float dvb2 = 2.0f * fmaf(s.vb[2], dvb[2], fmaf(s.vb[1], dvb[1], s.vb[0] * dvb[0]));
  const float dqbar = 0.6125f * dvb2;
  const float dalpha = (s.xa * dvb[2] - s.vb[2] * dvb[0]) * s.inv_den_a;
  const float dV = 0.5f * dvb2 * s.invV;
  const float dsb = (dvb[1] - s.sb * dV) * s.invV;
  const float dbeta = dsb * s.inv_cosb;

  float dC[6];
#pragma unroll
  for (int k = 0; k < 6; ++k) {
    float a = J[k][0] * dqbar;
    a = fmaf(J[k][1], dalpha, a);
    a = fmaf(J[k][2], dbeta, a);
    a = fmaf(J[k][3], dc.da, a);
    a = fmaf(J[k][4], dc.de, a);
    dC[k] = a;
  }
  dC[5] = fmaf(P.k_rudder, dc.dr, dC[5]);
  if (P.stall_scaling) {
    const float k = 10.0f;
    const float dsa = -k * s.sa * (1.0f - s.sa) * (s.alpha < 0.0f ? -dalpha : dalpha);
    const float dsb_s = -k * s.sbeta_s * (1.0f - s.sbeta_s) * (s.beta < 0.0f ? -dbeta : dbeta);
    dC[2] = dC[2] * s.sa * s.sbeta_s + s.Craw2 * (dsa * s.sbeta_s + s.sa * dsb_s);
    dC[4] = dC[4] * s.sa + s.Craw4 * dsa;
  }
  dC[0] = fmaf(P.k_flap_x, dc.df, dC[0]);
  dC[2] = fmaf(P.k_flap_z, dc.df, dC[2]);

  const float dqS = dqbar * P.S;
  float dF[3];
  dF[0] = (s.C[0] * dqS + s.qS * dC[0]) * s.sgn;
  dF[1] = s.C[1] * dqS + s.qS * dC[1];
  dF[2] = s.C[2] * dqS + s.qS * dC[2];
  float dMo[3];
  dMo[0] = (s.C[3] * dqbar + s.qbar * dC[3]) * P.Sb + (P.com[1] * dF[2] - P.com[2] * dF[1]);
  dMo[1] = (s.C[4] * dqbar + s.qbar * dC[4]) * P.Sc + (P.com[2] * dF[0] - P.com[0] * dF[2]);
  dMo[2] = (s.C[5] * dqbar + s.qbar * dC[5]) * P.Sb + (P.com[0] * dF[1] - P.com[1] * dF[0]);

  const float sc = s.inv_n2 * P.inv_mass;
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    float a = dM[3 * i] * s.F[0];
    a = fmaf(dM[3 * i + 1], s.F[1], a);
    a = fmaf(dM[3 * i + 2], s.F[2], a);
    a = fmaf(s.M[3 * i], dF[0], a);
    a = fmaf(s.M[3 * i + 1], dF[1], a);
    a = fmaf(s.M[3 * i + 2], dF[2], a);
    dvdot[i] = fmaf(a, sc, -s.acc[i] * rn);
  }
  dqdot[0] = 0.5f * ((dq[3] * w[0] + (dq[1] * w[2] - dq[2] * w[1])) +
                     (q[3] * dw[0] + (q[1] * dw[2] - q[2] * dw[1])));
  dqdot[1] = 0.5f * ((dq[3] * w[1] + (dq[2] * w[0] - dq[0] * w[2])) +
                     (q[3] * dw[1] + (q[2] * dw[0] - q[0] * dw[2])));
  dqdot[2] = 0.5f * ((dq[3] * w[2] + (dq[0] * w[1] - dq[1] * w[0])) +
                     (q[3] * dw[2] + (q[0] * dw[1] - q[1] * dw[0])));
  dqdot[3] = -0.5f * ((dq[0] * w[0] + dq[1] * w[1] + dq[2] * w[2]) +
                      (q[0] * dw[0] + q[1] * dw[1] + q[2] * dw[2]));
  float dIw[3];
#pragma unroll
  for (int i = 0; i < 3; ++i)
    dIw[i] = fmaf(P.I[3 * i + 2], dw[2], fmaf(P.I[3 * i + 1], dw[1], P.I[3 * i] * dw[0]));
  float t[3];
  t[0] = dMo[0] - (dw[1] * s.Iw[2] - dw[2] * s.Iw[1]) - (w[1] * dIw[2] - w[2] * dIw[1]);
  t[1] = dMo[1] - (dw[2] * s.Iw[0] - dw[0] * s.Iw[2]) - (w[2] * dIw[0] - w[0] * dIw[2]);
  t[2] = dMo[2] - (dw[0] * s.Iw[1] - dw[1] * s.Iw[0]) - (w[0] * dIw[1] - w[1] * dIw[0]);
#pragma unroll
  for (int i = 0; i < 3; ++i)
    dwdot[i] = fmaf(P.Iinv[3 * i + 2], t[2], fmaf(P.Iinv[3 * i + 1], t[1], P.Iinv[3 * i] * t[0]));
}

}  // namespace airc

// airc_jac.cuh -- K2, the Jacobian kernel of the NN path: (x+, A = dPhi/dx, B = dPhi/du [, dPhi/d(dt)])
// of one RK4 state_update per shooting node, in the STRUCTURED form
//
//   K_1 = [f_x | f_u](X_1),  S_{i+1} = S_1 + c_i h K_i,  K_{i+1} = f_x(X_{i+1}) S_{i+1} + f_u(X_{i+1}),
//   [A | B] = S_1 + h/6 (K_1 + 2 K_2 + 2 K_3 + K_4)                     (base.py:408-446 differentiated)
//
// What it replaces: ca.jacobian(state_update, x|u) evaluated over every node of a trajectory
// (/root/reference/src/aircraft/control/aircraft.py:85-95, base.py:51-52) and the constraint Jacobian
// blocks IPOPT takes from the defect constraints (control/base.py:275-286).
//
// Two phases per tile of NT nodes, one thread block per tile:
//   phase 1  ONE NODE PER THREAD PAIR, no replication and no shuffles: the primal RK4 step with the surrogate and
//            its 6x5 Jacobian, and for each of the four stage points the non-zero blocks of
//            f_x (10x10 over v, q, omega: f does not read p) and f_u (aileron, elevator, rudder, flaps;
//            thrust is inert, aircraft.py:320), assembled ONCE per node from the chain
//            (v,q) -> v_b -> (qbar, alpha, beta) -> C -> (F, M) -> (v_dot, omega_dot)   and parked in
//            shared memory: 84 floats per stage, 348 per node (92 / 380 with the d/d(dt) column).
//            The two threads of a pair (in two warps on the same scheduler) alternate the stages: f_jac_head
//            gives the primal k after a quarter of a stage's work, the next stage point goes to the partner,
//            and f_jac_tail (the blocks) overlaps the partner's head (airc_jac_kernel.cuh).
//   phase 2  EIGHT LANES PER NODE, one packed column pair (FFMA2) per lane: the tangent recursion above as
//            small matrix-vector products with the parked blocks as broadcast operands (~80 FFMA2 per stage
//            instead of ~230 for pushing a tangent through the whole of f); results go to a shared-memory
//            image of the output and leave with 128-bit coalesced stores in either layout.
// Column slots as in the kernel this replaces: 0..9 d/d(v,q,omega), 10..13 d/d(aileron, elevator, rudder,
// flaps), 14 d/d(dt) (time='progress', control/base.py:276), 15 idle.
// One substep per call (the setting of every MPC / MHTT script, main/control/control.py:166,
// main/mhe/mhtt.py:49); airc_api.cu routes substeps > 1 to the cooperative kernel of airc_coop.cuh.
//
// Everything except the kernel itself is AIRC_HD: oracle/host_emul.cpp compiles the same formulas with g++
// and the CPU suite checks them against jacfwd of the float64 oracle.
#pragma once
#include "airc_packed.cuh"
#include "airc_step.cuh"

namespace airc {

// ---- per-stage record in shared memory (floats); 4-wide blocks first so that rows are 16-byte aligned
constexpr int JR_PQ = 0;     // d(v_dot)_i / d q_j          [3][4]
constexpr int JR_RQ = 12;    // d(w_dot)_i / d q_j          [3][4]
constexpr int JR_FU = 24;    // rows v_dot(3), w_dot(3) x cols (aileron, elevator, rudder, flaps)   [6][4]
constexpr int JR_Q = 48;     // stage q[4]
constexpr int JR_PV = 52;    // d(v_dot)_i / d v_j          [3][3]
constexpr int JR_RV = 61;    // d(w_dot)_i / d v_j          [3][3]
constexpr int JR_RW = 70;    // d(w_dot)_i / d w_j          [3][3]
constexpr int JR_W = 79;     // stage omega[3]
constexpr int JR_K = 82;     // f at the stage point: v_dot(3), q_dot(4), w_dot(3) -- d/d(dt) layout only (else 82, 83: pad)
// A stage block is 84 floats, 92 when the d/d(dt) column is wanted.  Node record: 4 stage blocks + header
// [h, ap(3) | aq(4) | pad(4)]; (NODE/4) is ODD so that 128-bit accesses with a stride of one record
// (phase 1: lane = node) are bank-conflict free.
template <bool DT>
struct JRec {
  static constexpr int STAGE = DT ? 92 : 84;
  static constexpr int HDR = 4 * STAGE;
  static constexpr int NODE = HDR + 12;
  static_assert(STAGE % 4 == 0 && NODE % 4 == 0 && ((NODE / 4) & 1) == 1, "record layout");
};

struct alignas(16) Q4 {
  float a, b, c, d;
};

// f(x,u) at one stage point (base.py:385-406) together with the non-zero blocks of its Jacobian, written to
// r[0..81].  Mathematics as f_eval_net / f_tangent_t (airc_dynamics.cuh, airc_tangent.cuh), factored:
//   D  = d v_b / d(v,q)            3x7   (closed form of d(M^T v / |q|^2))
//   G  = d(qbar, alpha, beta)/d v_b 3x3 ;  Y = G D                                   3x7
//   Z  = d(F, m)/d(qbar, alpha, beta) through the surrogate Jacobian J               6x3
//   W  = [ M/(m |q|^2) ; I^-1 [com]x , I^-1 ] Z                                      6x3
//   [P; R](v,q) = W Y  + direct terms (d(M F)/dq at fixed F; gyroscopic d(w x Iw)/dw)
//
// f_jac is cut in two so that the PRIMAL derivative k = f(X) is known after the first quarter of the work
// (f_jac_head: geometry, the surrogate's forward pass, forces, k) and the next RK4 stage point can be handed to
// another warp while this one assembles the Jacobian blocks (f_jac_tail: D, Y, the surrogate's derivative pass from
// the parked logistic values, Z, W, the blocks).  Same operations in the same order as the fused form.
struct JacMid {
  float M[9], inv, vbr[3], vb[3], xa, invV, inv_rxz, ida, sb, qbar, alpha, beta;
  float C[6], Craw2, Craw4, qS, sgn, F[3], lq[3], sc, acc[3], Iw[3];
  F2 rr[NH / 2];                 // logistic values r_j of the 32 hidden units, two per entry
};

// the surrogate's forward pass (as surrogate<>, airc_dynamics.cuh), keeping r for the derivative pass
AIRC_HD void surrogate_fwd(const FusedNet& net, float qbar, float alpha, float beta, const Ctl& c, float C[6],
                           F2 rr[NH / 2]) {
  F2 acc[6];
#pragma unroll
  for (int k = 0; k < 6; ++k) acc[k] = F2{net.c0[k], 0.0f};
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
    const F2 r = sigm2_x2(h);
    rr[p] = r;
#pragma unroll
    for (int k = 0; k < 6; ++k) acc[k] = fma2b(F2{net.Wo[k][2 * p], net.Wo[k][2 * p + 1]}, r, acc[k]);
  }
#pragma unroll
  for (int k = 0; k < 6; ++k) C[k] = acc[k].lo + acc[k].hi;
}

// J = Wo diag(g) W1,  g_j = d r_j / d h_j = -ln2 r (1 - r)
AIRC_HD void surrogate_jac(const FusedNet& net, const F2 rr[NH / 2], float J[6][5]) {
  F2 Jacc[6][5];
#pragma unroll
  for (int k = 0; k < 6; ++k)
#pragma unroll
    for (int i = 0; i < 5; ++i) Jacc[k][i] = F2{0.0f, 0.0f};
#pragma unroll
  for (int p = 0; p < NH / 2; ++p) {
    const HidPair& w = net.hid[p];
    const F2 r = rr[p];
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
#pragma unroll
  for (int k = 0; k < 6; ++k)
#pragma unroll
    for (int i = 0; i < 5; ++i) J[k][i] = Jacc[k][i].lo + Jacc[k][i].hi;
}

AIRC_HD void f_jac_head(const DevConst& P, const float v[3], const float q[4], const float w[3], const Ctl& c,
                        float kv[3], float kq[4], float kw[3], JacMid& m) {
  // ---- v_b = M^T v / |q|^2 + eps                                      base.py:156-159
  float M[9], n2;
  rot_matrix(q, M, n2);
  const float inv = rcp_nr(n2);
  float vbr[3], vb[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    float a = M[i] * v[0];
    a = fmaf(M[3 + i], v[1], a);
    a = fmaf(M[6 + i], v[2], a);
    vbr[i] = a * inv;
    vb[i] = vbr[i] + P.eps;
  }
  const float xz2 = fmaf(vb[2], vb[2], vb[0] * vb[0]);
  const float vb2 = fmaf(vb[1], vb[1], xz2);
  const float qbar = 0.6125f * vb2;                                   // base.py:240
  const float xa = vb[0] + P.eps;
  float V, invV, rxz, inv_rxz;
  sqrt_rsqrt(vb2 + P.eps, V, invV);                                   // base.py:167
  sqrt_rsqrt(xz2 + P.eps, rxz, inv_rxz);                              // V cos(beta)
  float alpha, beta;
  alpha_beta(vb[2], xa, vb[1], rxz, alpha, beta);                     // base.py:176,234
  const float ida = rcp_nr(fmaf(xa, xa, vb[2] * vb[2]));
  const float sb = vb[1] * invV;


  // ---- surrogate, forward pass                                       models.py:143-155, coefficient_models.py:99-104
  float C[6];
  surrogate_fwd(P.net, qbar, alpha, beta, c, C, m.rr);
  C[5] = fmaf(P.k_rudder, c.dr, C[5]);
  m.Craw2 = C[2];
  m.Craw4 = C[4];
  if (P.stall_scaling) {                                              // aircraft.py:280-294
    const float k = 10.0f, a30 = 0.52359877559f;
    const float sa = 1.0f / (1.0f + expf(k * (fabsf(alpha) - a30)));
    const float sbs = 1.0f / (1.0f + expf(k * (fabsf(beta) - a30)));
    C[2] *= sa * sbs;
    C[4] *= sa;
  }
  C[0] = fmaf(P.k_flap_x, c.df, C[0]);                                // aircraft.py:297-299
  C[2] = fmaf(P.k_flap_z, c.df, C[2]);

  // ---- forces, moments, f                                            aircraft.py:313-328 ; base.py:255-297,344-383
  const float qS = qbar * P.S;
  const float sgn = sign_casadi(vb[0]);
  const float F[3] = {C[0] * qS * sgn, C[1] * qS, C[2] * qS};
  const float lq[3] = {qbar * P.Sb, qbar * P.Sc, qbar * P.Sb};
  float Mo[3];
  Mo[0] = C[3] * lq[0] + (P.com[1] * F[2] - P.com[2] * F[1]);
  Mo[1] = C[4] * lq[1] + (P.com[2] * F[0] - P.com[0] * F[2]);
  Mo[2] = C[5] * lq[2] + (P.com[0] * F[1] - P.com[1] * F[0]);
  const float sc = inv * P.inv_mass;
  float acc[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    float a = M[3 * i] * F[0];
    a = fmaf(M[3 * i + 1], F[1], a);
    a = fmaf(M[3 * i + 2], F[2], a);
    acc[i] = a * sc;
    kv[i] = acc[i] + P.g[i];
  }
  kq[0] = 0.5f * (q[3] * w[0] + (q[1] * w[2] - q[2] * w[1]));
  kq[1] = 0.5f * (q[3] * w[1] + (q[2] * w[0] - q[0] * w[2]));
  kq[2] = 0.5f * (q[3] * w[2] + (q[0] * w[1] - q[1] * w[0]));
  kq[3] = -0.5f * (q[0] * w[0] + q[1] * w[1] + q[2] * w[2]);
  float Iw[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) Iw[i] = fmaf(P.I[3 * i + 2], w[2], fmaf(P.I[3 * i + 1], w[1], P.I[3 * i] * w[0]));
  {
    const float t[3] = {Mo[0] - (w[1] * Iw[2] - w[2] * Iw[1]), Mo[1] - (w[2] * Iw[0] - w[0] * Iw[2]),
                        Mo[2] - (w[0] * Iw[1] - w[1] * Iw[0])};
#pragma unroll
    for (int i = 0; i < 3; ++i) kw[i] = fmaf(P.Iinv[3 * i + 2], t[2], fmaf(P.Iinv[3 * i + 1], t[1], P.Iinv[3 * i] * t[0]));
  }


#pragma unroll
  for (int i = 0; i < 9; ++i) m.M[i] = M[i];
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    m.vbr[i] = vbr[i]; m.vb[i] = vb[i]; m.F[i] = F[i]; m.lq[i] = lq[i]; m.acc[i] = acc[i]; m.Iw[i] = Iw[i];
  }
#pragma unroll
  for (int i = 0; i < 6; ++i) m.C[i] = C[i];
  m.inv = inv; m.xa = xa; m.invV = invV; m.inv_rxz = inv_rxz; m.ida = ida; m.sb = sb;
  m.qbar = qbar; m.alpha = alpha; m.beta = beta; m.qS = qS; m.sgn = sgn; m.sc = sc;
}

AIRC_HD void f_jac_tail(const DevConst& P, const float v[3], const float q[4], const float w[3], const Ctl& c,
                        const JacMid& m, float* r) {
  const float* M = m.M;
  const float* vbr = m.vbr;
  const float* vb = m.vb;
  const float* C = m.C;
  const float* F = m.F;
  const float* lq = m.lq;
  const float* acc = m.acc;
  const float* Iw = m.Iw;
  const float inv = m.inv, xa = m.xa, invV = m.invV, inv_rxz = m.inv_rxz, ida = m.ida, sb = m.sb;
  const float alpha = m.alpha, beta = m.beta, qS = m.qS, sgn = m.sgn, sc = m.sc;
  // ---- D = d v_b / d(v, q):  u = M^T v = (w^2-|qv|^2) v + 2 qv (qv.v) - 2 w (qv x v)
  float D[3][7];
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) D[i][j] = M[3 * j + i] * inv;
  {
    const float x = q[0], y = q[1], z = q[2], ww = q[3];
    const float s = fmaf(z, v[2], fmaf(y, v[1], x * v[0]));
    const float cx[3] = {y * v[2] - z * v[1], z * v[0] - x * v[2], x * v[1] - y * v[0]};
    const float qv[3] = {x, y, z};
    float du[4][3];                                                   // d u / d(x, y, z, w)
    // e_j x v
    const float ev[3][3] = {{0.0f, -v[2], v[1]}, {v[2], 0.0f, -v[0]}, {-v[1], v[0], 0.0f}};
#pragma unroll
    for (int j = 0; j < 3; ++j)
#pragma unroll
      for (int i = 0; i < 3; ++i) {
        float a = s;                                                  // i == j: the two cross terms cancel
        if (i != j) a = fmaf(-ww, ev[j][i], fmaf(v[j], qv[i], -qv[j] * v[i]));
        du[j][i] = a;
      }
#pragma unroll
    for (int i = 0; i < 3; ++i) du[3][i] = fmaf(ww, v[i], -cx[i]);
    const float i2 = 2.0f * inv;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float qj2 = q[j] * i2;                                    // d(|q|^2)/dq_j / |q|^2
#pragma unroll
      for (int i = 0; i < 3; ++i) D[i][3 + j] = fmaf(du[j][i], i2, -vbr[i] * qj2);
    }
  }
  // ---- Y = G D,  G = d(qbar, alpha, beta)/d v_b
  float Y[3][7];
  {
    const float g0[3] = {1.225f * vb[0], 1.225f * vb[1], 1.225f * vb[2]};
    const float g1x = -vb[2] * ida, g1z = xa * ida;
    const float t = -sb * invV * inv_rxz;                             // d beta = (d vb_y - sb dV)/(V cos beta)
    const float g2[3] = {t * vb[0], fmaf(t, vb[1], inv_rxz), t * vb[2]};
#pragma unroll
    for (int cc = 0; cc < 7; ++cc) {
      Y[0][cc] = fmaf(g0[2], D[2][cc], fmaf(g0[1], D[1][cc], g0[0] * D[0][cc]));
      Y[1][cc] = fmaf(g1z, D[2][cc], g1x * D[0][cc]);
      Y[2][cc] = fmaf(g2[2], D[2][cc], fmaf(g2[1], D[1][cc], g2[0] * D[0][cc]));
    }
  }

  // ---- surrogate, derivative pass
  float J[6][5];
  surrogate_jac(P.net, m.rr, J);
  if (P.stall_scaling) {                                              // aircraft.py:280-294
    const float k = 10.0f, a30 = 0.52359877559f;
    const float sa = 1.0f / (1.0f + expf(k * (fabsf(alpha) - a30)));
    const float sbs = 1.0f / (1.0f + expf(k * (fabsf(beta) - a30)));
    const float ca = -k * sa * (1.0f - sa) * (alpha < 0.0f ? -1.0f : 1.0f);
    const float cb = -k * sbs * (1.0f - sbs) * (beta < 0.0f ? -1.0f : 1.0f);
    const float s2 = sa * sbs;
#pragma unroll
    for (int i = 0; i < 5; ++i) { J[2][i] *= s2; J[4][i] *= sa; }
    J[2][1] = fmaf(m.Craw2 * sbs, ca, J[2][1]);
    J[2][2] = fmaf(m.Craw2 * sa, cb, J[2][2]);
    J[4][1] = fmaf(m.Craw4, ca, J[4][1]);
  }

  // ---- Z = d(F, m)/d(qbar, alpha, beta);  F_k = sigma_k qS C_k,  m_k = l_k qbar C_{3+k}
  const float sig[3] = {sgn, 1.0f, 1.0f}, ell[3] = {P.Sb, P.Sc, P.Sb};
  float Z[6][3];
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const float fq = sig[k] * qS, mq = lq[k];
#pragma unroll
    for (int a = 0; a < 3; ++a) { Z[k][a] = fq * J[k][a]; Z[3 + k][a] = mq * J[3 + k][a]; }
    Z[k][0] = fmaf(sig[k] * P.S, C[k], Z[k][0]);
    Z[3 + k][0] = fmaf(ell[k], C[3 + k], Z[3 + k][0]);
  }
  // 6-vector (dF, dm) -> (d v_dot, d w_dot) at fixed attitude and rates
  auto push = [&](const float dF[3], const float dm[3], float pv[3], float pw[3]) {
    const float t[3] = {dm[0] + (P.com[1] * dF[2] - P.com[2] * dF[1]), dm[1] + (P.com[2] * dF[0] - P.com[0] * dF[2]),
                        dm[2] + (P.com[0] * dF[1] - P.com[1] * dF[0])};
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      pv[i] = fmaf(M[3 * i + 2], dF[2], fmaf(M[3 * i + 1], dF[1], M[3 * i] * dF[0])) * sc;
      pw[i] = fmaf(P.Iinv[3 * i + 2], t[2], fmaf(P.Iinv[3 * i + 1], t[1], P.Iinv[3 * i] * t[0]));
    }
  };
  float WP[3][3], WR[3][3];
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    const float dF[3] = {Z[0][a], Z[1][a], Z[2][a]}, dm[3] = {Z[3][a], Z[4][a], Z[5][a]};
    float pv[3], pw[3];
    push(dF, dm, pv, pw);
#pragma unroll
    for (int i = 0; i < 3; ++i) { WP[i][a] = pv[i]; WR[i][a] = pw[i]; }
  }
  // ---- [P; R](v, q) = W Y (+ direct terms of the q columns)
#pragma unroll
  for (int i = 0; i < 3; ++i) {
#pragma unroll
    for (int cc = 0; cc < 7; ++cc) {
      const float pv = fmaf(WP[i][2], Y[2][cc], fmaf(WP[i][1], Y[1][cc], WP[i][0] * Y[0][cc]));
      const float pw = fmaf(WR[i][2], Y[2][cc], fmaf(WR[i][1], Y[1][cc], WR[i][0] * Y[0][cc]));
      if (cc < 3) { r[JR_PV + 3 * i + cc] = pv; r[JR_RV + 3 * i + cc] = pw; }
      else { r[JR_PQ + 4 * i + (cc - 3)] = pv; r[JR_RQ + 4 * i + (cc - 3)] = pw; }
    }
  }
  {  // d(M F)/dq_j at fixed F:  M F = (w^2-|qv|^2) F + 2 qv (qv.F) + 2 w (qv x F);  and d(1/|q|^2)
    const float x = q[0], y = q[1], z = q[2], ww = q[3];
    const float s = fmaf(z, F[2], fmaf(y, F[1], x * F[0]));
    const float cx[3] = {y * F[2] - z * F[1], z * F[0] - x * F[2], x * F[1] - y * F[0]};
    const float qv[3] = {x, y, z};
    const float eF[3][3] = {{0.0f, -F[2], F[1]}, {F[2], 0.0f, -F[0]}, {-F[1], F[0], 0.0f}};
    const float s2 = 2.0f * sc, i2 = 2.0f * inv;
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      const float qj2 = q[j] * i2;
#pragma unroll
      for (int i = 0; i < 3; ++i) {
        float a = s;
        if (i != j) a = fmaf(ww, eF[j][i], fmaf(F[j], qv[i], -qv[j] * F[i]));
        r[JR_PQ + 4 * i + j] += fmaf(a, s2, -acc[i] * qj2);
      }
    }
    const float qw2 = ww * i2;
#pragma unroll
    for (int i = 0; i < 3; ++i) r[JR_PQ + 4 * i + 3] += fmaf(fmaf(ww, F[i], cx[i]), s2, -acc[i] * qw2);
  }
  // ---- d(w_dot)/dw = -I^-1 (e_j x Iw + w x I e_j)                     base.py:376-377
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    const float Ij[3] = {P.I[j], P.I[3 + j], P.I[6 + j]};
    const float eI[3][3] = {{0.0f, -Iw[2], Iw[1]}, {Iw[2], 0.0f, -Iw[0]}, {-Iw[1], Iw[0], 0.0f}};
    const float t[3] = {eI[j][0] + (w[1] * Ij[2] - w[2] * Ij[1]), eI[j][1] + (w[2] * Ij[0] - w[0] * Ij[2]),
                        eI[j][2] + (w[0] * Ij[1] - w[1] * Ij[0])};
#pragma unroll
    for (int i = 0; i < 3; ++i)
      r[JR_RW + 3 * i + j] = -fmaf(P.Iinv[3 * i + 2], t[2], fmaf(P.Iinv[3 * i + 1], t[1], P.Iinv[3 * i] * t[0]));
  }
  // ---- control columns: aileron, elevator through J[:,3], J[:,4]; rudder (C_n only); flaps (CX, CZ)
#pragma unroll
  for (int cc = 0; cc < 2; ++cc) {
    const float dF[3] = {sig[0] * qS * J[0][3 + cc], qS * J[1][3 + cc], qS * J[2][3 + cc]};
    const float dm[3] = {lq[0] * J[3][3 + cc], lq[1] * J[4][3 + cc], lq[2] * J[5][3 + cc]};
    float pv[3], pw[3];
    push(dF, dm, pv, pw);
#pragma unroll
    for (int i = 0; i < 3; ++i) { r[JR_FU + 4 * i + cc] = pv[i]; r[JR_FU + 4 * (3 + i) + cc] = pw[i]; }
  }
  {
    const float dm2 = lq[2] * P.k_rudder;
#pragma unroll
    for (int i = 0; i < 3; ++i) { r[JR_FU + 4 * i + 2] = 0.0f; r[JR_FU + 4 * (3 + i) + 2] = P.Iinv[3 * i + 2] * dm2; }
    const float dF[3] = {sig[0] * qS * P.k_flap_x, 0.0f, qS * P.k_flap_z}, dm[3] = {0.0f, 0.0f, 0.0f};
    float pv[3], pw[3];
    push(dF, dm, pv, pw);
#pragma unroll
    for (int i = 0; i < 3; ++i) { r[JR_FU + 4 * i + 3] = pv[i]; r[JR_FU + 4 * (3 + i) + 3] = pw[i]; }
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) r[JR_Q + i] = q[i];
#pragma unroll
  for (int i = 0; i < 3; ++i) r[JR_W + i] = w[i];
}

AIRC_HD void f_jac(const DevConst& P, const float v[3], const float q[4], const float w[3], const Ctl& c,
                   float kv[3], float kq[4], float kw[3], float* r) {
  JacMid m;
  f_jac_head(P, v, q, w, c, kv, kq, kw, m);
  f_jac_tail(P, v, q, w, c, m, r);
}

// Phase 1 for one node: the primal RK4 step (x is advanced, NOT normalised) and the node record `rec`
// (JRec<DT>::NODE floats, 16-byte aligned).
template <bool DT>
AIRC_HD void jac_phase1(const DevConst& P, State& x, const Ctl& c, float h, float* rec) {
  using L = JRec<DT>;
  const float hh = 0.5f * h, h6 = h * (1.0f / 6.0f);
  float sv[3], sq[4], sw[3], ap[3], av[3], aq[4], aw[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) { sv[i] = x.v[i]; sw[i] = x.w[i]; ap[i] = av[i] = aw[i] = 0.0f; }
#pragma unroll
  for (int i = 0; i < 4; ++i) { sq[i] = x.q[i]; aq[i] = 0.0f; }
#pragma unroll 1
  for (int stage = 0; stage < 4; ++stage) {
    const float wgt = (stage == 0 || stage == 3) ? 1.0f : 2.0f;
    const float cn = (stage == 2) ? h : hh;
    float kv[3], kq[4], kw[3];
    float r[92];
    r[82] = r[83] = 0.0f;
    f_jac(P, sv, sq, sw, c, kv, kq, kw, r);
    if (DT) {
#pragma unroll
      for (int i = 0; i < 3; ++i) { r[JR_K + i] = kv[i]; r[JR_K + 7 + i] = kw[i]; }
#pragma unroll
      for (int i = 0; i < 4; ++i) r[JR_K + 3 + i] = kq[i];
    }
    Q4* dst = reinterpret_cast<Q4*>(rec + stage * L::STAGE);
#pragma unroll
    for (int i = 0; i < L::STAGE / 4; ++i) dst[i] = Q4{r[4 * i], r[4 * i + 1], r[4 * i + 2], r[4 * i + 3]};
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      ap[i] = fmaf(wgt, sv[i], ap[i]);
      av[i] = fmaf(wgt, kv[i], av[i]);
      aw[i] = fmaf(wgt, kw[i], aw[i]);
      sv[i] = fmaf(cn, kv[i], x.v[i]);
      sw[i] = fmaf(cn, kw[i], x.w[i]);
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      aq[i] = fmaf(wgt, kq[i], aq[i]);
      sq[i] = fmaf(cn, kq[i], x.q[i]);
    }
  }
  Q4* hd = reinterpret_cast<Q4*>(rec + L::HDR);
  hd[0] = Q4{h, ap[0], ap[1], ap[2]};
  hd[1] = Q4{aq[0], aq[1], aq[2], aq[3]};
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    x.p[i] = fmaf(h6, ap[i], x.p[i]);
    x.v[i] = fmaf(h6, av[i], x.v[i]);
    x.w[i] = fmaf(h6, aw[i], x.w[i]);
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) x.q[i] = fmaf(h6, aq[i], x.q[i]);
}

// Phase 2 for one (node, column pair): the tangent recursion over the parked blocks.  `pair` selects the
// column slots (2 pair, 2 pair + 1).  col[r] receives rows p(0:3) v(3:6) q(6:10) w(10:13) of the two columns.
// DT = false drops the arithmetic of the d/d(dt) slot (pair 7 is then not evaluated by the caller).
template <bool DT>
AIRC_HD void jac_phase2(const float* rec, int pair, bool normalise, F2 col[13]) {
  using L = JRec<DT>;
  const int s0 = 2 * pair, s1 = s0 + 1;
  auto seed = [&](int slot) { return F2{s0 == slot ? 1.0f : 0.0f, s1 == slot ? 1.0f : 0.0f}; };
  const bool ctl = (pair == 5) || (pair == 6);
  const int fu_off = JR_FU + ((pair == 6) ? 2 : 0);
  const Q4* hd = reinterpret_cast<const Q4*>(rec + L::HDR);
  const Q4 h0 = hd[0], h1 = hd[1];
  const float h = h0.a;
  const float hh = 0.5f * h, h6 = h * (1.0f / 6.0f);
  const F2 ddt = DT ? seed(14) : F2{0.0f, 0.0f};
  F2 Dv[3], Dq[4], Dw[3], tv[3], tq[4], tw[3], bp[3], bv[3], bq[4], bw[3];
  float av[3] = {0.0f, 0.0f, 0.0f}, aw[3] = {0.0f, 0.0f, 0.0f};   // sum_i w_i k_i (d/d(dt) column only)
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    Dv[i] = tv[i] = seed(i);
    Dw[i] = tw[i] = seed(7 + i);
    bp[i] = bv[i] = bw[i] = F2{0.0f, 0.0f};
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) { Dq[i] = tq[i] = seed(3 + i); bq[i] = F2{0.0f, 0.0f}; }
  float q1[4] = {0.0f, 0.0f, 0.0f, 0.0f};          // stage-1 q (the node's own), for the normalisation

#pragma unroll 1
  for (int stage = 0; stage < 4; ++stage) {
    const float wgt = (stage == 0 || stage == 3) ? 1.0f : 2.0f;
    const float cn = (stage == 2) ? h : hh;
    const Q4* r4 = reinterpret_cast<const Q4*>(rec + stage * L::STAGE);
    float r[92];
#pragma unroll
    for (int i = 0; i < L::STAGE / 4; ++i) {
      const Q4 t = r4[i];
      r[4 * i] = t.a; r[4 * i + 1] = t.b; r[4 * i + 2] = t.c; r[4 * i + 3] = t.d;
    }
    if (stage == 0) {
#pragma unroll
      for (int i = 0; i < 4; ++i) q1[i] = r[JR_Q + i];
    }
    // K = f_x S + f_u
    F2 dkv[3], dkq[4], dkw[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      // the control columns start from their f_u entries (lanes 5, 6), all others from zero
      const float* fu = rec + stage * L::STAGE + fu_off;
      F2 a = ctl ? F2{fu[4 * i], fu[4 * i + 1]} : F2{0.0f, 0.0f};
      F2 b = ctl ? F2{fu[4 * (3 + i)], fu[4 * (3 + i) + 1]} : F2{0.0f, 0.0f};
#pragma unroll
      for (int j = 0; j < 3; ++j) {
        a = tfma(r[JR_PV + 3 * i + j], tv[j], a);
        b = tfma(r[JR_RV + 3 * i + j], tv[j], b);
        b = tfma(r[JR_RW + 3 * i + j], tw[j], b);
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        a = tfma(r[JR_PQ + 4 * i + j], tq[j], a);
        b = tfma(r[JR_RQ + 4 * i + j], tq[j], b);
      }
      dkv[i] = a;
      dkw[i] = b;
    }
    {  // d(0.5 q (x) (w,0)) = 0.5 (dq (x) (w,0) + q (x) (dw,0))            base.py:292-295
      const float hq[4] = {0.5f * r[JR_Q], 0.5f * r[JR_Q + 1], 0.5f * r[JR_Q + 2], 0.5f * r[JR_Q + 3]};
      const float hw[3] = {0.5f * r[JR_W], 0.5f * r[JR_W + 1], 0.5f * r[JR_W + 2]};
      dkq[0] = tfma(hw[0], tq[3], tfma(hw[2], tq[1], tfma(-hw[1], tq[2],
               tfma(hq[3], tw[0], tfma(hq[1], tw[2], tmul(-hq[2], tw[1]))))));
      dkq[1] = tfma(hw[1], tq[3], tfma(hw[0], tq[2], tfma(-hw[2], tq[0],
               tfma(hq[3], tw[1], tfma(hq[2], tw[0], tmul(-hq[0], tw[2]))))));
      dkq[2] = tfma(hw[2], tq[3], tfma(hw[1], tq[0], tfma(-hw[0], tq[1],
               tfma(hq[3], tw[2], tfma(hq[0], tw[1], tmul(-hq[1], tw[0]))))));
      dkq[3] = tfma(-hw[0], tq[0], tfma(-hw[1], tq[1], tfma(-hw[2], tq[2],
               tfma(-hq[0], tw[0], tfma(-hq[1], tw[1], tmul(-hq[2], tw[2]))))));
    }
    // accumulate, next stage tangent  S = S_1 + c_i h K (+ c_i k d(h) for the step-size column)
    const F2 dcn = tmul((stage == 2) ? 1.0f : 0.5f, ddt);
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      bp[i] = tfma(wgt, tv[i], bp[i]);
      bv[i] = tfma(wgt, dkv[i], bv[i]);
      bw[i] = tfma(wgt, dkw[i], bw[i]);
      tv[i] = tfma(cn, dkv[i], Dv[i]);
      tw[i] = tfma(cn, dkw[i], Dw[i]);
      if (DT) {
        tv[i] = tfma(r[JR_K + i], dcn, tv[i]);
        tw[i] = tfma(r[JR_K + 7 + i], dcn, tw[i]);
        av[i] = fmaf(wgt, r[JR_K + i], av[i]);
        aw[i] = fmaf(wgt, r[JR_K + 7 + i], aw[i]);
      }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      bq[i] = tfma(wgt, dkq[i], bq[i]);
      tq[i] = tfma(cn, dkq[i], Dq[i]);
      if (DT) tq[i] = tfma(r[JR_K + 3 + i], dcn, tq[i]);
    }
  }
  // [A | B | dPhi/ddt] = S_1 + h/6 b (+ a/6 d(h))
  const float ap[3] = {h0.b, h0.c, h0.d}, aq[4] = {h1.a, h1.b, h1.c, h1.d};
  const F2 dh6 = tmul(1.0f / 6.0f, ddt);
  F2 Dp[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    Dp[i] = tmul(h6, bp[i]);
    Dv[i] = tfma(h6, bv[i], Dv[i]);
    Dw[i] = tfma(h6, bw[i], Dw[i]);
    if (DT) {
      Dp[i] = tfma(ap[i], dh6, Dp[i]);
      Dv[i] = tfma(av[i], dh6, Dv[i]);
      Dw[i] = tfma(aw[i], dh6, Dw[i]);
    }
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    Dq[i] = tfma(h6, bq[i], Dq[i]);
    if (DT) Dq[i] = tfma(aq[i], dh6, Dq[i]);
  }
  if (normalise) {                                                    // base.py:443-444 with its tangent
    float qn[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) qn[i] = fmaf(h6, aq[i], q1[i]);
    const float n2 = (qn[0] * qn[0] + qn[1] * qn[1]) + (qn[2] * qn[2] + qn[3] * qn[3]);
    float n, rn;
    sqrt_rsqrt(n2, n, rn);
#pragma unroll
    for (int i = 0; i < 4; ++i) qn[i] *= rn;
    const F2 d = tadd(tsum<F2>(qn[0], Dq[0], qn[1], Dq[1]), tsum<F2>(qn[2], Dq[2], qn[3], Dq[3]));
#pragma unroll
    for (int i = 0; i < 4; ++i) Dq[i] = tmul(rn, tfma(-qn[i], d, Dq[i]));
  }
#pragma unroll
  for (int i = 0; i < 3; ++i) { col[i] = Dp[i]; col[3 + i] = Dv[i]; col[10 + i] = Dw[i]; }
#pragma unroll
  for (int i = 0; i < 4; ++i) col[6 + i] = Dq[i];
}

}  // namespace airc

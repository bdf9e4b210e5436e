// airc_alt.cuh -- the other coefficient models of the reference's registry
// (/root/reference/src/aircraft/dynamics/coefficient_models.py:32-37), primal evaluation only:
//   AIRC_MODEL_POLY     PolynomialModel  coefficient_models.py:106-135  (what main/control/control.py:166,
//                       main/mhe/mhtt.py:49, main/rl/rl.py:36, main/dynamics/dynamics.py:35 run with)
//   AIRC_MODEL_LINEAR   LinearModel      coefficient_models.py:80-89
//   AIRC_MODEL_DEFAULT  DefaultModel     coefficient_models.py:41-78
// They plug into f_eval_net through the same policy interface as the network (eval + adjust) and are
// served by the one-node-per-thread kernels (state_derivative, state_update, diagnostics, rollout) and,
// for derivatives, by the jets of airc_jet.cuh / airc_hess.cuh; the packed rollout, the sampling sweep and
// the hand-derived tangents of K2 are NN-only.
// Included from the middle of airc_dynamics.cuh (needs DevConst, Stage, Ctl).
#pragma once

namespace airc {

constexpr int AIRC_MODEL_NN_ = 0, AIRC_MODEL_POLY_ = 1, AIRC_MODEL_LINEAR_ = 2, AIRC_MODEL_DEFAULT_ = 3;
constexpr int NPOLY = 34;   // monomials of degree 1..3 in (alpha, beta, aileron, elevator), sklearn order

// Lives in the bytes of DevConst::net when DevConst::model != 0 (the network is not needed then), so
// the kernels keep one signature and one constant block.
struct AltParams {
  float poly[6][NPOLY + 1];   // [k][0..33] weights, [k][34] intercept      (main/surrogates/polynomial_model.py:76-104)
  float lin[6][6];            // [k][q, alpha, beta, aileron, elevator, 1]  (data/networks/linearised.csv)
  float arm;                  // rudder_moment_arm                          (utils.py:215)
  float span;
  float k_rudder_model;       // the model's own rudder effectiveness * 6 pi/180 (poly: +0.01, others -0.1)
};
static_assert(sizeof(AltParams) <= sizeof(FusedNet), "AltParams must fit in the network's storage");

AIRC_HD const AltParams& alt_params(const DevConst& P) { return *reinterpret_cast<const AltParams*>(&P.net); }

// sklearn PolynomialFeatures(degree=3, include_bias=False) order for 4 inputs:
// x_i ; x_i x_j (i<=j) ; x_i x_j x_k (i<=j<=k)
AIRC_HD void poly_monomials(const float x[4], float m[NPOLY]) {
  int t = 0;
#pragma unroll
  for (int i = 0; i < 4; ++i) m[t++] = x[i];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = i; j < 4; ++j) m[t++] = x[i] * x[j];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = i; j < 4; ++j)
#pragma unroll
      for (int k = j; k < 4; ++k) m[t++] = x[i] * x[j] * x[k];
}
AIRC_HD float poly_dot(const float w[NPOLY + 1], const float m[NPOLY]) {
  float a = w[NPOLY];
#pragma unroll
  for (int t = 0; t < NPOLY; ++t) a = fmaf(w[t], m[t], a);
  return a;
}
// a polynomial at (alpha, 0, 0, 0): only the monomials alpha, alpha^2, alpha^3 (indices 0, 4, 14) survive
AIRC_HD float poly_alpha_only(const float w[NPOLY + 1], float a) {
  return fmaf(a, fmaf(a, fmaf(a, w[14], w[4]), w[0]), w[NPOLY]);
}

struct NetAlt {
  static constexpr bool kPolyTrig = false;   // table models keep the CUDA math library atan2f / asinf
  const AltParams* A;

  template <bool WITH_JAC>
  AIRC_HD void eval(float qbar, float alpha, float beta, const Ctl& c, float C[6], float J[6][5]) const {
    (void)J;
    // all three models are evaluated through P.model in adjust(), where the body rates are available;
    // here only the inputs every model shares
    C[0] = qbar; C[1] = alpha; C[2] = beta; C[3] = c.da; C[4] = c.de; C[5] = 0.0f;
  }

  AIRC_HD void adjust(const DevConst& P, Stage& s, const float* w, const Ctl& c) const {
    const float alpha = s.alpha, beta = s.beta;
    float C[6];
    if (P.model == AIRC_MODEL_POLY_) {
      const float u = s.vb[0], v = s.vb[1], ww = s.vb[2];
      const float xu = u + P.eps;
      const float a_el = atan2_acc(fmaf(A->arm, w[1], ww), xu);                    // aircraft.py:189-199
      const float a_lw = atan2_acc(fmaf(-0.25f * A->span, w[0], ww), xu);           // aircraft.py:202-211
      const float a_rw = atan2_acc(fmaf(0.25f * A->span, w[0], ww), xu);            // aircraft.py:213-222
      const float vr = fmaf(-A->arm, w[2], v);                                     // aircraft.py:224-233
      const float b_rud = asin_acc(vr * fast_rsqrt(fmaf(u, u, fmaf(vr, vr, fmaf(ww, ww, P.eps)))));
      float m[NPOLY];
      const float x0[4] = {alpha, beta, c.da, c.de};
      poly_monomials(x0, m);
#pragma unroll
      for (int k = 0; k < 4; ++k) C[k] = poly_dot(A->poly[k], m);
      C[3] = fmaf(0.125f * A->span, poly_alpha_only(A->poly[2], a_rw) - poly_alpha_only(A->poly[2], a_lw), C[3]);
      const float x1[4] = {a_el, beta, c.da, c.de};
      poly_monomials(x1, m);
      C[4] = poly_dot(A->poly[4], m);
      const float x2[4] = {alpha, b_rud, c.da, c.de};
      poly_monomials(x2, m);
      C[5] = poly_dot(A->poly[5], m);
    } else if (P.model == AIRC_MODEL_LINEAR_) {
#pragma unroll
      for (int k = 0; k < 6; ++k) {
        const float* L = A->lin[k];
        C[k] = fmaf(L[0], s.qbar, fmaf(L[1], alpha, fmaf(L[2], beta, fmaf(L[3], c.da, fmaf(L[4], c.de, L[5])))));
      }
    } else {                                                                     // DefaultModel
      const float d2r = 0.01745329252f;
      C[0] = -(0.02f + 0.3f * alpha * alpha);
      C[1] = -0.98f * beta;
      C[2] = -(5.0f * alpha);
      C[3] = fmaf(0.08f * 4.0f * d2r, c.da, -0.05f * w[0]);
      C[4] = fmaf(-1.2f * 5.0f * d2r, c.de, -0.5f * w[1]);
      C[5] = -0.05f * w[2];
    }
    // f_eval_net adds P.k_rudder * rudder afterwards (the NN model's -0.1 * 6 pi/180): make the sum the model's own
    C[5] = fmaf(A->k_rudder_model - P.k_rudder, c.dr, C[5]);
#pragma unroll
    for (int k = 0; k < 6; ++k) s.C[k] = C[k];
  }
};

}  // namespace airc

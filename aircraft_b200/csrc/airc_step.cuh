// airc_step.cuh -- classical RK4 step of the 13-state system, with NT forward-mode tangent
// columns carried alongside (NT = 0: primal only).
//
// Restated from /root/reference/src/aircraft/dynamics/base.py:408-446 (state_step):
//   k1 = f(x), k2 = f(x + h/2 k1), k3 = f(x + h/2 k2), k4 = f(x + h k3),
//   x+ = x + h/6 (k1 + 2 k2 + 2 k3 + k4)            -- all 13 components additively,
// optional q <- q/|q| afterwards (base.py:443-444).
#pragma once
#include "airc_dynamics.cuh"

namespace airc {

struct State {
  float p[3], v[3], q[4], w[3];
};

// one tangent column: d(state) plus the (constant) control direction it belongs to
struct Tangent {
  float p[3], v[3], q[4], w[3];
  Ctl dc;
};

template <int NT, int MODEL = 0>
AIRC_HD void rk4_step(const DevConst& P, State& x, const Ctl& c, float h, Tangent* T) {
  const float hh = 0.5f * h, h6 = h * (1.0f / 6.0f);
  // stage state (v,q,w only: f does not read p) and accumulators
  float sv[3], sq[4], sw[3];
  float ap[3], av[3], aq[4], aw[3];
  // tangent stage states and accumulators
  float tv[NT > 0 ? NT : 1][3], tq[NT > 0 ? NT : 1][4], tw[NT > 0 ? NT : 1][3];
  float bp[NT > 0 ? NT : 1][3], bv[NT > 0 ? NT : 1][3], bq[NT > 0 ? NT : 1][4],
      bw[NT > 0 ? NT : 1][3];
#pragma unroll
  for (int i = 0; i < 3; ++i) { sv[i] = x.v[i]; sw[i] = x.w[i]; ap[i] = av[i] = aw[i] = 0.0f; }
#pragma unroll
  for (int i = 0; i < 4; ++i) { sq[i] = x.q[i]; aq[i] = 0.0f; }
#pragma unroll
  for (int t = 0; t < NT; ++t) {
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      tv[t][i] = T[t].v[i]; tw[t][i] = T[t].w[i];
      bp[t][i] = bv[t][i] = bw[t][i] = 0.0f;
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) { tq[t][i] = T[t].q[i]; bq[t][i] = 0.0f; }
  }

#pragma unroll 1
  for (int stage = 0; stage < 4; ++stage) {
    const float wgt = (stage == 0 || stage == 3) ? 1.0f : 2.0f;  // RK4 weights 1,2,2,1
    const float cn = (stage == 2) ? h : hh;                       // next stage offset h/2,h/2,h
    float kv[3], kq[4], kw[3];
    Stage s;
    float J[6][5];
    f_eval<(NT > 0), MODEL>(P, sv, sq, sw, c, kv, kq, kw, s, J);
    // tangents first: they need the stage state before it is overwritten
#pragma unroll
    for (int t = 0; t < NT; ++t) {
      float dkv[3], dkq[4], dkw[3];
      f_tangent(P, sv, sq, sw, s, J, tv[t], tq[t], tw[t], T[t].dc, dkv, dkq, dkw);
#pragma unroll
      for (int i = 0; i < 3; ++i) {
        bp[t][i] = fmaf(wgt, tv[t][i], bp[t][i]);   // d(p_dot) = dv at this stage
        bv[t][i] = fmaf(wgt, dkv[i], bv[t][i]);
        bw[t][i] = fmaf(wgt, dkw[i], bw[t][i]);
        tv[t][i] = fmaf(cn, dkv[i], T[t].v[i]);
        tw[t][i] = fmaf(cn, dkw[i], T[t].w[i]);
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        bq[t][i] = fmaf(wgt, dkq[i], bq[t][i]);
        tq[t][i] = fmaf(cn, dkq[i], T[t].q[i]);
      }
    }
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      ap[i] = fmaf(wgt, sv[i], ap[i]);              // p_dot = v at this stage
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
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    x.p[i] = fmaf(h6, ap[i], x.p[i]);
    x.v[i] = fmaf(h6, av[i], x.v[i]);
    x.w[i] = fmaf(h6, aw[i], x.w[i]);
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) x.q[i] = fmaf(h6, aq[i], x.q[i]);
#pragma unroll
  for (int t = 0; t < NT; ++t) {
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      T[t].p[i] = fmaf(h6, bp[t][i], T[t].p[i]);
      T[t].v[i] = fmaf(h6, bv[t][i], T[t].v[i]);
      T[t].w[i] = fmaf(h6, bw[t][i], T[t].w[i]);
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) T[t].q[i] = fmaf(h6, bq[t][i], T[t].q[i]);
  }
}

// q <- q/|q| and its tangent (liecasadi Quaternion.normalize(); base.py:443-444, 473-474)
template <int NT>
AIRC_HD void normalise_q(State& x, Tangent* T) {
  const float n2 = (x.q[0] * x.q[0] + x.q[1] * x.q[1]) + (x.q[2] * x.q[2] + x.q[3] * x.q[3]);
  float n, rn;
  sqrt_rsqrt(n2, n, rn);
#pragma unroll
  for (int i = 0; i < 4; ++i) x.q[i] *= rn;
#pragma unroll
  for (int t = 0; t < NT; ++t) {
    const float d = (x.q[0] * T[t].q[0] + x.q[1] * T[t].q[1]) + (x.q[2] * T[t].q[2] + x.q[3] * T[t].q[3]);
#pragma unroll
    for (int i = 0; i < 4; ++i) T[t].q[i] = (T[t].q[i] - x.q[i] * d) * rn;
  }
}

// Phi(x,u,dt) of base.py:450-480: n substeps of h = dt/n without normalisation, one at the end.
template <int NT, int MODEL = 0>
AIRC_HD void state_update(const DevConst& P, State& x, const Ctl& c, float dt, int substeps,
                          bool normalise, Tangent* T) {
  static_assert(NT == 0 || MODEL == 0, "tangents exist for the NN model only");
  const float h = (substeps == 1) ? dt : dt / (float)substeps;
  for (int s = 0; s < substeps; ++s) rk4_step<NT, MODEL>(P, x, c, h, T);
  if (normalise) normalise_q<NT>(x, T);
}

}  // namespace airc

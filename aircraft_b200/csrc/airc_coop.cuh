// airc_coop.cuh -- building blocks of the cooperative Jacobian kernel (K2).
//
// L lanes of a warp work on one node.  Each lane
//   * evaluates 16/L of the 16 hidden-unit pairs of the surrogate and their contribution to the
//     6 outputs and to the 6x5 network Jacobian Wo.diag(g).A; the 36 partial sums are all-reduced
//     over the L lanes with xor-shuffles (a butterfly: every lane ends with bit-identical sums, so
//     the L redundant copies of the primal state stay identical);
//   * carries ONE packed tangent (two Jacobian columns) through the four RK4 stages with FFMA2.
// With L = 1 the group degenerates to one thread doing everything (no shuffles): that
// instantiation is host-compilable and is what oracle/host_emul.cpp checks on the CPU.
#pragma once
#include "airc_step.cuh"
#include "airc_tangent.cuh"

namespace airc {

template <int L>
AIRC_HD float group_sum(float v) {
#if defined(__CUDA_ARCH__)
#pragma unroll
  for (int off = L / 2; off > 0; off >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, off);
#endif
  return v;
}

// two sums per packed add: 2 shuffles + 1 FADD2 per butterfly level instead of 2 + 2
template <int L>
AIRC_HD F2 group_sum2(F2 v) {
#if defined(__CUDA_ARCH__)
#pragma unroll
  for (int off = L / 2; off > 0; off >>= 1)
    v = add2b(v, F2{__shfl_xor_sync(0xFFFFFFFFu, v.lo, off), __shfl_xor_sync(0xFFFFFFFFu, v.hi, off)});
#endif
  return v;
}

template <int L>
struct NetCoop {
  static constexpr bool kPolyTrig = true;
  const FusedNet* net;   // shared-memory copy when L > 1 (lane-dependent addresses)
  int lane;              // 0..L-1 within the node group

  // Primal only.  The six sums are formed in ONE association whatever L is: eight "virtual lanes" v, each
  // accumulating the pairs (v, v+8), combined by the xor butterfly 4, 2, 1 -- the steps with offset >= L happen
  // inside the thread, the rest by shuffles.  The small-batch rollout kernel picks L by batch size, so this keeps
  // a sharded batch bit-identical to the whole one.
  AIRC_HD void eval_primal(float qbar, float alpha, float beta, const Ctl& c, float C[6]) const {
    constexpr int V = 8, NV = V / L;
    const F2 da = bc2(c.da), de = bc2(c.de), qb = bc2(qbar), al = bc2(alpha), be = bc2(beta);
    F2 part[NV][3];
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const int vl = lane + j * L;
      F2 acc[6];
#pragma unroll
      for (int k = 0; k < 6; ++k) acc[k] = F2{0.0f, 0.0f};
#pragma unroll
      for (int m = 0; m < 2; ++m) {
        const int p = vl + m * V;
        const HidPair& w = net->hid[p];
        F2 h = F2{w.a0[0], w.a0[1]};
        h = fma2b(F2{w.Au[0][0], w.Au[0][1]}, da, h);
        h = fma2b(F2{w.Au[1][0], w.Au[1][1]}, de, h);
        h = fma2b(F2{w.Ax[0][0], w.Ax[0][1]}, qb, h);
        h = fma2b(F2{w.Ax[1][0], w.Ax[1][1]}, al, h);
        h = fma2b(F2{w.Ax[2][0], w.Ax[2][1]}, be, h);
        const F2 r = sigm2_x2(h);
#pragma unroll
        for (int k = 0; k < 6; ++k) acc[k] = fma2b(F2{net->Wo[k][2 * p], net->Wo[k][2 * p + 1]}, r, acc[k]);
      }
#pragma unroll
      for (int kk = 0; kk < 3; ++kk)
        part[j][kk] = F2{acc[2 * kk].lo + acc[2 * kk].hi, acc[2 * kk + 1].lo + acc[2 * kk + 1].hi};
    }
#pragma unroll
    for (int off = V / 2; off >= L; off >>= 1) {
      const int st = off / L;
#pragma unroll
      for (int j = 0; j < st; ++j)
#pragma unroll
        for (int kk = 0; kk < 3; ++kk) part[j][kk] = add2b(part[j][kk], part[j + st][kk]);
    }
#pragma unroll
    for (int kk = 0; kk < 3; ++kk) {
      const F2 a = group_sum2<L>(part[0][kk]);
      C[2 * kk] = a.lo + net->c0[2 * kk];
      C[2 * kk + 1] = a.hi + net->c0[2 * kk + 1];
    }
  }

  template <bool WITH_JAC>
  AIRC_HD void eval(float qbar, float alpha, float beta, const Ctl& c, float C[6], float J[6][5]) const {
    if constexpr (!WITH_JAC) {
      eval_primal(qbar, alpha, beta, c, C);
      return;
    }
    F2 acc[6], Jacc[6][5];
#pragma unroll
    for (int k = 0; k < 6; ++k) {
      acc[k] = F2{0.0f, 0.0f};
#pragma unroll
      for (int i = 0; i < 5; ++i) Jacc[k][i] = F2{0.0f, 0.0f};
    }
    const F2 da = bc2(c.da), de = bc2(c.de), qb = bc2(qbar), al = bc2(alpha), be = bc2(beta);
#pragma unroll
    for (int m = 0; m < (NH / 2) / L; ++m) {
      const int p = lane + m * L;
      const HidPair& w = net->hid[p];
      F2 h = F2{w.a0[0], w.a0[1]};
      h = fma2b(F2{w.Au[0][0], w.Au[0][1]}, da, h);
      h = fma2b(F2{w.Au[1][0], w.Au[1][1]}, de, h);
      h = fma2b(F2{w.Ax[0][0], w.Ax[0][1]}, qb, h);
      h = fma2b(F2{w.Ax[1][0], w.Ax[1][1]}, al, h);
      h = fma2b(F2{w.Ax[2][0], w.Ax[2][1]}, be, h);
      const F2 r = sigm2_x2(h);
      const F2 g = mul2b(mul2b(r, bc2(-0.69314718056f)), F2{1.0f - r.lo, 1.0f - r.hi});
      F2 wo[6];
#pragma unroll
      for (int k = 0; k < 6; ++k) {
        wo[k] = F2{net->Wo[k][2 * p], net->Wo[k][2 * p + 1]};
        acc[k] = fma2b(wo[k], r, acc[k]);
      }
#pragma unroll
      for (int i = 0; i < 5; ++i) {
        const F2 wi = (i < 3) ? F2{w.Ax[i][0], w.Ax[i][1]} : F2{w.Au[i - 3][0], w.Au[i - 3][1]};
        const F2 gw = mul2b(g, wi);
#pragma unroll
        for (int k = 0; k < 6; ++k) Jacc[k][i] = fma2b(wo[k], gw, Jacc[k][i]);
      }
    }
    // 36 sums = 18 packed pairs: (C[k], J[k][0]) and (J[k][1], J[k][2]), (J[k][3], J[k][4])
#pragma unroll
    for (int k = 0; k < 6; ++k) {
      const F2 a = group_sum2<L>(F2{acc[k].lo + acc[k].hi, Jacc[k][0].lo + Jacc[k][0].hi});
      const F2 b = group_sum2<L>(F2{Jacc[k][1].lo + Jacc[k][1].hi, Jacc[k][2].lo + Jacc[k][2].hi});
      const F2 d = group_sum2<L>(F2{Jacc[k][3].lo + Jacc[k][3].hi, Jacc[k][4].lo + Jacc[k][4].hi});
      C[k] = a.lo + net->c0[k];
      J[k][0] = a.hi; J[k][1] = b.lo; J[k][2] = b.hi; J[k][3] = d.lo; J[k][4] = d.hi;
    }
  }
  AIRC_HD void adjust(const DevConst&, Stage&, const float*, const Ctl&) const {}
};

// Primal RK4 step with the network evaluated through a policy (NetCoop: shared by the lanes of a group).
// Same arithmetic as rk4_step<0> apart from the network's summation order.
// speed2 (optional) receives |v_b|^2 of the step's first stage point, i.e. of the node itself (the loss's speed term).
template <class Net>
AIRC_HD void rk4_step_net(const DevConst& P, const Net& net, State& x, const Ctl& c, float h, float* speed2 = nullptr) {
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
    Stage st;
    float J[6][5];
    f_eval_net<false>(P, net, sv, sq, sw, c, kv, kq, kw, st, J);
    if (stage == 0 && speed2) *speed2 = fmaf(st.vb[2], st.vb[2], fmaf(st.vb[1], st.vb[1], st.vb[0] * st.vb[0]));
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
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    x.p[i] = fmaf(h6, ap[i], x.p[i]);
    x.v[i] = fmaf(h6, av[i], x.v[i]);
    x.w[i] = fmaf(h6, aw[i], x.w[i]);
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) x.q[i] = fmaf(h6, aq[i], x.q[i]);
}

template <class Net>
AIRC_HD void state_update_net(const DevConst& P, const Net& net, State& x, const Ctl& c, float dt, int substeps,
                              bool normalise, float* speed2 = nullptr) {
  const float h = (substeps == 1) ? dt : dt / (float)substeps;
  rk4_step_net(P, net, x, c, h, speed2);
  for (int s = 1; s < substeps; ++s) rk4_step_net(P, net, x, c, h);
  if (normalise) normalise_q<0>(x, nullptr);
}

template <class T>
struct TangentT {
  T p[3], v[3], q[4], w[3];
  CtlT<T> dc;
  T ddt;   // d(dt) along this column: 1 for the dPhi/d(dt) column (control/base.py:276, time='progress'), else 0
};

// One RK4 step of the primal with ONE tangent of type T alongside (base.py:408-446 and its
// forward-mode derivative).  f does not read p, so d(p_dot) = dv.
// dh = d(h)/d(dt) * D.ddt (h = dt/substeps): the step-size direction of this column.  The stage
// states X_{i+1} = x + c_i h k_i and the update x+ = x + h/6 sum w_i k_i depend on h explicitly.
template <class T, class Net>
AIRC_HD void rk4_step_lin(const DevConst& P, const Net& net, State& x, const Ctl& c, float h, TangentT<T>& D,
                          const T& dh) {
  const float hh = 0.5f * h, h6 = h * (1.0f / 6.0f);
  float sv[3], sq[4], sw[3], ap[3], av[3], aq[4], aw[3];
  T tv[3], tq[4], tw[3], bp[3], bv[3], bq[4], bw[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    sv[i] = x.v[i]; sw[i] = x.w[i]; ap[i] = av[i] = aw[i] = 0.0f;
    tv[i] = D.v[i]; tw[i] = D.w[i]; bp[i] = bv[i] = bw[i] = tzero<T>();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) { sq[i] = x.q[i]; aq[i] = 0.0f; tq[i] = D.q[i]; bq[i] = tzero<T>(); }

#pragma unroll 1
  for (int stage = 0; stage < 4; ++stage) {
    const float wgt = (stage == 0 || stage == 3) ? 1.0f : 2.0f;
    const float cn = (stage == 2) ? h : hh;
    const T dcn = tmul((stage == 2) ? 1.0f : 0.5f, dh);      // d(c_i h) along this column
    float kv[3], kq[4], kw[3], J[6][5];
    Stage s;
    f_eval_net<true>(P, net, sv, sq, sw, c, kv, kq, kw, s, J);
    T dkv[3], dkq[4], dkw[3];
    f_tangent_t<T>(P, sv, sq, sw, s, J, tv, tq, tw, D.dc, dkv, dkq, dkw);
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      bp[i] = tfma(wgt, tv[i], bp[i]);
      bv[i] = tfma(wgt, dkv[i], bv[i]);
      bw[i] = tfma(wgt, dkw[i], bw[i]);
      tv[i] = tfma(kv[i], dcn, tfma(cn, dkv[i], D.v[i]));
      tw[i] = tfma(kw[i], dcn, tfma(cn, dkw[i], D.w[i]));
      ap[i] = fmaf(wgt, sv[i], ap[i]);
      av[i] = fmaf(wgt, kv[i], av[i]);
      aw[i] = fmaf(wgt, kw[i], aw[i]);
      sv[i] = fmaf(cn, kv[i], x.v[i]);
      sw[i] = fmaf(cn, kw[i], x.w[i]);
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      bq[i] = tfma(wgt, dkq[i], bq[i]);
      tq[i] = tfma(kq[i], dcn, tfma(cn, dkq[i], D.q[i]));
      aq[i] = fmaf(wgt, kq[i], aq[i]);
      sq[i] = fmaf(cn, kq[i], x.q[i]);
    }
  }
  const T dh6 = tmul(1.0f / 6.0f, dh);                       // d(h/6) along this column
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    D.p[i] = tfma(ap[i], dh6, tfma(h6, bp[i], D.p[i]));
    D.v[i] = tfma(av[i], dh6, tfma(h6, bv[i], D.v[i]));
    D.w[i] = tfma(aw[i], dh6, tfma(h6, bw[i], D.w[i]));
    x.p[i] = fmaf(h6, ap[i], x.p[i]);
    x.v[i] = fmaf(h6, av[i], x.v[i]);
    x.w[i] = fmaf(h6, aw[i], x.w[i]);
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    D.q[i] = tfma(aq[i], dh6, tfma(h6, bq[i], D.q[i]));
    x.q[i] = fmaf(h6, aq[i], x.q[i]);
  }
}

// q <- q/|q| with its tangent (base.py:443-444, 473-474)
template <class T>
AIRC_HD void normalise_q_lin(State& x, TangentT<T>& D) {
  const float n2 = (x.q[0] * x.q[0] + x.q[1] * x.q[1]) + (x.q[2] * x.q[2] + x.q[3] * x.q[3]);
  float n, rn;
  sqrt_rsqrt(n2, n, rn);
#pragma unroll
  for (int i = 0; i < 4; ++i) x.q[i] *= rn;
  const T d = tadd(tsum<T>(x.q[0], D.q[0], x.q[1], D.q[1]), tsum<T>(x.q[2], D.q[2], x.q[3], D.q[3]));
#pragma unroll
  for (int i = 0; i < 4; ++i) D.q[i] = tmul(rn, tfma(-x.q[i], d, D.q[i]));
}

template <class T, class Net>
AIRC_HD void state_update_lin(const DevConst& P, const Net& net, State& x, const Ctl& c, float dt,
                              int substeps, bool normalise, TangentT<T>& D) {
  const float h = (substeps == 1) ? dt : dt / (float)substeps;
  const T dh = (substeps == 1) ? D.ddt : tmul(1.0f / (float)substeps, D.ddt);
  for (int s = 0; s < substeps; ++s) rk4_step_lin<T>(P, net, x, c, h, D, dh);
  if (normalise) normalise_q_lin<T>(x, D);
}

}  // namespace airc

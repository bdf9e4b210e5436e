// airc_hess.cuh -- K4 (second-order jets): Hessian of the multiplier-weighted step, H = d2 (lambda^T Phi) / dz2 with
// z = (x[13], u[7], dt), and its gradient, at n shooting nodes.
//
// Replaces CasADi's second-order AD behind IPOPT's exact nlp_hess_l
// (/root/reference/src/aircraft/config.py:29; constraints control/base.py:275-286; the dt row and
// column belong to time='progress', control/base.py:276,306-307).
//
// 15 directions are active: v(3) q(4) w(3) aileron elevator rudder flaps dt.  p never enters f and
// thrust is inert (aircraft.py:320), so their rows/columns of H are exactly zero and are left to the
// caller's memset.  The 120 entries (i <= j) of the symmetric 15x15 block are computed by 64 threads
// per node: thread t owns one outer direction i and a PAIR of inner directions (j, j+1), packed
// into the two lanes of FFMA2 (airc_jet.cuh).
#pragma once
#include "airc_jet.cuh"
#include "airc_step.cuh"
#if defined(__CUDACC__)
#include "airc_kernels.cuh"
#endif

namespace airc {

constexpr int NHACT = 15;       // active directions
constexpr int NZ = 21;          // z = (x 13, u 7, dt)
constexpr int HESS_LANES = 64;  // sum_{i<15} ceil((15-i)/2)

// thread t of a node -> outer direction i, inner pair (j0, j1); j1 == NHACT means "no second entry"
AIRC_HD void hess_task(int t, int& i, int& j0, int& j1) {
  int acc = 0;
  for (i = 0; i < NHACT; ++i) {
    const int cnt = (NHACT - i + 1) >> 1;
    if (t < acc + cnt) {
      j0 = i + 2 * (t - acc);
      j1 = j0 + 1;
      return;
    }
    acc += cnt;
  }
  i = j0 = j1 = NHACT;
}

// active direction -> index in z
AIRC_HD int hess_zindex(int d) {
  return d < 13 ? 3 + d : (d == 13 ? 19 : 20);
}

struct HessSeed {
  int i, j0, j1;
  AIRC_HD Jet operator()(float val, int slot) const {
    return Jet{val, slot == i ? 1.0f : 0.0f, F2{slot == j0 ? 1.0f : 0.0f, slot == j1 ? 1.0f : 0.0f}, F2{0.0f, 0.0f}};
  }
};

// How the p rows are carried through the step: f never reads p, p+ = p + sum over substeps of
// h/6 (v1 + 2 v2 + 2 v3 + v4), so the Hessian kernel keeps only lambda_p . p+ (one jet), the
// Jacobian-by-jets kernel the three rows.
struct PContract {
  float l[3];
  Jet acc, tot;
  AIRC_HD void begin() { acc = jconst(0.0f); }
  AIRC_HD void stage(float wgt, const Jet sv[3]) { acc = jfma(wgt, jfma(l[2], sv[2], jfma(l[1], sv[1], l[0] * sv[0])), acc); }
  AIRC_HD void end(const Jet& h6) { tot = jmad(h6, acc, tot); }
};
struct PRows {
  Jet acc[3], tot[3];
  AIRC_HD void begin() { acc[0] = acc[1] = acc[2] = jconst(0.0f); }
  AIRC_HD void stage(float wgt, const Jet sv[3]) {
#pragma unroll
    for (int i = 0; i < 3; ++i) acc[i] = jfma(wgt, sv[i], acc[i]);
  }
  AIRC_HD void end(const Jet& h6) {
#pragma unroll
    for (int i = 0; i < 3; ++i) tot[i] = jmad(h6, acc[i], tot[i]);
  }
};

// ---- the network shared by the 64 threads of a node (short horizons) --------------------------------------
// The value-level sums over the 32 hidden units are the same in all 64 threads of a node (they differ in their
// directions only): C = c0 + Wo r, J = dC/dy (6 x 5) and T_k = d2 C_k / dy2 (six symmetric 5 x 5: 90 numbers),
// y = (aileron, elevator, qbar, alpha, beta).  Thread t takes hidden unit t mod 32 and one half of the 126 terms,
// the terms go through shared memory, thread t adds columns t and t + 64 over the 32 units, and every thread
// reads the 126 sums back.  A jet of C_k is then
//   v = C_k,  o = J_k y.o,  g = J_k y.g,  og = J_k y.og + (y.o^T T_k) y.g.
// This shortens the dependency chain of one f evaluation from ~3 000 to ~1 900 instructions, which is what a
// short horizon (one warp per scheduler) waits for.
constexpr int NJS_SUMS = 6 + 30 + 90, NJS_ROW = NJS_SUMS + 1, NJS_HALF = 6 + 30 + 30;

// the terms [0, NJS_HALF) (HALF 0) or [NJS_HALF, NJS_SUMS) (HALF 1) of one hidden unit; w = (a0, a[5], Wo[6])
template <int HALF>
AIRC_HD void net_unit_terms(const float w[12], const float y[5], float* out) {
  float h = w[0];
#pragma unroll
  for (int i = 0; i < 5; ++i) h = fmaf(w[1 + i], y[i], h);
  const float r = sigm2(h);
  const float r1 = -0.69314718056f * r * (1.0f - r);
  const float r2 = -0.69314718056f * (1.0f - 2.0f * r) * r1;
  float qq[15];
  {
    int e = 0;
#pragma unroll
    for (int i = 0; i < 5; ++i) {
      const float t = r2 * w[1 + i];
#pragma unroll
      for (int l = i; l < 5; ++l, ++e) qq[e] = t * w[1 + l];
    }
  }
  if (HALF == 0) {
#pragma unroll
    for (int k = 0; k < 6; ++k) {
      out[k] = w[6 + k] * r;
      const float wr = w[6 + k] * r1;
#pragma unroll
      for (int i = 0; i < 5; ++i) out[6 + 5 * k + i] = wr * w[1 + i];
    }
  }
#pragma unroll
  for (int k = (HALF ? 2 : 0); k < (HALF ? 6 : 2); ++k)
#pragma unroll
    for (int e = 0; e < 15; ++e) out[36 + 15 * k + e] = w[6 + k] * qq[e];
}

AIRC_HD void net_unit_weights(const DevConst& P, int j, float w[12]) {
  const HidPair& hp = P.net.hid[j >> 1];
  const int e = j & 1;
  w[0] = hp.a0[e];
  w[1] = hp.Au[0][e]; w[2] = hp.Au[1][e];
  w[3] = hp.Ax[0][e]; w[4] = hp.Ax[1][e]; w[5] = hp.Ax[2][e];
#pragma unroll
  for (int k = 0; k < 6; ++k) w[6 + k] = P.net.Wo[k][j];
}

struct NetSharedJet {
  float* prod;    // device: [32][NJS_ROW] in shared memory; host: unused
  float* sums;    // device: [NJS_SUMS] in shared memory; host: unused
  int t;          // thread of the node, 0..63
  float w[12];    // device: the weights of hidden unit t mod 32

  AIRC_HD void operator()(const DevConst& P, const Jet& da, const Jet& de, const Jet& qbar, const Jet& alpha,
                          const Jet& beta, Jet C[6]) const {
    const Jet* yin[5] = {&da, &de, &qbar, &alpha, &beta};
    const float y[5] = {da.v, de.v, qbar.v, alpha.v, beta.v};
    float S[NJS_SUMS];
#if defined(__CUDA_ARCH__)
    {
      float* row = prod + (t & 31) * NJS_ROW;
      if (t < 32) net_unit_terms<0>(w, y, row);
      else net_unit_terms<1>(w, y, row);
      __syncthreads();
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        const int m = t + 64 * c;
        if (m < NJS_SUMS) {
          float a = prod[m];
#pragma unroll 8
          for (int j = 1; j < 32; ++j) a += prod[j * NJS_ROW + m];
          sums[m] = a;
        }
      }
      __syncthreads();
#pragma unroll
      for (int m = 0; m < NJS_SUMS; ++m) S[m] = sums[m];
    }
#else
    {   // the same terms, added in the same order (units 0..31)
      float term[NJS_SUMS], ww[12];
      for (int j = 0; j < 32; ++j) {
        net_unit_weights(P, j, ww);
        net_unit_terms<0>(ww, y, term);
        net_unit_terms<1>(ww, y, term);
        for (int m = 0; m < NJS_SUMS; ++m) S[m] = j ? S[m] + term[m] : term[m];
      }
    }
#endif
#pragma unroll
    for (int k = 0; k < 6; ++k) {
      const float* Jk = S + 6 + 5 * k;
      const float* Tk = S + 36 + 15 * k;
      Jet r;
      r.v = P.net.c0[k] + S[k];
      r.o = Jk[0] * yin[0]->o;
      r.g = tmul(Jk[0], yin[0]->g);
      r.og = tmul(Jk[0], yin[0]->og);
#pragma unroll
      for (int i = 1; i < 5; ++i) {
        r.o = fmaf(Jk[i], yin[i]->o, r.o);
        r.g = tfma(Jk[i], yin[i]->g, r.g);
        r.og = tfma(Jk[i], yin[i]->og, r.og);
      }
#pragma unroll
      for (int l = 0; l < 5; ++l) {   // u_l = sum_i T_k[i][l] y_i.o
        float u = 0.0f;
#pragma unroll
        for (int i = 0; i < 5; ++i) {
          const int lo = i < l ? i : l, hi = i < l ? l : i;
          u = fmaf(Tk[lo * 5 - (lo * (lo - 1)) / 2 + (hi - lo)], yin[i]->o, u);
        }
        r.og = tfma(u, yin[l]->g, r.og);
      }
      C[k] = r;
    }
  }
};

// Phi(x,u,dt) on jets (base.py:408-480): v, q, w in place; the p rows through PA.
template <int MODEL, class PA, class NET = NetInlineJet>
AIRC_HD void state_update_jet(const DevConst& P, Jet v[3], Jet q[4], Jet w[3], const CtlJ& c, const Jet& h,
                              int substeps, bool normalise, PA& pa, const NET& net = NET()) {
  const Jet hh = 0.5f * h, h6 = (1.0f / 6.0f) * h;
  for (int sub = 0; sub < substeps; ++sub) {
    Jet sv[3], sq[4], sw[3], av[3], aq[4], aw[3];
    pa.begin();
#pragma unroll
    for (int i = 0; i < 3; ++i) { sv[i] = v[i]; sw[i] = w[i]; av[i] = jconst(0.0f); aw[i] = jconst(0.0f); }
#pragma unroll
    for (int i = 0; i < 4; ++i) { sq[i] = q[i]; aq[i] = jconst(0.0f); }
#pragma unroll 1
    for (int stage = 0; stage < 4; ++stage) {
      const float wgt = (stage == 0 || stage == 3) ? 1.0f : 2.0f;
      Jet kv[3], kq[4], kw[3];
      f_jet<MODEL, Jet, NET>(P, sv, sq, sw, c, kv, kq, kw, net);
      pa.stage(wgt, sv);
      const Jet cn = (stage == 2) ? h : hh;
#pragma unroll
      for (int i = 0; i < 3; ++i) {
        av[i] = jfma(wgt, kv[i], av[i]);
        aw[i] = jfma(wgt, kw[i], aw[i]);
        sv[i] = jmad(cn, kv[i], v[i]);
        sw[i] = jmad(cn, kw[i], w[i]);
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        aq[i] = jfma(wgt, kq[i], aq[i]);
        sq[i] = jmad(cn, kq[i], q[i]);
      }
    }
    pa.end(h6);
#pragma unroll
    for (int i = 0; i < 3; ++i) { v[i] = jmad(h6, av[i], v[i]); w[i] = jmad(h6, aw[i], w[i]); }
#pragma unroll
    for (int i = 0; i < 4; ++i) q[i] = jmad(h6, aq[i], q[i]);
  }
  if (normalise) {   // q <- q/|q| (base.py:443-444, 473-474)
    const Jet rn = jrsqrt((q[0] * q[0] + q[1] * q[1]) + (q[2] * q[2] + q[3] * q[3]));
#pragma unroll
    for (int i = 0; i < 4; ++i) q[i] = q[i] * rn;
  }
}

AIRC_HD void seed_state_jet(const State& x0, const Ctl& c0, float dt, int substeps, const HessSeed& seed,
                            Jet v[3], Jet q[4], Jet w[3], CtlJ& c, Jet& h) {
#pragma unroll
  for (int i = 0; i < 3; ++i) { v[i] = seed(x0.v[i], i); w[i] = seed(x0.w[i], 7 + i); }
#pragma unroll
  for (int i = 0; i < 4; ++i) q[i] = seed(x0.q[i], 3 + i);
  c = CtlJ{seed(c0.da, 10), seed(c0.de, 11), seed(c0.dr, 12), seed(c0.df, 13)};
  h = (1.0f / (float)substeps) * seed(dt, 14);
}

// lambda^T Phi(x,u,dt) on jets.  lam[13] in state order.  Returns the jet of the scalar.
template <int MODEL = 0, class NET = NetInlineJet>
AIRC_HD Jet lagrangian_step_jet(const DevConst& P, const State& x0, const Ctl& c0, float dt, int substeps,
                                bool normalise, const float lam[13], const HessSeed& seed, const NET& net = NET()) {
  Jet v[3], q[4], w[3], h;
  CtlJ c;
  seed_state_jet(x0, c0, dt, substeps, seed, v, q, w, c, h);
  PContract pa;
  pa.l[0] = lam[0]; pa.l[1] = lam[1]; pa.l[2] = lam[2];
  pa.tot = jconst((lam[0] * x0.p[0] + lam[1] * x0.p[1]) + lam[2] * x0.p[2]);
  state_update_jet<MODEL, PContract, NET>(P, v, q, w, c, h, substeps, normalise, pa, net);
  Jet L = pa.tot;
#pragma unroll
  for (int i = 0; i < 3; ++i) { L = jfma(lam[3 + i], v[i], L); L = jfma(lam[10 + i], w[i], L); }
#pragma unroll
  for (int i = 0; i < 4; ++i) L = jfma(lam[6 + i], q[i], L);
  return L;
}

// x+ with two first-order directions (j0, j1) alongside: out[r].v = x+[r], out[r].g = (d/dz_j0, d/dz_j1).
// The outer / second-order parts of the jets are seeded with zeros and never read, so the compiler drops
// them: this is plain two-column forward mode written once for every coefficient model.
template <int MODEL>
AIRC_HD void state_update_dual(const DevConst& P, const State& x0, const Ctl& c0, float dt, int substeps,
                               bool normalise, int j0, int j1, Jet out[13]) {
  Jet v[3], q[4], w[3], h;
  CtlJ c;
  const HessSeed seed{-1, j0, j1};
  seed_state_jet(x0, c0, dt, substeps, seed, v, q, w, c, h);
  PRows pa;
#pragma unroll
  for (int i = 0; i < 3; ++i) pa.tot[i] = jconst(x0.p[i]);
  state_update_jet<MODEL>(P, v, q, w, c, h, substeps, normalise, pa);
#pragma unroll
  for (int i = 0; i < 3; ++i) { out[i] = pa.tot[i]; out[3 + i] = v[i]; out[10 + i] = w[i]; }
#pragma unroll
  for (int i = 0; i < 4; ++i) out[6 + i] = q[i];
}

#if defined(__CUDACC__)
struct HessArgs {
  VecView x, u, lam;         // [13|7|13] per node
  const float* dt_per_node;  // optional
  float dt;
  long long n;
  int substeps, normalise;
  float* H;                  // SoA [441][n] / AoS [n][21][21]; zero-filled by the caller of the kernel
  float* grad;               // optional, SoA [21][n] / AoS [n][21]
  float* value;              // optional [n]: lambda^T Phi
};

// SHARED (network model, short horizons): one node per 64-thread block, the network through NetSharedJet.
template <bool SOA, int MODEL = 0, bool SHARED = false>
__global__ void __launch_bounds__(128)
hessian_kernel(const __grid_constant__ DevConst P, const HessArgs a) {
  const long long gt = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long node = gt / HESS_LANES;
  if (node >= a.n) return;   // SHARED: blockDim.x == HESS_LANES, so a block leaves as a whole
  const int t = (int)(gt % HESS_LANES);
  HessSeed seed;
  hess_task(t, seed.i, seed.j0, seed.j1);
  State s;
  load_state(a.x, node, s);
  const Ctl c = load_ctl(a.u.base + node * a.u.sb, a.u.sc);
  float lam[13];
#pragma unroll
  for (int r = 0; r < 13; ++r) lam[r] = __ldg(a.lam.base + node * a.lam.sb + r * a.lam.sc);
  const float dt = a.dt_per_node ? __ldg(a.dt_per_node + node) : a.dt;
  Jet L;
  if constexpr (SHARED) {
    __shared__ float prod[32 * NJS_ROW];
    __shared__ float sums[NJS_SUMS];
    NetSharedJet net;
    net.prod = prod; net.sums = sums; net.t = t;
    net_unit_weights(P, t & 31, net.w);
    L = lagrangian_step_jet<MODEL, NetSharedJet>(P, s, c, dt, a.substeps, a.normalise != 0, lam, seed, net);
  } else {
    L = lagrangian_step_jet<MODEL>(P, s, c, dt, a.substeps, a.normalise != 0, lam, seed);
  }
  const long long es = SOA ? a.n : 1;
  float* Hn = a.H + (SOA ? node : node * (NZ * NZ));
  const int zi = hess_zindex(seed.i), z0 = hess_zindex(seed.j0);
  Hn[(zi * NZ + z0) * es] = L.og.lo;
  Hn[(z0 * NZ + zi) * es] = L.og.lo;
  if (seed.j1 < NHACT) {
    const int z1 = hess_zindex(seed.j1);
    Hn[(zi * NZ + z1) * es] = L.og.hi;
    Hn[(z1 * NZ + zi) * es] = L.og.hi;
  }
  if (seed.j0 == seed.i) {   // the first task of every outer direction also owns the gradient entry
    if (a.grad) {
      float* g = a.grad + (SOA ? node : node * NZ);
      g[zi * es] = L.o;
      if (seed.i == 0) {     // inactive rows: d/dp = lambda_p (dPhi/dp = [I;0]), d/dthrust = 0
#pragma unroll
        for (int r = 0; r < 3; ++r) { g[r * es] = lam[r]; g[(16 + r) * es] = 0.0f; }
      }
    }
    if (a.value && seed.i == 0) a.value[node] = L.v;
  }
}

// Jacobian by two-column forward mode on jets, for the coefficient models K2's hand-derived tangents do not
// cover (poly / linear / default).  8 lanes per node, lane l owns directions (2l, 2l+1) of the 15 active
// ones (v q w | aileron elevator rudder flaps | dt); outputs and layouts exactly as linearize_coop_kernel.
template <bool SOA, int MODEL>
__global__ void __launch_bounds__(128)
linearize_jet_kernel(const __grid_constant__ DevConst P, const LinArgs a) {
  const long long gt = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long node = gt >> 3;
  if (node >= a.n) return;
  const int lane = (int)(gt & 7);
  State s;
  load_state(a.x, node, s);
  const Ctl c = load_ctl(a.u.base + node * a.u.sb, a.u.sc);
  const float dt = a.dt_per_node ? __ldg(a.dt_per_node + node) : a.dt;
  Jet out[13];
  state_update_dual<MODEL>(P, s, c, dt, a.substeps, a.normalise != 0, 2 * lane, 2 * lane + 1, out);
  const long long es = SOA ? a.n : 1;
  float* const Ab = a.A ? a.A + (SOA ? node : node * 169) : nullptr;
  float* const Bb = a.Bm ? a.Bm + (SOA ? node : node * 91) : nullptr;
#pragma unroll
  for (int half = 0; half < 2; ++half) {
    const int d = 2 * lane + half;
    if (d < 10) {
      if (Ab)
#pragma unroll
        for (int r = 0; r < 13; ++r) Ab[(r * 13 + 3 + d) * es] = half ? out[r].g.hi : out[r].g.lo;
    } else if (d < 14) {
      const int col = d == 13 ? 6 : d - 10;
      if (Bb)
#pragma unroll
        for (int r = 0; r < 13; ++r) Bb[(r * 7 + col) * es] = half ? out[r].g.hi : out[r].g.lo;
    } else if (d == 14) {
      if (a.Dt) {
        float* o = a.Dt + (SOA ? node : node * 13);
#pragma unroll
        for (int r = 0; r < 13; ++r) o[r * es] = out[r].g.lo;
      }
    }
  }
  if (lane == 7) {   // x+, the defect, and the constant columns dPhi/dp = [I;0], dPhi/d(thrust) = 0
    State n;
#pragma unroll
    for (int i = 0; i < 3; ++i) { n.p[i] = out[i].v; n.v[i] = out[3 + i].v; n.w[i] = out[10 + i].v; }
#pragma unroll
    for (int i = 0; i < 4; ++i) n.q[i] = out[6 + i].v;
    if (a.xn.base) store_state(a.xn, node, n);
    if (a.r.base) {
      State t;
      load_state(a.xt, node, t);
#pragma unroll
      for (int i = 0; i < 3; ++i) { t.p[i] -= n.p[i]; t.v[i] -= n.v[i]; t.w[i] -= n.w[i]; }
#pragma unroll
      for (int i = 0; i < 4; ++i) t.q[i] -= n.q[i];
      store_state(a.r, node, t);
    }
    if (Ab)
#pragma unroll
      for (int r = 0; r < 13; ++r)
#pragma unroll
        for (int cc = 0; cc < 3; ++cc) Ab[(r * 13 + cc) * es] = (r == cc) ? 1.0f : 0.0f;
    if (Bb)
#pragma unroll
      for (int r = 0; r < 13; ++r)
#pragma unroll
        for (int cc = 3; cc < 6; ++cc) Bb[(r * 7 + cc) * es] = 0.0f;
  }
}
#endif

}  // namespace airc

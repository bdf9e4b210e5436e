// airc_kernels.cuh -- sm_100a kernels of the aircraft hot path.
//
//   K0  state_derivative_kernel   f(x,u) at n nodes                       (base.py:385-406)
//   K0' state_update_kernel       Phi(x,u,dt) at n nodes                  (base.py:450-480)
//   K1  rollout_kernel            B trajectories x H steps, state in registers for the whole
//                                 horizon, optional trajectory write-out and fused cost (K3)
//   K2  linearize_coop_kernel     (x+, dPhi/dx, dPhi/du) at n nodes, forward mode, 8 lanes per node -- serves
//                                 substeps > 1; the one-substep case runs the tile kernel of airc_jac_kernel.cuh
//   K3' argmin_kernel             (min, argmin) over costs, warp shuffles + 64-bit atomicMin
//
// All global accesses go through explicit element strides so that the same kernels serve the
// native SoA layout (batch index fastest => every warp access is one coalesced 128-byte line)
// and the AoS convenience layout.
#pragma once
#include <cuda_runtime.h>

#include "airc_step.cuh"
#include "airc_coop.cuh"

namespace airc {

struct VecView {      // a batched small vector: element (b, c) at base[b*sb + c*sc]
  const float* base;
  long long sb, sc;
};
struct VecViewW {
  float* base;
  long long sb, sc;
};

AIRC_D void load_state(const VecView& X, long long b, State& s) {
  const float* p = X.base + b * X.sb;
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    s.p[i] = __ldg(p + i * X.sc);
    s.v[i] = __ldg(p + (3 + i) * X.sc);
    s.w[i] = __ldg(p + (10 + i) * X.sc);
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) s.q[i] = __ldg(p + (6 + i) * X.sc);
}

AIRC_D void store_state(const VecViewW& X, long long b, const State& s) {
  float* p = X.base + b * X.sb;
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    p[i * X.sc] = s.p[i];
    p[(3 + i) * X.sc] = s.v[i];
    p[(10 + i) * X.sc] = s.w[i];
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) p[(6 + i) * X.sc] = s.q[i];
}

AIRC_D Ctl load_ctl(const float* u, long long sc) {
  Ctl c;
  c.da = __ldg(u);
  c.de = __ldg(u + sc);
  c.dr = __ldg(u + 2 * sc);
  c.df = __ldg(u + 6 * sc);
  return c;
}

// ------------------------------------------------------------------------------------ K0
struct NodeArgs {
  VecView x, u;
  VecViewW out;
  const float* dt_per_node;
  float dt;
  long long n;
  int substeps;
  int normalise;
};

template <int MODEL>
__global__ void __launch_bounds__(128)
state_derivative_kernel(const __grid_constant__ DevConst P, const NodeArgs a) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= a.n) return;
  State s;
  load_state(a.x, i, s);
  const Ctl c = load_ctl(a.u.base + i * a.u.sb, a.u.sc);
  State d;
  Stage st;
  float J[6][5];
  f_eval<false, MODEL>(P, s.v, s.q, s.w, c, d.v, d.q, d.w, st, J);
#pragma unroll
  for (int k = 0; k < 3; ++k) d.p[k] = s.v[k];
  store_state(a.out, i, d);
}

template <int MODEL>
__global__ void __launch_bounds__(128)
state_update_kernel(const __grid_constant__ DevConst P, const NodeArgs a) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= a.n) return;
  State s;
  load_state(a.x, i, s);
  const Ctl c = load_ctl(a.u.base + i * a.u.sb, a.u.sc);
  const float dt = a.dt_per_node ? __ldg(a.dt_per_node + i) : a.dt;
  state_update<0, MODEL>(P, s, c, dt, a.substeps, a.normalise != 0, nullptr);
  store_state(a.out, i, s);
}

// ------------------------------------------------------------------------------------ K0''  (diagnostics)
// The side functions the reference exposes next to the dynamics and evaluates on every node for
// envelope constraints, losses, logging and plots (control/aircraft.py:44-59,140-142,
// main/control/control.py:63-64, plotting/plotting.py:349,363), in one pass per node:
//   out[0:3]  v_frd_rel        base.py:147-162        out[7:13]  coefficients   aircraft.py:255-307
//   out[3]    airspeed         base.py:164-170        out[13:16] forces_frd     aircraft.py:309-321
//   out[4]    alpha            base.py:172-177        out[16:19] moments_frd    aircraft.py:323-330, base.py:253-278
//   out[5]    beta             base.py:229-235        out[19:22] phi, theta, psi base.py:179-197
//   out[6]    qbar             base.py:237-241
constexpr int NDIAG = 22;
template <int MODEL>
__global__ void __launch_bounds__(128)
diagnostics_kernel(const __grid_constant__ DevConst P, const NodeArgs a) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= a.n) return;
  State s;
  load_state(a.x, i, s);
  const Ctl c = load_ctl(a.u.base + i * a.u.sb, a.u.sc);
  State d;
  Stage st;
  float J[6][5];
  f_eval<false, MODEL>(P, s.v, s.q, s.w, c, d.v, d.q, d.w, st, J);
  float o[NDIAG];
  const float vb2 = fmaf(st.vb[2], st.vb[2], fmaf(st.vb[1], st.vb[1], st.vb[0] * st.vb[0]));
  float V, rV;
  sqrt_rsqrt(vb2 + P.eps, V, rV);
#pragma unroll
  for (int k = 0; k < 3; ++k) { o[k] = st.vb[k]; o[13 + k] = st.F[k]; }
  o[3] = V; o[4] = st.alpha; o[5] = st.beta; o[6] = st.qbar;
#pragma unroll
  for (int k = 0; k < 6; ++k) o[7 + k] = st.C[k];
  o[16] = st.C[3] * (st.qbar * P.Sb) + (P.com[1] * st.F[2] - P.com[2] * st.F[1]);
  o[17] = st.C[4] * (st.qbar * P.Sc) + (P.com[2] * st.F[0] - P.com[0] * st.F[2]);
  o[18] = st.C[5] * (st.qbar * P.Sb) + (P.com[0] * st.F[1] - P.com[1] * st.F[0]);
  const float x = s.q[0], y = s.q[1], z = s.q[2], w = s.q[3];
  o[19] = atan2f(2.0f * (w * x + y * z), 1.0f - 2.0f * (x * x + y * y));
  o[20] = asinf(2.0f * (w * y - z * x));
  o[21] = atan2f(2.0f * (w * z + x * y), 1.0f - 2.0f * (y * y + z * z));
  float* p = a.out.base + i * a.out.sb;
#pragma unroll
  for (int k = 0; k < NDIAG; ++k) p[k * a.out.sc] = o[k];
}

// ------------------------------------------------------------------------------------ K1
struct CostConst {
  float goal0, goal1, vel_param, w_goal, w_du, neg_log2e_over_eps, w_height, w_speed_over_H,
      w_vx, w_vyz;
};

struct RolloutArgs {
  VecView x0;
  const float* U;            // element (b,k,c) at U[b*u_sb + k*u_sk + c*u_sc]
  long long u_sb, u_sk, u_sc;
  float* X;                  // optional trajectory, element (b,k,c) at X[b*X_sb + k*X_sk + c*X_sc]
  long long X_sb, X_sk, X_sc;
  VecViewW xf;               // optional final state
  float* cost;               // optional
  CostConst cc;
  float dt;
  long long B;
  int H;
  int substeps;
  int normalise;
};

// |v_b|^2 of the node, for the speed term of the loss (control.py:63-64 evaluates
// aircraft.v_frd_rel on every node of the trajectory).
AIRC_D float body_speed2(const DevConst& P, const State& s) {
  float M[9], n2;
  rot_matrix(s.q, M, n2);
  const float inv = rcp_nr(n2);
  float acc = 0.0f;
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    float a = M[i] * s.v[0];
    a = fmaf(M[3 + i], s.v[1], a);
    a = fmaf(M[6 + i], s.v[2], a);
    a = fmaf(a, inv, P.eps);
    acc = fmaf(a, a, acc);
  }
  return acc;
}

// 4 blocks of 128 threads per SM = 128 registers per thread, no spills; 5 or 6 blocks per SM
// (96 / 80 registers) spill and are 9-18 % slower (profiles/r1_variants.txt).
#ifndef AIRC_ROLLOUT_MINBLOCKS
#define AIRC_ROLLOUT_MINBLOCKS 4
#endif
#ifndef AIRC_ROLLOUT_BS
#define AIRC_ROLLOUT_BS 128
#endif
template <bool COST, bool TRAJ, int MODEL = 0>
__global__ void __launch_bounds__(AIRC_ROLLOUT_BS, AIRC_ROLLOUT_MINBLOCKS)
rollout_kernel(const __grid_constant__ DevConst P, const RolloutArgs a) {
  const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= a.B) return;
  State s;
  load_state(a.x0, b, s);
  const float z0 = s.p[2];
  const float* u = a.U + b * a.u_sb;
  float* Xb = TRAJ ? a.X + b * a.X_sb : nullptr;
  const VecViewW Xv{Xb, 0, a.X_sc};
  if (TRAJ) store_state(Xv, 0, s);

  float sum_du = 0.0f, sum_v2 = 0.0f;
  float up[7];
#pragma unroll 1
  for (int k = 0; k < a.H; ++k) {
    const float* uk = u + k * a.u_sk;
    const Ctl c = load_ctl(uk, a.u_sc);
    if (COST) {
      // smooth-L0 penalty on control increments, all 7 rows (control.py:22-23,47-48)
      float un[7];
      un[0] = c.da; un[1] = c.de; un[2] = c.dr; un[6] = c.df;
#pragma unroll
      for (int i = 3; i < 6; ++i) un[i] = __ldg(uk + i * a.u_sc);
      if (k > 0) {
#pragma unroll
        for (int i = 0; i < 7; ++i) {
          const float d = un[i] - up[i];
          sum_du += 1.0f - fast_ex2(d * d * a.cc.neg_log2e_over_eps);
        }
      }
#pragma unroll
      for (int i = 0; i < 7; ++i) up[i] = un[i];
      sum_v2 += body_speed2(P, s);
    }
    state_update<0, MODEL>(P, s, c, a.dt, a.substeps, a.normalise != 0, nullptr);
    if (TRAJ) {
      const VecViewW Xk{Xb + (long long)(k + 1) * a.X_sk, 0, a.X_sc};
      store_state(Xk, 0, s);
    }
  }
  if (a.xf.base) store_state(a.xf, b, s);
  if (COST) {
    const float ex = s.p[0] - a.cc.goal0, ey = s.p[1] - a.cc.goal1, ez = s.p[2] - z0;
    float J = a.cc.w_goal * (ex * ex + ey * ey);
    J = fmaf(a.cc.w_du, sum_du, J);
    J = fmaf(a.cc.w_height, ez * ez, J);
    J = fmaf(-a.cc.w_speed_over_H, sum_v2, J);
    J = fmaf(a.cc.vel_param * a.cc.w_vx, s.v[0], J);
    J = fmaf(a.cc.w_vyz, s.v[1] * s.v[1] + s.v[2] * s.v[2], J);
    a.cost[b] = J;
  }
}

// ------------------------------------------------------------------------------------ K1 (small batches)
// Latency case (BASELINE config 2: 4096 trajectories leave three quarters of the sub-partitions idle with one
// trajectory per thread): L = 8, 4 or 2 lanes share one trajectory, chosen so that the batch still runs as one
// warp per sub-partition.  Each lane evaluates 16/L of the hidden-unit pairs, the six coefficient sums are
// all-reduced by xor-shuffles (NetCoop), the rest of f is replicated (it is the latency chain, not the
// throughput, that matters here).  Lane 0 stores.
template <bool COST, bool TRAJ, int L>
__global__ void __launch_bounds__(128)
rollout_coop_kernel(const __grid_constant__ DevConst P, const RolloutArgs a) {
  __shared__ FusedNet snet;
  {
    const float* src = reinterpret_cast<const float*>(&P.net);
    float* dst = reinterpret_cast<float*>(&snet);
    for (int i = threadIdx.x; i < (int)(sizeof(FusedNet) / sizeof(float)); i += blockDim.x) dst[i] = src[i];
  }
  __syncthreads();
  const long long gt = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long braw = gt / L;
  const int lane = (int)(gt % L);
  const bool live = braw < a.B;
  const long long b = live ? braw : a.B - 1;          // whole warps stay alive for the shuffles
  const bool writer = live && lane == 0;
  State s;
  load_state(a.x0, b, s);
  const float z0 = s.p[2];
  const float* u = a.U + b * a.u_sb;
  float* Xb = TRAJ ? a.X + b * a.X_sb : nullptr;
  if (TRAJ && writer) store_state(VecViewW{Xb, 0, a.X_sc}, 0, s);
  const NetCoop<L> net{&snet, lane};
  float sum_du = 0.0f, sum_v2 = 0.0f;
  float up[7];
  // The controls of step k + 1 are fetched BEFORE step k is integrated: in this kernel nothing else hides the
  // global-memory latency (one or two warps per scheduler), and it would otherwise start every step.
  Ctl cn = {};
  float tn[3] = {0.0f, 0.0f, 0.0f};
  if (a.H > 0) {
    cn = load_ctl(u, a.u_sc);
    if (COST) {
#pragma unroll
      for (int i = 0; i < 3; ++i) tn[i] = __ldg(u + (3 + i) * a.u_sc);
    }
  }
#pragma unroll 1
  for (int k = 0; k < a.H; ++k) {
    const Ctl c = cn;
    const float t3[3] = {tn[0], tn[1], tn[2]};
    if (k + 1 < a.H) {
      const float* un1 = u + (long long)(k + 1) * a.u_sk;
      cn = load_ctl(un1, a.u_sc);
      if (COST) {
#pragma unroll
        for (int i = 0; i < 3; ++i) tn[i] = __ldg(un1 + (3 + i) * a.u_sc);
      }
    }
    if (COST) {
      float un[7];
      un[0] = c.da; un[1] = c.de; un[2] = c.dr; un[6] = c.df;
#pragma unroll
      for (int i = 3; i < 6; ++i) un[i] = t3[i - 3];
      if (k > 0) {
#pragma unroll
        for (int i = 0; i < 7; ++i) {
          const float d = un[i] - up[i];
          sum_du += 1.0f - fast_ex2(d * d * a.cc.neg_log2e_over_eps);
        }
      }
#pragma unroll
      for (int i = 0; i < 7; ++i) up[i] = un[i];
    }
    float sp2 = 0.0f;
    state_update_net(P, net, s, c, a.dt, a.substeps, a.normalise != 0, COST ? &sp2 : nullptr);
    if (COST) sum_v2 += sp2;
    if (TRAJ && writer) store_state(VecViewW{Xb + (long long)(k + 1) * a.X_sk, 0, a.X_sc}, 0, s);
  }
  if (!writer) return;
  if (a.xf.base) store_state(a.xf, b, s);
  if (COST) {
    const float ex = s.p[0] - a.cc.goal0, ey = s.p[1] - a.cc.goal1, ez = s.p[2] - z0;
    float J = a.cc.w_goal * (ex * ex + ey * ey);
    J = fmaf(a.cc.w_du, sum_du, J);
    J = fmaf(a.cc.w_height, ez * ez, J);
    J = fmaf(-a.cc.w_speed_over_H, sum_v2, J);
    J = fmaf(a.cc.vel_param * a.cc.w_vx, s.v[0], J);
    J = fmaf(a.cc.w_vyz, s.v[1] * s.v[1] + s.v[2] * s.v[2], J);
    a.cost[b] = J;
  }
}

// ------------------------------------------------------------------------------------ K2
struct LinArgs {
  VecView x, u;
  VecViewW xn;               // optional
  float* A;                  // optional; element (node, r, c) at A[node*A_sn + (r*13+c)*A_se]
  long long A_sn, A_se;
  float* Bm;                 // optional; element (node, r, c) at Bm[node*B_sn + (r*7+c)*B_se]
  long long B_sn, B_se;
  float* Dt;                 // optional dPhi/d(dt): SoA [13][n] / AoS [n][13]
  VecView xt;                // optional: the next node's state guess ...
  VecViewW r;                // ... and the multiple-shooting defect r = xt - Phi(x,u,dt) (control/base.py:275-279)
  const float* dt_per_node;
  float dt;
  long long n;
  int substeps;
  int normalise;
};

// ------------------------------------------------------------------------------------ K2 (cooperative)
// L = 8 lanes per node, 4 nodes per warp.  Column slots 2*lane, 2*lane+1 (one packed tangent):
//   slots 0..9   d/d(v, q, omega)      = A columns 3..12            (lanes 0..4)
//   slots 10..13 d/d(aileron, elevator, rudder, flaps) = B columns 0,1,2,6   (lanes 5, 6)
//   slot 14      d/d(dt)  (the step-size column of time='progress' mode, control/base.py:276)   (lane 7)
//   lane 7 also writes x+ and the constant columns
//   dPhi/dp = [I;0] and dPhi/d(thrust) = 0 (f does not read p; thrust is inert, aircraft.py:320).
// The surrogate and its 6x5 Jacobian are shared: two hidden-unit pairs per lane, all-reduced by
// shuffles (NetCoop).  The network lives in shared memory because each lane reads different pairs.
#ifndef AIRC_LIN_MINBLOCKS
#define AIRC_LIN_MINBLOCKS 2
#endif
// SOA is a template parameter so that the 26 column stores per lane use compile-time (AoS) or
// single-stride (SoA: element stride n) addressing instead of generic 64-bit stride arithmetic.
template <int L, bool SOA>
__global__ void __launch_bounds__(128, AIRC_LIN_MINBLOCKS)
linearize_coop_kernel(const __grid_constant__ DevConst P, const LinArgs a) {
  __shared__ FusedNet snet;
  {
    const float* src = reinterpret_cast<const float*>(&P.net);
    float* dst = reinterpret_cast<float*>(&snet);
    for (int i = threadIdx.x; i < (int)(sizeof(FusedNet) / sizeof(float)); i += blockDim.x) dst[i] = src[i];
  }
  __syncthreads();
  const long long gt = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long node_raw = gt / L;
  const int lane = (int)(gt % L);
  const bool live = node_raw < a.n;
  const long long node = live ? node_raw : a.n - 1;   // keep whole warps alive for the shuffles
  State s;
  load_state(a.x, node, s);
  const Ctl c = load_ctl(a.u.base + node * a.u.sb, a.u.sc);
  const float dt = a.dt_per_node ? __ldg(a.dt_per_node + node) : a.dt;

  const int s0 = 2 * lane, s1 = 2 * lane + 1;
  TangentT<F2> D;
  auto seed = [&](int slot) { return F2{s0 == slot ? 1.0f : 0.0f, s1 == slot ? 1.0f : 0.0f}; };
#pragma unroll
  for (int i = 0; i < 3; ++i) { D.p[i] = F2{0.0f, 0.0f}; D.v[i] = seed(i); D.w[i] = seed(7 + i); }
#pragma unroll
  for (int i = 0; i < 4; ++i) D.q[i] = seed(3 + i);
  D.dc.da = seed(10); D.dc.de = seed(11); D.dc.dr = seed(12); D.dc.df = seed(13);
  D.ddt = seed(14);

  const NetCoop<L> net{&snet, lane};
  state_update_lin<F2>(P, net, s, c, dt, a.substeps, a.normalise != 0, D);
  if (!live) return;

  F2 col[13];
#pragma unroll
  for (int i = 0; i < 3; ++i) { col[i] = D.p[i]; col[3 + i] = D.v[i]; col[10 + i] = D.w[i]; }
#pragma unroll
  for (int i = 0; i < 4; ++i) col[6 + i] = D.q[i];

  // element (node, e) of A / Bm: SoA base[e*n + node], AoS base[node*E + e]
  const long long es = SOA ? a.n : 1;                       // element stride
  float* const Ab = a.A ? a.A + (SOA ? node : node * 169) : nullptr;
  float* const Bb = a.Bm ? a.Bm + (SOA ? node : node * 91) : nullptr;
  if (lane < 5) {
    if (Ab) {
      float* o = Ab + (3 + s0) * es;
#pragma unroll
      for (int r = 0; r < 13; ++r, o += 13 * es) { o[0] = col[r].lo; o[es] = col[r].hi; }
    }
  } else if (lane < 7) {
    if (Bb) {
      const int c0 = (lane == 5) ? 0 : 2, c1 = (lane == 5) ? 1 : 6;
      float* o = Bb;
#pragma unroll
      for (int r = 0; r < 13; ++r, o += 7 * es) { o[c0 * es] = col[r].lo; o[c1 * es] = col[r].hi; }
    }
  } else {
    if (a.xn.base) store_state(a.xn, node, s);
    if (a.r.base) {
      State t;
      load_state(a.xt, node, t);
#pragma unroll
      for (int i = 0; i < 3; ++i) { t.p[i] -= s.p[i]; t.v[i] -= s.v[i]; t.w[i] -= s.w[i]; }
#pragma unroll
      for (int i = 0; i < 4; ++i) t.q[i] -= s.q[i];
      store_state(a.r, node, t);
    }
    if (a.Dt) {
      float* o = a.Dt + (SOA ? node : node * 13);
#pragma unroll
      for (int r = 0; r < 13; ++r) o[r * es] = col[r].lo;
    }
    if (Ab) {
      float* o = Ab;
#pragma unroll
      for (int r = 0; r < 13; ++r, o += 13 * es)
#pragma unroll
        for (int cc = 0; cc < 3; ++cc) o[cc * es] = (r == cc) ? 1.0f : 0.0f;
    }
    if (Bb) {
      float* o = Bb + 3 * es;
#pragma unroll
      for (int r = 0; r < 13; ++r, o += 7 * es)
#pragma unroll
        for (int cc = 0; cc < 3; ++cc) o[cc * es] = 0.0f;
    }
  }
}

// ------------------------------------------------------------------------------------ K3'
// Order-independent (min, argmin): key = (orderable bits of cost) << 32 | index-in-chunk.
// The 64-bit minimum is exact and ties resolve to the lower index; NaN costs sort last.
AIRC_D unsigned int float_orderable(float f) {
  unsigned int u = __float_as_uint(f);
  if (f != f) return 0xFFFFFFFFu;
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
AIRC_D float orderable_float(unsigned int k) {
  const unsigned int u = (k & 0x80000000u) ? (k & 0x7FFFFFFFu) : ~k;
  return __uint_as_float(u);
}

__global__ void __launch_bounds__(256)
argmin_kernel(const float* __restrict__ cost, long long B, unsigned long long* __restrict__ key) {
  unsigned long long best = 0xFFFFFFFFFFFFFFFFull;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < B;
       i += (long long)gridDim.x * blockDim.x) {
    const unsigned long long k =
        ((unsigned long long)float_orderable(__ldg(cost + i)) << 32) | (unsigned long long)(unsigned int)i;
    best = k < best ? k : best;
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
    const unsigned long long o = __shfl_xor_sync(0xFFFFFFFFu, best, off);
    best = o < best ? o : best;
  }
  __shared__ unsigned long long sm[8];
  const int wid = threadIdx.x >> 5, ln = threadIdx.x & 31;
  if (ln == 0) sm[wid] = best;
  __syncthreads();
  if (wid == 0) {
    best = (ln < (int)(blockDim.x >> 5)) ? sm[ln] : 0xFFFFFFFFFFFFFFFFull;
#pragma unroll
    for (int off = 4; off > 0; off >>= 1) {
      const unsigned long long o = __shfl_xor_sync(0xFFFFFFFFu, best, off);
      best = o < best ? o : best;
    }
    if (ln == 0) atomicMin(key, best);
  }
}

__global__ void argmin_finish_kernel(const unsigned long long* key, float* out_min, long long* out_idx) {
  const unsigned long long k = *key;
  if (out_min) *out_min = orderable_float((unsigned int)(k >> 32));
  if (out_idx) *out_idx = (long long)(k & 0xFFFFFFFFull);
}

}  // namespace airc

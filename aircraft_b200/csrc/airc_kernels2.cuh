// airc_kernels2.cuh -- rollout-family kernels on the packed two-trajectories-per-thread
// arithmetic of airc_packed.cuh.
//
//   rollout2_kernel<COST,TRAJ>  K1: B trajectories x H steps from explicit control sequences
//   sweep2_kernel<CKPT>         K1 + in-kernel control sampling + fused loss + warp-shuffle arg-min
//   sweep_winner2_kernel        parallel-in-time re-roll of the winner from the checkpoints
//
// Thread t of a block of NT threads owns trajectories base+t (lo half) and base+NT+t (hi half),
// base = blockIdx.x * 2 NT: both halves load/store fully coalesced 128-byte lines in the SoA
// layout and no alignment is required of B.  Out-of-range halves are clamped to B-1 (they
// recompute the last trajectory) and masked at the stores.
#pragma once
#include "airc_kernels.cuh"
#include "airc_packed.cuh"
#include "airc_sampler.cuh"

namespace airc {

AIRC_D void load_state2(const VecView& X, long long b0, long long b1, State2& s) {
  const float* p0 = X.base + b0 * X.sb;
  const float* p1 = X.base + b1 * X.sb;
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    s.p[i] = F2{__ldg(p0 + i * X.sc), __ldg(p1 + i * X.sc)};
    s.v[i] = F2{__ldg(p0 + (3 + i) * X.sc), __ldg(p1 + (3 + i) * X.sc)};
    s.w[i] = F2{__ldg(p0 + (10 + i) * X.sc), __ldg(p1 + (10 + i) * X.sc)};
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) s.q[i] = F2{__ldg(p0 + (6 + i) * X.sc), __ldg(p1 + (6 + i) * X.sc)};
}

// one half of a packed state to memory (element stride sc)
template <bool HI>
AIRC_D void store_half(float* p, long long sc, const State2& s) {
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    p[i * sc] = HI ? s.p[i].hi : s.p[i].lo;
    p[(3 + i) * sc] = HI ? s.v[i].hi : s.v[i].lo;
    p[(10 + i) * sc] = HI ? s.w[i].hi : s.w[i].lo;
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) p[(6 + i) * sc] = HI ? s.q[i].hi : s.q[i].lo;
}

AIRC_D float rollout_cost_finish(const CostConst& cc, float px, float py, float pz, float z0, float vx,
                                 float vy, float vz, float sum_du, float sum_v2) {
  const float ex = px - cc.goal0, ey = py - cc.goal1, ez = pz - z0;
  float J = cc.w_goal * (ex * ex + ey * ey);
  J = fmaf(cc.w_du, sum_du, J);
  J = fmaf(cc.w_height, ez * ez, J);
  J = fmaf(-cc.w_speed_over_H, sum_v2, J);
  J = fmaf(cc.vel_param * cc.w_vx, vx, J);
  J = fmaf(cc.w_vyz, vy * vy + vz * vz, J);
  return J;
}

#ifndef AIRC_R2_BS
#define AIRC_R2_BS 128
#endif
#ifndef AIRC_R2_MINBLOCKS
#define AIRC_R2_MINBLOCKS 3
#endif
// Register budget: <= 168 registers = 3 warps per SM sub-partition (the register file is per
// sub-partition: 16384 / (32 x regs)), i.e. 3 blocks of 128 threads per SM.  144 registers do NOT give
// 14 warps/SM, and 128 registers spill (profiles/r1_variants_packed.txt).  AIRC_R2_MAXNREG is the
// experiment knob that established this.
#ifdef AIRC_R2_MAXNREG
#define AIRC_R2_BOUNDS __maxnreg__(AIRC_R2_MAXNREG)
#else
#define AIRC_R2_BOUNDS __launch_bounds__(AIRC_R2_BS, AIRC_R2_MINBLOCKS)
#endif

// ------------------------------------------------------------------------------------ K1 (packed)
template <bool COST, bool TRAJ>
__global__ void AIRC_R2_BOUNDS
rollout2_kernel(const __grid_constant__ DevConst P, const __grid_constant__ UnitNet N, const RolloutArgs a) {
  const long long base = (long long)blockIdx.x * (2 * blockDim.x);
  const long long r0 = base + threadIdx.x, r1 = r0 + blockDim.x;
  if (r0 >= a.B) return;
  const bool live1 = r1 < a.B;
  const long long b0 = r0, b1 = live1 ? r1 : a.B - 1;
  State2 s;
  load_state2(a.x0, b0, b1, s);
  const F2 z0 = s.p[2];
  const float* u0 = a.U + b0 * a.u_sb;
  const float* u1 = a.U + b1 * a.u_sb;
  float* X0 = TRAJ ? a.X + b0 * a.X_sb : nullptr;
  float* X1 = TRAJ ? a.X + b1 * a.X_sb : nullptr;
  if (TRAJ) {
    store_half<false>(X0, a.X_sc, s);
    if (live1) store_half<true>(X1, a.X_sc, s);
  }
  F2 sum_du = bc2(0.0f), sum_v2 = bc2(0.0f);
  F2 up[7];
#pragma unroll 1
  for (int k = 0; k < a.H; ++k) {
    const float* k0 = u0 + k * a.u_sk;
    const float* k1 = u1 + k * a.u_sk;
    Ctl2 c;
    c.da = F2{__ldg(k0), __ldg(k1)};
    c.de = F2{__ldg(k0 + a.u_sc), __ldg(k1 + a.u_sc)};
    c.dr = F2{__ldg(k0 + 2 * a.u_sc), __ldg(k1 + 2 * a.u_sc)};
    c.df = F2{__ldg(k0 + 6 * a.u_sc), __ldg(k1 + 6 * a.u_sc)};
    if (COST) {
      // smooth-L0 penalty on control increments, all 7 rows (control.py:22-23,47-48)
      F2 un[7];
      un[0] = c.da; un[1] = c.de; un[2] = c.dr; un[6] = c.df;
#pragma unroll
      for (int i = 3; i < 6; ++i) un[i] = F2{__ldg(k0 + i * a.u_sc), __ldg(k1 + i * a.u_sc)};
      if (k > 0) {
#pragma unroll
        for (int i = 0; i < 7; ++i) sum_du = add2(sum_du, smooth_l0_2(un[i], up[i], a.cc.neg_log2e_over_eps));
      }
#pragma unroll
      for (int i = 0; i < 7; ++i) up[i] = un[i];
    }
    F2 speed2;
    state_update2(P, N, s, c, a.dt, a.substeps, a.normalise != 0, speed2);
    if (COST) sum_v2 = add2(sum_v2, speed2);
    if (TRAJ) {
      store_half<false>(X0 + (long long)(k + 1) * a.X_sk, a.X_sc, s);
      if (live1) store_half<true>(X1 + (long long)(k + 1) * a.X_sk, a.X_sc, s);
    }
  }
  if (a.xf.base) {
    store_half<false>(a.xf.base + b0 * a.xf.sb, a.xf.sc, s);
    if (live1) store_half<true>(a.xf.base + b1 * a.xf.sb, a.xf.sc, s);
  }
  if (COST) {
    a.cost[b0] = rollout_cost_finish(a.cc, s.p[0].lo, s.p[1].lo, s.p[2].lo, z0.lo, s.v[0].lo, s.v[1].lo,
                                     s.v[2].lo, sum_du.lo, sum_v2.lo);
    if (live1)
      a.cost[b1] = rollout_cost_finish(a.cc, s.p[0].hi, s.p[1].hi, s.p[2].hi, z0.hi, s.v[0].hi, s.v[1].hi,
                                       s.v[2].hi, sum_du.hi, sum_v2.hi);
  }
}

}  // namespace airc

// airc_sweep2.cuh -- the sampling sweep on the packed arithmetic, TIME-SLICED over the horizon.
//
// Why: a trajectory needs ~72 registers of live state per stage, so one SM holds at most 12 warps x
// 64 trajectories; a B200 holds W = 148 x 12 warps = 113 664 trajectories at once.  BASELINE config
// 4's shard (125 000) is 1.1 "waves": launched plainly, the 11 336 left-over trajectories run
// almost alone for a whole horizon and cost as much as the full first wave (2.96 ms instead of
// 2.2 ms x 1.1).  Here the horizon is cut into S segments and the (segment, group) tasks are taken by the blocks in
// segment-major order as warp slots free up; a group's state round-trips through
// a 23-float slice record in HBM between segments (traffic ~1 % of the kernel's time).  All warps
// stay busy until the very end: makespan ~ ceil(G S / W)/S waves instead of ceil(G / W).
// A task (s, g) waits on flag[g] >= s; its predecessor (s-1, g) has a lower block index and is normally resident
// or done.  The wait is BOUNDED: if the hand-over does not arrive, the warp rolls its group forward from x0 itself
// (and does not publish, so the slice record keeps a single writer chain): termination and results do not depend
// on the hardware's dispatch order, only the speed does.
#pragma once
#include "airc_kernels2.cuh"
#include "airc_sweep.cuh"

namespace airc {

struct Sweep2Args {
  const float* x0;            // [13]
  const float* Unom;          // [H][7]
  float* cost;                // [B] or null
  float* ckpt;                // [(NS-1)][17][B] or null (state + walk offsets before step j*C)
  float* slice;               // [23][B]: state 13, walk offsets 4, previous controls 4, sum_du, sum_v2
  int* flag;                  // [G] segments completed per group
  unsigned long long* key;    // arg-min key
  const float* thrust_du;
  CostConst cc;
  Sampler smp;
  float dt;
  long long B;
  int H, C, substeps, normalise;
  int G, S, L;                // groups, segments, steps per segment
  int GB;                     // blocks per segment = ceil(G / warps per block)
  int gs;                     // trajectories per group: 64, or 32 for small sweeps (hi halves idle:
                              // twice the warps for the same work when the machine is not full)
};

// resets the group flags and the arg-min key; sums the thrust rows' smooth-L0 term
__global__ void sweep2_prep_kernel(const float* __restrict__ Unom, int H, float neg_log2e_over_eps,
                                   unsigned long long* key, float* thrust_du, int* flag, int G) {
  for (int i = threadIdx.x; i < G; i += blockDim.x) flag[i] = 0;
  if (threadIdx.x < 32) {
    float acc = 0.0f;
    for (int k = 1 + (int)threadIdx.x; k < H; k += 32) {
#pragma unroll
      for (int c = 3; c < 6; ++c) {
        const float d = Unom[k * 7 + c] - Unom[(k - 1) * 7 + c];
        acc += 1.0f - fast_ex2(d * d * neg_log2e_over_eps);
      }
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) acc += __shfl_xor_sync(0xFFFFFFFFu, acc, off);
    if (threadIdx.x == 0) {
      *key = 0xFFFFFFFFFFFFFFFFull;
      *thrust_du = acc;
    }
  }
}

AIRC_D F2 ldcg2(const float* p, long long i0, long long i1) { return F2{__ldcg(p + i0), __ldcg(p + i1)}; }
AIRC_D void stcg2(float* p, long long i0, long long i1, const F2& v, bool live0, bool live1) {
  if (live0) __stcg(p + i0, v.lo);
  if (live1) __stcg(p + i1, v.hi);
}

#ifndef AIRC_S2_BS
#define AIRC_S2_BS 128
#endif
#ifndef AIRC_S2_MINBLOCKS
#define AIRC_S2_MINBLOCKS 3
#endif
// polls (128 ns sleep each + one L2 round trip) before a task stops waiting for its predecessor: ~50 ms, two
// orders of magnitude above a segment's duration.  0 in the A/B variant that exercises the fallback on every task.
#ifndef AIRC_S2_SPIN_LIMIT
#define AIRC_S2_SPIN_LIMIT 32768u
#endif

// every WARP is an independent worker (no block-level synchronisation); lanes carry trajectories gs*g + lane (lo) and, when gs = 64, gs*g + 32 + lane (hi)
template <bool CKPT>
__global__ void __launch_bounds__(AIRC_S2_BS, AIRC_S2_MINBLOCKS)
sweep2_kernel(const __grid_constant__ DevConst P, const __grid_constant__ UnitNet N, const Sweep2Args a) {
  // One (segment, 4 groups) task per block, taken from blockIdx in segment-major order: the hardware dispatches
  // blocks in index order as slots free up, which IS the dynamic queue, and a task's predecessor (seg-1, same
  // groups) has a lower index, so normally it is resident or done when the task starts.  CUDA does not
  // GUARANTEE that order, so nothing here depends on it for correctness or termination: the wait below is
  // bounded, and a warp whose predecessor has not delivered in time rolls its group forward from x0 itself
  // (same arithmetic, bit-identical state).
  // (Measured alternatives, profiles/r2_variants_sweep.txt: tasks handed out by an atomic ticket -- 5 % slower
  // although the instruction mix is the same; a software queue in a persistent loop -- the step loop's bounds
  // become warp-variant for the compiler and the weights leave the uniform registers, 30 %.)
  const int lane = threadIdx.x & 31;
  const int seg = (int)(blockIdx.x / (unsigned int)a.GB);
  const int g = (int)(blockIdx.x - (unsigned int)seg * (unsigned int)a.GB) * (int)(blockDim.x >> 5) + (int)(threadIdx.x >> 5);
  if (g >= a.G) return;
  {
    const long long r0 = (long long)a.gs * g + lane, r1 = r0 + 32;
    const bool live0 = r0 < a.B, live1 = a.gs == 64 && r1 < a.B;
    const long long b0 = live0 ? r0 : a.B - 1, b1 = live1 ? r1 : a.B - 1;
    const int kown = seg * a.L, k1 = min(a.H, kown + a.L);   // this task owns steps [kown, k1)
    int k0 = kown;
    bool from_x0 = seg == 0, fell_back = false;
    if (!from_x0) {
      // Bounded wait for the predecessor's hand-over.  Every condition is a warp VOTE, i.e. warp-uniform BY
      // CONSTRUCTION: with a spin loop that only lane 0 runs, ptxas treats everything after it as divergent and
      // stops using uniform registers for the network weights (LDC + R.F32 broadcasts instead of LDCU + UR.F32:
      // 30 % slower).
      unsigned int spins = 0;
      while (__any_sync(0xFFFFFFFFu, *(volatile int*)(a.flag + g) < seg && spins < AIRC_S2_SPIN_LIMIT)) {
        __nanosleep(128);
        ++spins;
      }
      if (__any_sync(0xFFFFFFFFu, *(volatile int*)(a.flag + g) < seg)) {
        from_x0 = fell_back = true;        // not delivered: do the earlier segments ourselves
        k0 = 0;
      }
    }

    State2 s;
    F2 d[4], up[4], sum_du, sum_v2;
    if (from_x0) {
      State s1;
      load_state_uniform(a.x0, s1);
#pragma unroll
      for (int i = 0; i < 3; ++i) { s.p[i] = bc2(s1.p[i]); s.v[i] = bc2(s1.v[i]); s.w[i] = bc2(s1.w[i]); }
#pragma unroll
      for (int i = 0; i < 4; ++i) { s.q[i] = bc2(s1.q[i]); d[i] = bc2(0.0f); up[i] = bc2(0.0f); }
      sum_du = bc2(0.0f);
      sum_v2 = bc2(0.0f);
    } else {
      __threadfence();
      const float* sl = a.slice;
#pragma unroll
      for (int i = 0; i < 3; ++i) {
        s.p[i] = ldcg2(sl + (long long)i * a.B, b0, b1);
        s.v[i] = ldcg2(sl + (long long)(3 + i) * a.B, b0, b1);
        s.w[i] = ldcg2(sl + (long long)(10 + i) * a.B, b0, b1);
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        s.q[i] = ldcg2(sl + (long long)(6 + i) * a.B, b0, b1);
        d[i] = ldcg2(sl + (long long)(13 + i) * a.B, b0, b1);
        up[i] = ldcg2(sl + (long long)(17 + i) * a.B, b0, b1);
      }
      sum_du = ldcg2(sl + 21LL * a.B, b0, b1);
      sum_v2 = ldcg2(sl + 22LL * a.B, b0, b1);
    }
    const unsigned int tk0 = traj_key(a.smp.seed_lo_mixed, a.smp.seed_hi, a.smp.index_offset + (unsigned long long)b0);
    const unsigned int tk1 = traj_key(a.smp.seed_lo_mixed, a.smp.seed_hi, a.smp.index_offset + (unsigned long long)b1);
#pragma unroll 1
    for (int k = k0; k < k1; ++k) {
      if (CKPT) {
        // (a task that fell back rewrites the earlier segments' checkpoints with the same bits: skipping them with
        //  `k >= kown` costs 4 % of the whole kernel -- the loop counter then leaves the uniform datapath)
        if (k > 0 && (k % a.C) == 0) {
          float* ck = a.ckpt + (long long)(k / a.C - 1) * 17 * a.B;
          if (live0) store_half<false>(ck + b0, a.B, s);
          if (live1) store_half<true>(ck + b1, a.B, s);
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            if (live0) ck[(13 + c) * a.B + b0] = d[c].lo;
            if (live1) ck[(13 + c) * a.B + b1] = d[c].hi;
          }
        }
      }
      float dl[4] = {d[0].lo, d[1].lo, d[2].lo, d[3].lo}, dh[4] = {d[0].hi, d[1].hi, d[2].hi, d[3].hi};
      float ul[4], uh[4];
      sample_step(a.smp, tk0, k, a.Unom + k * 7, dl, ul);
      sample_step(a.smp, tk1, k, a.Unom + k * 7, dh, uh);
      F2 un[4];
#pragma unroll
      for (int c = 0; c < 4; ++c) { d[c] = F2{dl[c], dh[c]}; un[c] = F2{ul[c], uh[c]}; }
      if (k > 0) {
#pragma unroll
        for (int c = 0; c < 4; ++c) sum_du = add2(sum_du, smooth_l0_2(un[c], up[c], a.cc.neg_log2e_over_eps));
      }
#pragma unroll
      for (int c = 0; c < 4; ++c) up[c] = un[c];
      const Ctl2 c{un[0], un[1], un[2], un[3]};
      F2 speed2;
      state_update2(P, N, s, c, a.dt, a.substeps, a.normalise != 0, speed2);
      sum_v2 = add2(sum_v2, speed2);
    }
    if (k1 < a.H) {
     if (!fell_back) {
      // hand the group over to whoever picks up (seg+1, g).  A task that fell back does NOT publish: the slice
      // record has one writer chain (seg 0 -> 1 -> ...), so a late predecessor can never overwrite a later
      // hand-over; the successors of a fallen-back task time out in turn and roll forward themselves.
      float* sl = a.slice;
#pragma unroll
      for (int i = 0; i < 3; ++i) {
        stcg2(sl + (long long)i * a.B, b0, b1, s.p[i], live0, live1);
        stcg2(sl + (long long)(3 + i) * a.B, b0, b1, s.v[i], live0, live1);
        stcg2(sl + (long long)(10 + i) * a.B, b0, b1, s.w[i], live0, live1);
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        stcg2(sl + (long long)(6 + i) * a.B, b0, b1, s.q[i], live0, live1);
        stcg2(sl + (long long)(13 + i) * a.B, b0, b1, d[i], live0, live1);
        stcg2(sl + (long long)(17 + i) * a.B, b0, b1, up[i], live0, live1);
      }
      stcg2(sl + 21LL * a.B, b0, b1, sum_du, live0, live1);
      stcg2(sl + 22LL * a.B, b0, b1, sum_v2, live0, live1);
      __threadfence();
      __syncwarp();
      if (lane == 0) *(volatile int*)(a.flag + g) = seg + 1;
     }
    } else {
      const float z0 = __ldg(a.x0 + 2), tdu = __ldg(a.thrust_du);
      const float J0 = rollout_cost_finish(a.cc, s.p[0].lo, s.p[1].lo, s.p[2].lo, z0, s.v[0].lo, s.v[1].lo,
                                           s.v[2].lo, sum_du.lo + tdu, sum_v2.lo);
      const float J1 = rollout_cost_finish(a.cc, s.p[0].hi, s.p[1].hi, s.p[2].hi, z0, s.v[0].hi, s.v[1].hi,
                                           s.v[2].hi, sum_du.hi + tdu, sum_v2.hi);
      if (a.cost) {
        if (live0) a.cost[b0] = J0;
        if (live1) a.cost[b1] = J1;
      }
      // K3: (min, arg-min) by warp shuffles, one 64-bit atomicMin per warp
      const unsigned long long none = 0xFFFFFFFFFFFFFFFFull;
      const unsigned long long key0 = live0 ? (((unsigned long long)float_orderable(J0) << 32) | (unsigned int)b0) : none;
      const unsigned long long key1 = live1 ? (((unsigned long long)float_orderable(J1) << 32) | (unsigned int)b1) : none;
      unsigned long long best = key0 < key1 ? key0 : key1;
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) {
        const unsigned long long o = __shfl_xor_sync(0xFFFFFFFFu, best, off);
        best = o < best ? o : best;
      }
      if (lane == 0) atomicMin(a.key, best);
    }
  }
}

// Thread j re-rolls steps [j*C, (j+1)*C) of the winner from checkpoint j with the SAME packed
// arithmetic (both halves carry the winner), so the record is bit-identical to what the sweep saw.
__global__ void sweep_winner2_kernel(const __grid_constant__ DevConst P, const __grid_constant__ UnitNet N,
                                     const WinnerArgs a) {
  const unsigned long long key = *a.key;
  const long long b = (long long)(key & 0xFFFFFFFFull);
  float* traj = a.record + AIRC_SWEEP_RECORD_HEADER;
  float* ctrl = traj + 13 * (a.H + 1);
  const int j = (int)(blockIdx.x * blockDim.x + threadIdx.x);
  if (j == 0) {
    a.record[0] = orderable_float((unsigned int)(key >> 32));
    a.record[1] = __uint_as_float((unsigned int)(key >> 32));
    const unsigned long long g = a.smp.index_offset + (unsigned long long)b;
    a.record[2] = __uint_as_float((unsigned int)(g & 0xFFFFFFFFull));
    a.record[3] = __uint_as_float((unsigned int)(g >> 32));
  }
  const int k0 = j * a.C;
  if (k0 >= a.H && j != 0) return;
  if (key == 0xFFFFFFFFFFFFFFFFull || b >= a.B) return;   // empty sweep: header only
  State2 s;
  float d[4] = {0.f, 0.f, 0.f, 0.f};
  if (j == 0) {
    State s1;
    load_state_uniform(a.x0, s1);
#pragma unroll
    for (int i = 0; i < 3; ++i) { s.p[i] = bc2(s1.p[i]); s.v[i] = bc2(s1.v[i]); s.w[i] = bc2(s1.w[i]); }
#pragma unroll
    for (int i = 0; i < 4; ++i) s.q[i] = bc2(s1.q[i]);
    store_half<false>(traj, 1, s);
  } else {
    const float* ck = a.ckpt + (long long)(j - 1) * 17 * a.B + b;
    const VecView cv{ck, 0, a.B};
    load_state2(cv, 0, 0, s);
#pragma unroll
    for (int c = 0; c < 4; ++c) d[c] = ck[(13 + c) * a.B];
  }
  const unsigned int tkey = traj_key(a.smp.seed_lo_mixed, a.smp.seed_hi, a.smp.index_offset + (unsigned long long)b);
  const int k1 = min(a.H, k0 + a.C);
#pragma unroll 1
  for (int k = k0; k < k1; ++k) {
    float un[4];
    sample_step(a.smp, tkey, k, a.Unom + k * 7, d, un);
    float* uo = ctrl + k * 7;
    uo[0] = un[0]; uo[1] = un[1]; uo[2] = un[2]; uo[6] = un[3];
#pragma unroll
    for (int c = 3; c < 6; ++c) uo[c] = a.Unom[k * 7 + c];
    const Ctl2 c{bc2(un[0]), bc2(un[1]), bc2(un[2]), bc2(un[3])};
    F2 speed2;
    state_update2(P, N, s, c, a.dt, a.substeps, a.normalise != 0, speed2);
    store_half<false>(traj + 13 * (k + 1), 1, s);
  }
}

}  // namespace airc

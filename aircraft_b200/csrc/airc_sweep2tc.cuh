// airc_sweep2tc.cuh -- the time-sliced sampling sweep (airc_sweep2.cuh) with the surrogate on the tensor cores
// (airc_tc.cuh).  Same tasks, same hand-over protocol, same arithmetic around the network; what changes is the
// shape of a block: ONE block of 384 threads per SM = three independent GROUPS of 128 threads (4 warps, 256
// trajectories), because tensor memory is allocated per block (all 512 columns here, 160 per group) and an
// M = 128 MMA is fed by four warps, one lane quarter each.  A group takes the task a 128-thread block of the
// FFMA2 kernel would take (virtual block = 3 blockIdx + group, segment-major), synchronises only with itself
// (named barrier + mbarrier per group) and never leaves early: warps without trajectories run on clamped
// indices with their stores masked, so that the group's barriers always see 128 threads.
#pragma once
#include "airc_sweep2.cuh"
#include "airc_tc.cuh"

namespace airc {

#define AIRC_S2TC_BS 384

struct alignas(16) S2tcShared {
  TcNetImage W;
  unsigned long long mbar[3];
  uint32_t tmem_base;
  uint32_t pad_;
  float2 acc[13 * AIRC_S2TC_BS]; // RK4 weighted sums, [component][thread]
};

template <bool CKPT>
__global__ void __launch_bounds__(AIRC_S2TC_BS, 1)
sweep2tc_kernel(const __grid_constant__ DevConst P, const __grid_constant__ TcNetImage W, const Sweep2Args a) {
  __shared__ S2tcShared sh;
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xFFFFFFFFu, tid >> 5, 0);          // warp-uniform by construction
  const int grp = warp >> 2, wig = warp & 3;                       // group, warp in group (= tensor-memory lane quarter)

  // ---- one-time set-up: weights to shared memory, tensor memory, one mbarrier per group
  {
    const float4* src = reinterpret_cast<const float4*>(&W);
    float4* dst = reinterpret_cast<float4*>(&sh.W);
    for (int i = tid; i < (int)(sizeof(TcNetImage) / 16); i += AIRC_S2TC_BS) dst[i] = src[i];
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" :: "r"(tc::smem_u32(&sh.tmem_base)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  if (tid == 0) {
#pragma unroll
    for (int i = 0; i < 3; ++i) tc::mbar_init(tc::smem_u32(&sh.mbar[i]), 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");      // the weights are read by the tensor core (async proxy)
  tc::fence_before();
  __syncthreads();
  tc::fence_after();

  tc::Group G;
  G.tm_group = sh.tmem_base + (uint32_t)(grp * tc::COLS_PER_GROUP);
  G.tm_lanes = G.tm_group + ((uint32_t)(wig * 32) << 16);
  G.mbar = tc::smem_u32(&sh.mbar[grp]);
  G.parity = 0;
  G.d1h = tc::smem_desc(tc::smem_u32(sh.W.B1h), tc::TC_LBO1, 128); G.d1l = tc::smem_desc(tc::smem_u32(sh.W.B1l), tc::TC_LBO1, 128);
  G.d2h = tc::smem_desc(tc::smem_u32(sh.W.B2h), tc::TC_LBO2, 128); G.d2l = tc::smem_desc(tc::smem_u32(sh.W.B2l), tc::TC_LBO2, 128);
  G.bar_id = 1 + grp;
  G.issuer = wig == 0;
  SurTc sur{G, sh.W.c0};

  // ---- the group's task (segment-major over virtual blocks of 4 warps, as in sweep2_kernel)
  const long long vb = (long long)blockIdx.x * 3 + grp;
  const int seg = (int)(vb / a.GB);
  if (seg < a.S) {                                                  // group-uniform
    const int g_raw = (int)(vb - (long long)seg * a.GB) * 4 + wig;
    const bool warp_live = g_raw < a.G;
    const int g = warp_live ? g_raw : a.G - 1;
    const long long r0 = 64LL * g + lane, r1 = r0 + 32;
    const bool live0 = warp_live && r0 < a.B, live1 = warp_live && r1 < a.B;
    const long long b0 = live0 ? r0 : a.B - 1, b1 = live1 ? r1 : a.B - 1;
    const int kown = seg * a.L, k1 = min(a.H, kown + a.L);
    int k0 = kown;
    bool from_x0 = seg == 0, fell_back = false;
    if (!from_x0) {
      unsigned int spins = 0;
      while (__any_sync(0xFFFFFFFFu, *(volatile int*)(a.flag + g) < seg && spins < AIRC_S2_SPIN_LIMIT)) {
        __nanosleep(128);
        ++spins;
      }
      // the decision must be the same for the whole group (its barriers need all four warps in every step)
      const unsigned int late = __any_sync(0xFFFFFFFFu, *(volatile int*)(a.flag + g) < seg) ? 1u : 0u;
      unsigned int any_late;
      asm volatile("{\n\t.reg .pred p, q;\n\tsetp.ne.u32 q, %1, 0;\n\tbar.red.or.pred p, %2, %3, q;\n\tselp.u32 %0, 1, 0, p;\n\t}\n"
                   : "=r"(any_late) : "r"(late), "r"(G.bar_id), "r"(tc::GROUP_THREADS) : "memory");
      if (any_late) {
        from_x0 = fell_back = true;
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
      state_update2_smem(P, sur, s, c, a.dt, a.substeps, a.normalise != 0, speed2, sh.acc + tid, AIRC_S2TC_BS);
      sum_v2 = add2(sum_v2, speed2);
    }
    if (k1 < a.H) {
      if (!fell_back && warp_live) {
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
      const unsigned long long none = 0xFFFFFFFFFFFFFFFFull;
      const unsigned long long key0 = live0 ? (((unsigned long long)float_orderable(J0) << 32) | (unsigned int)b0) : none;
      const unsigned long long key1 = live1 ? (((unsigned long long)float_orderable(J1) << 32) | (unsigned int)b1) : none;
      unsigned long long best = key0 < key1 ? key0 : key1;
#pragma unroll
      for (int off = 16; off > 0; off >>= 1) {
        const unsigned long long o = __shfl_xor_sync(0xFFFFFFFFu, best, off);
        best = o < best ? o : best;
      }
      if (lane == 0 && best != none) atomicMin(a.key, best);
    }
  }

  // ---- teardown: every MMA of every group has completed (each group waited for its last commit)
  tc::fence_before();
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" :: "r"(sh.tmem_base) : "memory");
}

}  // namespace airc

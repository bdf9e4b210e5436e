// airc_sweep.cuh -- sampling-MPC sweep (BASELINE config 4) without a control tensor.
//
//   sweep_prep_kernel    resets the arg-min key, sums the smooth-L0 term of the (unsampled)
//                        thrust rows of the nominal sequence once for all trajectories
//   sweep_kernel         one trajectory per thread: controls sampled from a counter-based
//                        generator, fused RK4 rollout + loss (main/control/control.py:35-70),
//                        sparse state checkpoints, warp-shuffle (min,arg-min) + one atomicMin
//                        per warp
//   sweep_winner_kernel  one thread per checkpoint segment re-rolls the winning trajectory in
//                        parallel-in-time (C steps of latency instead of H) and writes the record
//   select_record_kernel best of n records (after the all-gather across ranks)
//   sweep_controls_kernel materialises sampled control sequences (tests / inspection)
//
// The generator is specified in include/aircraft_b200.h (AircSweepParams); oracle/cpu_ref.cpp
// restates it independently.
#pragma once
#include "airc_kernels.cuh"
#include "airc_sampler.cuh"

namespace airc {

struct SweepArgs {
  const float* x0;            // [13]
  const float* Unom;          // [H][7]
  float* cost;                // [B] or null
  float* ckpt;                // [(S-1)][17][B] or null: state (13) + walk offsets (4) before step j*C
  unsigned long long* key;    // arg-min key (orderable(cost) << 32 | local index)
  const float* thrust_du;     // scalar written by sweep_prep_kernel
  CostConst cc;
  Sampler smp;
  float dt;
  long long B;
  int H, C, substeps, normalise;
};

__global__ void sweep_prep_kernel(const float* __restrict__ Unom, int H, float neg_log2e_over_eps,
                                  unsigned long long* key, float* thrust_du) {
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

AIRC_D void load_state_uniform(const float* x, State& s) {
#pragma unroll
  for (int i = 0; i < 3; ++i) { s.p[i] = __ldg(x + i); s.v[i] = __ldg(x + 3 + i); s.w[i] = __ldg(x + 10 + i); }
#pragma unroll
  for (int i = 0; i < 4; ++i) s.q[i] = __ldg(x + 6 + i);
}

template <bool CKPT, int MODEL = 0>
__global__ void __launch_bounds__(AIRC_ROLLOUT_BS, AIRC_ROLLOUT_MINBLOCKS)
sweep_kernel(const __grid_constant__ DevConst P, const SweepArgs a) {
  const long long braw = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const bool live = braw < a.B;
  const long long b = live ? braw : a.B - 1;     // whole warps stay alive for the shuffles
  State s;
  load_state_uniform(a.x0, s);
  const float z0 = s.p[2];
  const unsigned int tkey = traj_key(a.smp.seed_lo_mixed, a.smp.seed_hi, a.smp.index_offset + (unsigned long long)b);
  float d[4] = {0.f, 0.f, 0.f, 0.f}, up[4] = {0.f, 0.f, 0.f, 0.f};
  float sum_du = 0.0f, sum_v2 = 0.0f;
  int until_ckpt = a.C;
  float* ck = CKPT ? a.ckpt + b : nullptr;
#pragma unroll 1
  for (int k = 0; k < a.H; ++k) {
    if (CKPT) {
      if (until_ckpt == 0) {
        until_ckpt = a.C;
        const VecViewW cv{ck, 0, a.B};
        store_state(cv, 0, s);
#pragma unroll
        for (int c = 0; c < 4; ++c) ck[(13 + c) * a.B] = d[c];
        ck += 17 * a.B;
      }
      --until_ckpt;
    }
    float un[4];
    sample_step(a.smp, tkey, k, a.Unom + k * 7, d, un);
    if (k > 0) {
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const float e = un[c] - up[c];
        sum_du += 1.0f - fast_ex2(e * e * a.cc.neg_log2e_over_eps);
      }
    }
#pragma unroll
    for (int c = 0; c < 4; ++c) up[c] = un[c];
    sum_v2 += body_speed2(P, s);
    const Ctl c{un[0], un[1], un[2], un[3]};
    state_update<0, MODEL>(P, s, c, a.dt, a.substeps, a.normalise != 0, nullptr);
  }
  const float ex = s.p[0] - a.cc.goal0, ey = s.p[1] - a.cc.goal1, ez = s.p[2] - z0;
  float J = a.cc.w_goal * (ex * ex + ey * ey);
  J = fmaf(a.cc.w_du, sum_du + __ldg(a.thrust_du), J);
  J = fmaf(a.cc.w_height, ez * ez, J);
  J = fmaf(-a.cc.w_speed_over_H, sum_v2, J);
  J = fmaf(a.cc.vel_param * a.cc.w_vx, s.v[0], J);
  J = fmaf(a.cc.w_vyz, s.v[1] * s.v[1] + s.v[2] * s.v[2], J);
  if (live && a.cost) a.cost[b] = J;
  // K3: (min, arg-min) by warp shuffles, one 64-bit atomicMin per warp
  unsigned long long best = live
      ? (((unsigned long long)float_orderable(J) << 32) | (unsigned long long)(unsigned int)b)
      : 0xFFFFFFFFFFFFFFFFull;
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
    const unsigned long long o = __shfl_xor_sync(0xFFFFFFFFu, best, off);
    best = o < best ? o : best;
  }
  if ((threadIdx.x & 31) == 0) atomicMin(a.key, best);
}

struct WinnerArgs {
  const float* x0;
  const float* Unom;
  const float* ckpt;          // as in SweepArgs, or null (then one thread re-rolls all H steps)
  const unsigned long long* key;
  Sampler smp;
  float* record;              // [4 + 13(H+1) + 7H]
  float dt;
  long long B;
  int H, C, substeps, normalise;
};

// Thread j owns steps [j*C, min(H,(j+1)*C)): it starts from checkpoint j (x0 for j = 0) and
// repeats the sweep kernel's arithmetic, so the record is the trajectory the cost was computed on.
template <int MODEL = 0>
__global__ void sweep_winner_kernel(const __grid_constant__ DevConst P, const WinnerArgs a) {
  const unsigned long long key = *a.key;
  const long long b = (long long)(key & 0xFFFFFFFFull);
  float* traj = a.record + AIRC_SWEEP_RECORD_HEADER;
  float* ctrl = traj + 13 * (a.H + 1);
  const int j = (int)(blockIdx.x * blockDim.x + threadIdx.x);
  if (j == 0) {
    a.record[0] = orderable_float((unsigned int)(key >> 32));
    a.record[1] = __uint_as_float((unsigned int)(key >> 32));
    const long long g = (long long)(a.smp.index_offset + (unsigned long long)b);
    a.record[2] = __uint_as_float((unsigned int)((unsigned long long)g & 0xFFFFFFFFull));
    a.record[3] = __uint_as_float((unsigned int)((unsigned long long)g >> 32));
  }
  const int k0 = j * a.C;
  if (k0 >= a.H && !(j == 0)) return;
  if (key == 0xFFFFFFFFFFFFFFFFull || b >= a.B) return;   // empty sweep: header only
  State s;
  float d[4] = {0.f, 0.f, 0.f, 0.f};
  if (j == 0) {
    load_state_uniform(a.x0, s);
    const VecViewW t0{traj, 0, 1};
    store_state(t0, 0, s);
  } else {
    const float* ck = a.ckpt + (long long)(j - 1) * 17 * a.B + b;
    const VecView cv{ck, 0, a.B};
    load_state(cv, 0, s);
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
    const Ctl c{un[0], un[1], un[2], un[3]};
    state_update<0, MODEL>(P, s, c, a.dt, a.substeps, a.normalise != 0, nullptr);
    const VecViewW tk{traj + 13 * (k + 1), 0, 1};
    store_state(tk, 0, s);
  }
}

// records[r*stride ...]: pick the lowest (orderable cost, global index); copy it to out.
__global__ void select_record_kernel(const float* __restrict__ records, int n, long long stride,
                                     float* __restrict__ out) {
  int best = 0;
  unsigned int bc = 0xFFFFFFFFu;
  unsigned long long bi = 0xFFFFFFFFFFFFFFFFull;
  for (int r = 0; r < n; ++r) {
    const float* rec = records + (long long)r * stride;
    const unsigned int c = __float_as_uint(rec[1]);
    const unsigned long long i =
        (unsigned long long)__float_as_uint(rec[2]) | ((unsigned long long)__float_as_uint(rec[3]) << 32);
    if (c < bc || (c == bc && i < bi)) { bc = c; bi = i; best = r; }
  }
  const float* src = records + (long long)best * stride;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < stride;
       i += (long long)gridDim.x * blockDim.x)
    out[i] = src[i];
}

__global__ void sweep_controls_kernel(const float* __restrict__ Unom, int H, const Sampler smp,
                                      long long first, long long count, int soa, float* __restrict__ U) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count) return;
  const unsigned int tkey = traj_key(smp.seed_lo_mixed, smp.seed_hi, smp.index_offset + (unsigned long long)(first + i));
  float d[4] = {0.f, 0.f, 0.f, 0.f};
  for (int k = 0; k < H; ++k) {
    float un[4];
    sample_step(smp, tkey, k, Unom + k * 7, d, un);
    float row[7] = {un[0], un[1], un[2], Unom[k * 7 + 3], Unom[k * 7 + 4], Unom[k * 7 + 5], un[3]};
#pragma unroll
    for (int c = 0; c < 7; ++c) {
      if (soa) U[((long long)k * 7 + c) * count + i] = row[c];
      else U[(i * H + k) * 7 + c] = row[c];
    }
  }
}

}  // namespace airc

// airc_sampler.cuh -- the counter-based control sampler of the sampling-MPC sweep
// (specification: include/aircraft_b200.h, AircSweepParams).  AIRC_HD, so that
// oracle/host_emul.cpp can compile this very code for the CPU-only tests; the independent
// restatement lives in oracle/cpu_ref.cpp.
#pragma once
#include "airc_math.cuh"

namespace airc {

AIRC_HD unsigned int mix32(unsigned int x) {
  x ^= x >> 16;
  x *= 0x7feb352dU;
  x ^= x >> 15;
  x *= 0x846ca68bU;
  x ^= x >> 16;
  return x;
}

AIRC_HD unsigned int traj_key(unsigned int seed_lo_mixed, unsigned int seed_hi, unsigned long long g) {
  return mix32((unsigned int)g + seed_lo_mixed) ^ seed_hi;
}

AIRC_HD float byte_sum_centered(unsigned int w) {
#if defined(__CUDA_ARCH__)
  return (float)((int)__dp4a(w, 0x01010101U, 0U) - 510);
#else
  return (float)((int)((w & 255U) + ((w >> 8) & 255U) + ((w >> 16) & 255U) + (w >> 24)) - 510);
#endif
}

struct Sampler {
  float s0[4], s[4], lo[4], hi[4];     // s0, s already divided by the generator's std 147.80053
  unsigned int seed_lo_mixed, seed_hi;
  unsigned long long index_offset;
};

AIRC_HD int live_chan(int c) { return c < 3 ? c : 6; }   // aileron, elevator, rudder, flaps

// advance the random-walk offsets d to step k and return the four live controls of that step
AIRC_HD void sample_step(const Sampler& S, unsigned int tkey, int k, const float* unom_k, float d[4],
                         float un[4]) {
  const unsigned int cnt = tkey + (unsigned int)k * 0x9E3779B9U;
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    const float n = byte_sum_centered(mix32(cnt + (unsigned int)c * 0x7F4A7C15U));
    // explicit rounding intrinsics: the compiler must not contract s0*n + nom into one FMA, or
    // the sequence would no longer be the specified (host-reproducible) one
#if defined(__CUDA_ARCH__)
    d[c] = (k == 0) ? __fmul_rn(S.s0[c], n) : __fmaf_rn(S.s[c], n, d[c]);
    const float t = __fadd_rn(__ldg(unom_k + live_chan(c)), d[c]);
#else
    d[c] = (k == 0) ? S.s0[c] * n : fmaf(S.s[c], n, d[c]);
    const float t = unom_k[live_chan(c)] + d[c];
#endif
    un[c] = min_nan(max_nan(t, S.lo[c]), S.hi[c]);     // NaN in the nominal sequence stays NaN
  }
}

}  // namespace airc

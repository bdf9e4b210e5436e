// airc_tc.cuh -- the surrogate's two layers on the 5th-generation tensor cores (tcgen05, TF32, accumulators in
// tensor memory) for the packed sweep kernels.
//
// Why: in the FFMA2 kernels the network is 352 of the ~550 packed FMAs of one f evaluation and the FMA pipe is
// 70 % busy next to a 66 % busy XU pipe (profiles/r2n_ncu_summary_sweep2.txt): the kernel is bound by everything
// at once.  The layers are tiny GEMMs over the trajectories of a thread group,
//     H[256 x 32] = X[256 x 8] . W1^T          X = (1, aileron, elevator, qbar, alpha, beta, 0, 0)
//     C[256 x 6]  = R[256 x 32] . Wo^T + c0    R = 1 / (1 + 2^H)         (coefficient_models.py:99-104)
// so they go to the tensor pipe, which is otherwise idle, as M = 128 tiles (one row per thread, two tiles per
// group of 128 threads: the lo and hi trajectories of the packed arithmetic).
//
// Precision: TF32 keeps 11 significand bits, the parity bar needs FP32.  Every product is therefore formed from
// split operands, a = a_hi + a_lo, b = b_hi + b_lo, as three MMAs a_hi b_hi + a_lo b_hi + a_hi b_lo accumulated in
// FP32 (the dropped a_lo b_lo term and the truncation of the lo parts are ~2^-21 relative).  The weights are split
// on the host (round to nearest); activations are split by the threads (the tensor core itself ignores the low 13
// bits of an operand, so a_hi is the raw value and a_lo = a - trunc(a)).
//
// Data flow per f evaluation (tcgen05.mma, A operand in tensor memory, B = weights in shared memory):
//   threads: tcgen05.st X_hi | X_lo -> fence -> group barrier
//   elected thread of the group's warp 0: 2 x 3 MMAs (M128 N32 K8) -> tcgen05.commit -> mbarrier
//   threads: mbarrier wait -> tcgen05.ld H (16 columns at a time) -> logistic -> tcgen05.st R | R_lo
//            (R overwrites the columns H was read from) -> fence -> group barrier
//   elected thread: 2 x 4 x 3 MMAs (M128 N16 K8) -> commit
//   threads: mbarrier wait -> tcgen05.ld C.
// Tensor-memory columns per group (160): per tile [0,16) X_hi | X_lo, later C; [16,48) H, later R; [48,80) R_lo.
//
// One row's result depends on that row's operands only (and on nothing else in the tile), so a trajectory's bits
// do not depend on its neighbours or on how the batch is sharded.
#pragma once
#include "airc_packed.cuh"

namespace airc {

// ---------------------------------------------------------------- weights image (host-built)
constexpr int TC_K1 = 8;        // layer 1 inputs incl. the constant 1 (bias) and two zero pads
constexpr int TC_N2 = 16;       // layer 2 outputs padded to the smallest N of an M = 128 MMA
// K-major, no swizzle: core matrix = 8 rows x 16 bytes, contiguous (128 B); 8-row groups 128 B apart (SBO),
// 16-byte K chunks (N/8) x 128 B apart (LBO).  Index in floats of element (n, k) of a [N][K] matrix:
AIRC_HD constexpr int tc_b_index(int n, int k, int N) {
  return (k >> 2) * (N / 8) * 32 + (n >> 3) * 32 + (n & 7) * 4 + (k & 3);
}
struct alignas(16) TcNetImage {
  float B1h[NH * TC_K1], B1l[NH * TC_K1];       // [32][8]
  float B2h[TC_N2 * NH], B2l[TC_N2 * NH];       // [16][32]
  float c0[8];
};

inline float tc_round_tf32(float x) {            // round to nearest (ties away), 10 explicit mantissa bits
  uint32_t u;
  memcpy(&u, &x, 4);
  u = (u + 0x1000u) & 0xFFFFE000u;
  float r;
  memcpy(&r, &u, 4);
  return r;
}

inline void make_tc_net(const UnitNet& n, TcNetImage& im) {
  memset(&im, 0, sizeof(im));
  for (int j = 0; j < NH; ++j) {
    const UnitW& w = n.u[j];
    const float row[TC_K1] = {w.a0, w.Au0, w.Au1, w.Ax0, w.Ax1, w.Ax2, 0.0f, 0.0f};
    for (int k = 0; k < TC_K1; ++k) {
      const float hi = tc_round_tf32(row[k]);
      im.B1h[tc_b_index(j, k, NH)] = hi;
      im.B1l[tc_b_index(j, k, NH)] = tc_round_tf32(row[k] - hi);
    }
    const float wo[6] = {w.wo0, w.wo1, w.wo2, w.wo3, w.wo4, w.wo5};
    for (int o = 0; o < 6; ++o) {
      const float hi = tc_round_tf32(wo[o]);
      im.B2h[tc_b_index(o, j, TC_N2)] = hi;
      im.B2l[tc_b_index(o, j, TC_N2)] = tc_round_tf32(wo[o] - hi);
    }
  }
  for (int o = 0; o < 6; ++o) im.c0[o] = n.c0[o];
}

#if defined(__CUDACC__)
namespace tc {

// ---------------------------------------------------------------- PTX
AIRC_D uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

AIRC_D uint64_t smem_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  return (uint64_t)((saddr >> 4) & 0x3FFFu) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16) |
         ((uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32) | ((uint64_t)1 << 46);   // version 1, no swizzle
}
// D F32 (bit 4), A and B TF32 (2 at bits 7, 10), both K-major, N >> 3 at bit 17, M >> 4 at bit 24
__host__ __device__ constexpr uint32_t idesc_tf32(int M, int N) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
AIRC_D void mma_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
               "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}\n"
               :: "r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}
AIRC_D void commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(bar) : "memory");
}
AIRC_D void fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
AIRC_D void fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
AIRC_D void wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
AIRC_D void wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
AIRC_D bool elect_one() {
  uint32_t p;
  asm volatile("{\n\t.reg .pred q;\n\telect.sync _|q, 0xFFFFFFFF;\n\tselp.u32 %0, 1, 0, q;\n\t}\n" : "=r"(p));
  return p != 0;
}
AIRC_D void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(bar), "r"(count) : "memory");
}
// Wait for the phase with the given parity.  The MMA it waits for completes within microseconds; the spin bound
// turns a programming error into a trap (reported as a CUDA error) instead of a hung device.
AIRC_D void mbar_wait(uint32_t bar, uint32_t parity) {
  for (uint32_t spin = 0; spin < (1u << 26); ++spin) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}\n"
                 : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    if (ok) return;
  }
  __trap();
}
AIRC_D void group_arrive(int id, int nthreads) { asm volatile("bar.arrive %0, %1;" :: "r"(id), "r"(nthreads) : "memory"); }
AIRC_D void group_sync(int id, int nthreads) { asm volatile("bar.sync %0, %1;" :: "r"(id), "r"(nthreads) : "memory"); }

AIRC_D void st16(uint32_t taddr, const float* v) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
               :: "r"(taddr), "f"(v[0]), "f"(v[1]), "f"(v[2]), "f"(v[3]), "f"(v[4]), "f"(v[5]), "f"(v[6]), "f"(v[7]),
                  "f"(v[8]), "f"(v[9]), "f"(v[10]), "f"(v[11]), "f"(v[12]), "f"(v[13]), "f"(v[14]), "f"(v[15]) : "memory");
}
AIRC_D void ld16(uint32_t taddr, float* v) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
               : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7]),
                 "=f"(v[8]), "=f"(v[9]), "=f"(v[10]), "=f"(v[11]), "=f"(v[12]), "=f"(v[13]), "=f"(v[14]), "=f"(v[15])
               : "r"(taddr) : "memory");
}
AIRC_D void ld8(uint32_t taddr, float* v) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7])
               : "r"(taddr) : "memory");
}
AIRC_D float trunc_tf32(float x) { return __uint_as_float(__float_as_uint(x) & 0xFFFFE000u); }

// ---------------------------------------------------------------- per-group context
constexpr int GROUP_THREADS = 128;
constexpr int COLS_PER_TILE = 80, COLS_PER_GROUP = 2 * COLS_PER_TILE;
constexpr int COL_X = 0, COL_H = 16, COL_RLO = 48;
constexpr uint32_t TC_LBO1 = (NH / 8) * 128, TC_LBO2 = (TC_N2 / 8) * 128;   // bytes between 16-byte K chunks of W1 / Wo     // within a tile; C re-uses COL_X, R re-uses COL_H

struct Group {
  uint32_t tm_group;      // tensor-memory address of the group's first column, lane 0
  uint32_t tm_lanes;      // the same with this warp's lane quarter in the lane field
  uint32_t mbar;          // shared-memory address of the group's mbarrier
  uint32_t parity;        // phase the next wait expects
  uint64_t d1h, d1l, d2h, d2l;   // shared-memory matrix descriptors of the weights (K step 0)
  int bar_id;             // named barrier of the group (1 + group index; 0 is __syncthreads)
  bool issuer;            // this warp issues the group's MMAs
};

// all threads of the group: operands are in tensor memory -> hand over to the issuing warp
AIRC_D void operands_ready(const Group& g) {
  wait_st();
  fence_before();
  if (g.issuer) group_sync(g.bar_id, GROUP_THREADS);
  else group_arrive(g.bar_id, GROUP_THREADS);
}
AIRC_D void results_wait(Group& g) {
  mbar_wait(g.mbar, g.parity);
  g.parity ^= 1u;
  fence_after();
}

// layer 1 for both tiles: H = X_hi W1_hi + X_lo W1_hi + X_hi W1_lo
AIRC_D void issue_layer1(const Group& g) {
  if (!g.issuer) return;
  fence_after();
  if (elect_one()) {
    const uint32_t id = idesc_tf32(128, NH);
#pragma unroll
    for (int t = 0; t < 2; ++t) {
      const uint32_t base = g.tm_group + t * COLS_PER_TILE;
      mma_ts(base + COL_H, base + COL_X, g.d1h, id, 0u);
      mma_ts(base + COL_H, base + COL_X + 8, g.d1h, id, 1u);
      mma_ts(base + COL_H, base + COL_X, g.d1l, id, 1u);
    }
    commit(g.mbar);
  }
  __syncwarp();
}
// layer 2 for both tiles: C = R W_hi + R_lo W_hi + R W_lo over four K steps of 8
AIRC_D void issue_layer2(const Group& g) {
  if (!g.issuer) return;
  fence_after();
  if (elect_one()) {
    const uint32_t id = idesc_tf32(128, TC_N2);
    constexpr uint64_t KSTEP = (2 * TC_LBO2) >> 4;     // one K step = two 16-byte chunks further on (address field, 16-byte units)
#pragma unroll
    for (int t = 0; t < 2; ++t) {
      const uint32_t base = g.tm_group + t * COLS_PER_TILE;
#pragma unroll
      for (int ks = 0; ks < NH / 8; ++ks) {
        const uint64_t dh = g.d2h + ks * KSTEP, dl = g.d2l + ks * KSTEP;
        mma_ts(base + COL_X, base + COL_H + 8 * ks, dh, id, ks ? 1u : 0u);
        mma_ts(base + COL_X, base + COL_RLO + 8 * ks, dh, id, 1u);
        mma_ts(base + COL_X, base + COL_H + 8 * ks, dl, id, 1u);
      }
    }
    commit(g.mbar);
  }
  __syncwarp();
}

}  // namespace tc

// Logistic of FOUR hidden units of one trajectory with ONE reciprocal (Montgomery's trick): 4 EX2 + 1 RCP instead of
// 4 + 4.  With the layers on the tensor cores the XU pipe is what the kernel waits for (MUFU: 40 % of the stall
// samples, profiles/r2p_ncu_summary_sweep2tc.txt), and issue slots are plentiful.  The pre-activations are clamped
// at 30 (NaN-preserving) so that the product of four denominators stays finite: r(30) = 9e-10, i.e. tanh is off by
// 2e-9 beyond |a| > 10.4 where it has long rounded to +-1.  All four units belong to the SAME trajectory: a
// trajectory's bits do not depend on its neighbours.
AIRC_D void sigm4_shared(const F2& h01, const F2& h23, F2& r01, F2& r23) {
  const F2 d01 = add2(F2{fast_ex2(min_nan(h01.lo, 30.0f)), fast_ex2(min_nan(h01.hi, 30.0f))}, bc2(1.0f));
  const F2 d23 = add2(F2{fast_ex2(min_nan(h23.lo, 30.0f)), fast_ex2(min_nan(h23.hi, 30.0f))}, bc2(1.0f));
  const F2 p = mul2(d01, d23);                       // (d0 d2, d1 d3)
  const float inv = fast_rcp(p.lo * p.hi);
  const F2 q{inv * p.hi, inv * p.lo};                // (1 / (d0 d2), 1 / (d1 d3))
  r01 = mul2(q, d23);                                // (1 / d0, 1 / d1)
  r23 = mul2(q, d01);                                // (1 / d2, 1 / d3)
}

// The surrogate for two trajectories per thread on the tensor cores, in two halves so that the caller can put work
// that does not need the coefficients between them (it then overlaps the second layer's round trip).  Called by ALL
// threads of the group in lockstep (group-wide barriers inside).  Same contract as surrogate2 (airc_packed.cuh).
AIRC_D void surrogate2_tc_begin(tc::Group& g, const F2& qbar, const F2& alpha, const F2& beta, const Ctl2& c) {
  using namespace tc;
  // ---- X = (1, aileron, elevator, qbar, alpha, beta, 0, 0), hi | lo
  {
    const F2 in[5] = {c.da, c.de, qbar, alpha, beta};
    float x0[16], x1[16];
    x0[0] = x1[0] = 1.0f;
    x0[8] = x1[8] = 0.0f;
#pragma unroll
    for (int i = 0; i < 5; ++i) {
      const F2 hi{trunc_tf32(in[i].lo), trunc_tf32(in[i].hi)};
      const F2 lo = sub2(in[i], hi);
      x0[1 + i] = in[i].lo; x1[1 + i] = in[i].hi;      // the tensor core reads the top 19 bits: raw value = hi part
      x0[9 + i] = lo.lo; x1[9 + i] = lo.hi;
    }
    x0[6] = x0[7] = x1[6] = x1[7] = 0.0f;
    x0[14] = x0[15] = x1[14] = x1[15] = 0.0f;
    st16(g.tm_lanes + COL_X, x0);
    st16(g.tm_lanes + COLS_PER_TILE + COL_X, x1);
  }
  operands_ready(g);
  issue_layer1(g);
  results_wait(g);
  // ---- R = 1 / (1 + 2^H), 16 columns at a time (the next 16 are in flight meanwhile); R goes back where H was,
  //      R_lo beside it
  float buf[2][16];
  ld16(g.tm_lanes + COL_H, buf[0]);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const uint32_t col = g.tm_lanes + (i >> 1) * COLS_PER_TILE + 16 * (i & 1);
    float* h = buf[i & 1];
    float rl[16];
    wait_ld();
    if (i < 3) ld16(g.tm_lanes + ((i + 1) >> 1) * COLS_PER_TILE + 16 * ((i + 1) & 1) + COL_H, buf[(i + 1) & 1]);
#pragma unroll
    for (int j = 0; j < 16; j += 4) {
      F2 r01, r23;
      sigm4_shared(F2{h[j], h[j + 1]}, F2{h[j + 2], h[j + 3]}, r01, r23);
      const F2 l01 = sub2(r01, F2{trunc_tf32(r01.lo), trunc_tf32(r01.hi)});
      const F2 l23 = sub2(r23, F2{trunc_tf32(r23.lo), trunc_tf32(r23.hi)});
      h[j] = r01.lo; h[j + 1] = r01.hi; h[j + 2] = r23.lo; h[j + 3] = r23.hi;
      rl[j] = l01.lo; rl[j + 1] = l01.hi; rl[j + 2] = l23.lo; rl[j + 3] = l23.hi;
    }
    st16(col + COL_H, h);
    st16(col + COL_RLO, rl);
  }
  operands_ready(g);
  issue_layer2(g);
}
AIRC_D void surrogate2_tc_end(tc::Group& g, const float* __restrict__ c0_smem, F2 C[6]) {
  using namespace tc;
  results_wait(g);
  float c0v[8], c1v[8];
  ld8(g.tm_lanes + COL_X, c0v);
  ld8(g.tm_lanes + COLS_PER_TILE + COL_X, c1v);
  wait_ld();
#pragma unroll
  for (int k = 0; k < 6; ++k) C[k] = adds(F2{c0v[k], c1v[k]}, c0_smem[k]);
}

struct SurTc {
  tc::Group& g;
  const float* c0;          // output-layer constants in shared memory
  AIRC_D void begin(const F2& qbar, const F2& alpha, const F2& beta, const Ctl2& c) {
    surrogate2_tc_begin(g, qbar, alpha, beta, c);
  }
  AIRC_D void end(F2 C[6]) { surrogate2_tc_end(g, c0, C); }
};

// rk4_step2 (airc_packed.cuh) with the 13 weighted sums of the stage derivatives parked in SHARED memory between
// stages (thread-private slots acc[i * stride], 64-bit accesses, conflict-free): they are touched once per stage,
// and the 26 registers they free are what the logistic phase needs to keep 16 hidden units in flight without
// spilling.  Same operations in the same order as rk4_step2: bit-identical results for the same surrogate.
template <class Sur>
AIRC_D void rk4_step2_smem(const DevConst& P, Sur& sur, State2& x, const Ctl2& c, float h, F2& speed2,
                           float2* __restrict__ acc, int stride) {
  auto ld = [&](int i) { const float2 t = acc[i * stride]; return F2{t.x, t.y}; };
  auto st = [&](int i, const F2& v) { acc[i * stride] = make_float2(v.lo, v.hi); };
  const float hh = 0.5f * h, h6 = h * (1.0f / 6.0f);
  F2 sv[3], sq[4], sw[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) { sv[i] = x.v[i]; sw[i] = x.w[i]; }
#pragma unroll
  for (int i = 0; i < 4; ++i) sq[i] = x.q[i];
#pragma unroll
  for (int i = 0; i < 13; ++i) st(i, bc2(0.0f));
#pragma unroll 1
  for (int stage = 0; stage < 4; ++stage) {
    const float wgt = (stage == 0 || stage == 3) ? 1.0f : 2.0f;
    const float cn = (stage == 2) ? h : hh;
    F2 kv[3], kq[4], kw[3], vb2;
    f_eval2(P, sur, sv, sq, sw, c, kv, kq, kw, vb2);
    if (stage == 0) speed2 = vb2;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      st(i, fmas(wgt, sv[i], ld(i)));                  // p
      st(3 + i, fmas(wgt, kv[i], ld(3 + i)));      // v
      st(10 + i, fmas(wgt, kw[i], ld(10 + i)));    // w
      sv[i] = fmas(cn, kv[i], x.v[i]);
      sw[i] = fmas(cn, kw[i], x.w[i]);
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      st(6 + i, fmas(wgt, kq[i], ld(6 + i)));
      sq[i] = fmas(cn, kq[i], x.q[i]);
    }
  }
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    x.p[i] = fmas(h6, ld(i), x.p[i]);
    x.v[i] = fmas(h6, ld(3 + i), x.v[i]);
    x.w[i] = fmas(h6, ld(10 + i), x.w[i]);
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) x.q[i] = fmas(h6, ld(6 + i), x.q[i]);
}

template <class Sur>
AIRC_D void state_update2_smem(const DevConst& P, Sur& sur, State2& x, const Ctl2& c, float dt, int substeps,
                               bool normalise, F2& speed2, float2* __restrict__ acc, int stride) {
  const float h = (substeps == 1) ? dt : dt / (float)substeps;
  F2 sp2;
  rk4_step2_smem(P, sur, x, c, h, speed2, acc, stride);
  for (int s = 1; s < substeps; ++s) rk4_step2_smem(P, sur, x, c, h, sp2, acc, stride);
  if (normalise) normalise_q2(x);
}
#endif  // __CUDACC__

}  // namespace airc

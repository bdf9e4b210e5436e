// airc_jac_kernel.cuh -- the tile kernel of K2 (mathematics and design notes: airc_jac.cuh).
//
// One block = one tile of NT nodes, 2 NT threads.
//   phase 1: all 2 NT threads; threads t and t + NT share node tile*NT + t and alternate its RK4 stages (head =
//            primal k, handed to the partner through shared memory; tail = the f_x / f_u blocks of the stage point,
//            parked in the node record: 348 floats, 380 with the d/d(dt) column, 128-bit stores); x+ (and the
//            multiple-shooting defect) go straight to global memory.  The code needs ~250 registers per thread and
//            the records fill the shared memory, so NT = 128 nodes and 8 warps per SM is what fits.
//   phase 2: 4 passes of NP = NT/4 nodes; in a pass every group of 8 lanes runs the tangent recursion of one
//            node (one packed column pair per lane), drops its 13 x 2 results into a shared-memory IMAGE of
//            the pass's output (A, B, dPhi/ddt in the caller's layout; the constant columns dPhi/dp = [I;0]
//            and dPhi/d(thrust) = 0 are written into the image once per block), and the whole block streams
//            the image out with 128-bit stores: AoS -- NP x 169 (91, 13) contiguous floats; SoA -- one
//            16-byte chunk = four consecutive nodes of one matrix element.
// Shared memory: NT = 64 (shipped for SoA and d/d(dt)): 87 KB of records + 17..21 KB of image, two blocks per SM;
// NT = 32 (shipped for AoS): four blocks per SM; NT = 128: one block per SM (A/B variant only).
#pragma once
#include "airc_jac.cuh"
#include "airc_kernels.cuh"

namespace airc {

template <int NT, bool SOA, bool DT>
struct JacTile {
  static constexpr int TPB = 2 * NT;                    // threads per block
  static constexpr int NP = TPB / 8;                    // nodes per pass
  static constexpr int PASSES = NT / NP;                // 4
  static constexpr int EL = DT ? 273 : 260;             // image elements per node: A 169, B 91 (, dPhi/ddt 13)
  // SoA image row pitch: +4 keeps 16-byte rows and spreads the lanes' stores over the banks; the d/d(dt)
  // variant has no room left for the padding (stores then see bank conflicts; the rarer call)
  static constexpr int NPS = DT ? NP : NP + 4;
  static constexpr int IMG = SOA ? NPS * EL : NP * EL;
  static constexpr int OFF_B = SOA ? 169 * NPS : 169 * NP;
  static constexpr int OFF_D = SOA ? 260 * NPS : 260 * NP;
  static constexpr size_t SMEM = (size_t)(NT * JRec<DT>::NODE + IMG) * sizeof(float);
  static_assert(NP % 4 == 0 && (TPB / (NP / 4)) == 32, "write-out mapping");
  static_assert(2 * 23 * NT <= IMG, "phase 1's exchange area aliases the image");
};

// full-pass write-out, SoA: E elements x NP nodes; thread -> (element e0 + 32 k, 16-byte chunk `part`)
template <int E, int NP, int NPS>
AIRC_D void jk_out_soa(float* __restrict__ g, const float* __restrict__ im, long long n, int tid) {
  constexpr int CH = NP / 4, IT = (E + 31) / 32;
  const int e0 = tid / CH, part = tid - e0 * CH;
  float4 v[IT];
#pragma unroll
  for (int k = 0; k < IT; ++k)
    if (e0 + 32 * k < E) v[k] = *reinterpret_cast<const float4*>(im + (e0 + 32 * k) * NPS + 4 * part);
  float* d = g + (long long)e0 * n + 4 * part;
#pragma unroll
  for (int k = 0; k < IT; ++k, d += 32 * n)
    if (e0 + 32 * k < E) *reinterpret_cast<float4*>(d) = v[k];
}

// full-pass write-out, AoS: COUNT4 consecutive 16-byte chunks
template <int COUNT4, int TPB>
AIRC_D void jk_out_lin(float* __restrict__ g, const float* __restrict__ im, int tid) {
  constexpr int IT = (COUNT4 + TPB - 1) / TPB;
  float4 v[IT];
#pragma unroll
  for (int k = 0; k < IT; ++k)
    if (tid + k * TPB < COUNT4) v[k] = reinterpret_cast<const float4*>(im)[tid + k * TPB];
#pragma unroll
  for (int k = 0; k < IT; ++k)
    if (tid + k * TPB < COUNT4) reinterpret_cast<float4*>(g)[tid + k * TPB] = v[k];
}

// ------------------------------------------------------------------ phase 1: TWO warps per 32 nodes
// Thread t and thread t + NT work on the same node and alternate the RK4 stages (t: stages 1 and 3, t + NT: 2 and
// 4).  f_jac_head delivers the primal k after a quarter of a stage's work; the thread then hands the next stage
// point (and the running RK4 sums) to its partner through shared memory and assembles its Jacobian blocks
// (f_jac_tail) WHILE the partner already runs the next stage's head: both warps of a pair sit on the same
// scheduler, so the lone-warp dependency stalls of one fill with the other's instructions.  Hand-over = a named
// barrier per warp pair and direction (producer bar.arrive, consumer bar.sync, 64 threads); the exchange area
// aliases the output image, which phase 2 initialises afterwards.
template <int NT, bool DT>
AIRC_D void jk_phase1(const DevConst& P, const LinArgs& a, long long node0, int tid, float* __restrict__ recs,
                      float* __restrict__ xch, int bar0) {
  using L = JRec<DT>;
  constexpr int WP = NT / 32;                          // warp pairs
  const int pnode = tid & (NT - 1), half = tid / NT, pw = (tid >> 5) & (WP - 1);
  const long long node = node0 + pnode;
  const bool live = node < a.n;
  float* const rec = recs + pnode * L::NODE;
  const int bar_ab = bar0 + pw, bar_ba = bar0 + WP + pw;     // first half -> second half, and back
  State s;
  Ctl c = {};
  float dt = 0.0f;
  if (live) {
    load_state(a.x, node, s);
    c = load_ctl(a.u.base + node * a.u.sb, a.u.sc);
    dt = a.dt_per_node ? __ldg(a.dt_per_node + node) : a.dt;
  }
  const float hh = 0.5f * dt, h6 = dt * (1.0f / 6.0f);
  float sv[3], sq[4], sw[3], ap[3], av[3], aq[4], aw[3];
#pragma unroll 1
  for (int it = 0; it < 2; ++it) {
    const int stage = 2 * it + half;
    const float wgt = (stage == 0 || stage == 3) ? 1.0f : 2.0f;
    const float cn = (stage == 2) ? dt : hh;
    // ---- receive the stage point and the sums so far
    if (stage == 0) {
#pragma unroll
      for (int i = 0; i < 3; ++i) { sv[i] = s.v[i]; sw[i] = s.w[i]; ap[i] = av[i] = aw[i] = 0.0f; }
#pragma unroll
      for (int i = 0; i < 4; ++i) { sq[i] = s.q[i]; aq[i] = 0.0f; }
    } else {
      asm volatile("bar.sync %0, 64;" :: "r"(half ? bar_ab : bar_ba) : "memory");
      const float* mi = xch + (half ? 0 : 23 * NT) + pnode;
#pragma unroll
      for (int i = 0; i < 3; ++i) {
        sv[i] = mi[i * NT]; sw[i] = mi[(7 + i) * NT];
        ap[i] = mi[(10 + i) * NT]; av[i] = mi[(13 + i) * NT]; aw[i] = mi[(20 + i) * NT];
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) { sq[i] = mi[(3 + i) * NT]; aq[i] = mi[(16 + i) * NT]; }
    }
    // ---- head: k = f(X_stage)
    float kv[3], kq[4], kw[3];
    JacMid m;
    if (live) f_jac_head(P, sv, sq, sw, c, kv, kq, kw, m);
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      ap[i] = fmaf(wgt, sv[i], ap[i]);
      av[i] = fmaf(wgt, kv[i], av[i]);
      aw[i] = fmaf(wgt, kw[i], aw[i]);
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) aq[i] = fmaf(wgt, kq[i], aq[i]);
    // ---- hand over (stages 1..3), or finish the step (stage 4)
    if (stage < 3) {
      float* mo = xch + (half ? 23 * NT : 0) + pnode;
#pragma unroll
      for (int i = 0; i < 3; ++i) {
        mo[i * NT] = fmaf(cn, kv[i], s.v[i]); mo[(7 + i) * NT] = fmaf(cn, kw[i], s.w[i]);
        mo[(10 + i) * NT] = ap[i]; mo[(13 + i) * NT] = av[i]; mo[(20 + i) * NT] = aw[i];
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) { mo[(3 + i) * NT] = fmaf(cn, kq[i], s.q[i]); mo[(16 + i) * NT] = aq[i]; }
      asm volatile("bar.arrive %0, 64;" :: "r"(half ? bar_ba : bar_ab) : "memory");
    } else if (live) {
      Q4* hd = reinterpret_cast<Q4*>(rec + L::HDR);
      hd[0] = Q4{dt, ap[0], ap[1], ap[2]};
      hd[1] = Q4{aq[0], aq[1], aq[2], aq[3]};
#pragma unroll
      for (int i = 0; i < 3; ++i) {
        s.p[i] = fmaf(h6, ap[i], s.p[i]);
        s.v[i] = fmaf(h6, av[i], s.v[i]);
        s.w[i] = fmaf(h6, aw[i], s.w[i]);
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) s.q[i] = fmaf(h6, aq[i], s.q[i]);
      if (a.normalise) normalise_q<0>(s, nullptr);
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
    }
    // ---- tail: the Jacobian blocks of this stage point -> node record
    if (live) {
      float r[92];
      r[82] = r[83] = 0.0f;
      f_jac_tail(P, sv, sq, sw, c, m, r);
      if (DT) {
#pragma unroll
        for (int i = 0; i < 3; ++i) { r[JR_K + i] = kv[i]; r[JR_K + 7 + i] = kw[i]; }
#pragma unroll
        for (int i = 0; i < 4; ++i) r[JR_K + 3 + i] = kq[i];
      }
      Q4* dst = reinterpret_cast<Q4*>(rec + stage * L::STAGE);
#pragma unroll
      for (int i = 0; i < L::STAGE / 4; ++i) dst[i] = Q4{r[4 * i], r[4 * i + 1], r[4 * i + 2], r[4 * i + 3]};
    }
  }
}

// Phase 2 for the NT nodes whose records are in `recs`, by 2 NT threads (tid 0 .. 2 NT - 1); `sync` is the barrier
// of exactly those threads.
template <int NT, bool SOA, bool DT, class Sync>
AIRC_D void jk_phase2(const LinArgs& a, long long node0, int tid, const float* __restrict__ recs, float* __restrict__ img,
                      int vec_ok, Sync sync) {
  using T = JacTile<NT, SOA, DT>;
  using L = JRec<DT>;
  constexpr int TPB2 = T::TPB;
  float* const imgB = img + T::OFF_B;
  float* const imgD = img + T::OFF_D;
  // ---- the constant columns of the image, once per block: dPhi/dp = [I;0] (f does not read p) and
  //      dPhi/d(thrust) = 0 (thrust is inert, aircraft.py:320)
  for (int i = tid; i < T::NP * 39; i += TPB2) {
    const int j = i / 39, rem = i - j * 39, r = rem / 3, cc = rem - r * 3;
    if (SOA) {
      img[(r * 13 + cc) * T::NPS + j] = (r == cc) ? 1.0f : 0.0f;
      imgB[(r * 7 + 3 + cc) * T::NPS + j] = 0.0f;
    } else {
      img[j * 169 + r * 13 + cc] = (r == cc) ? 1.0f : 0.0f;
      imgB[j * 91 + r * 7 + 3 + cc] = 0.0f;
    }
  }

  // ------------------------------------------------------------------ phase 2: eight lanes per node
  const int lane8 = tid & 7, grp = tid >> 3;
  // where this lane's column pair goes in the image: row r of the lo / hi column at off + r * rst
  int off_lo, off_hi, rst;
  {
    int e_lo, e_hi, w;                                   // element indices of row 0, row width
    if (lane8 < 5) { e_lo = 3 + 2 * lane8; e_hi = e_lo + 1; w = 13; }
    else if (lane8 == 5) { e_lo = 169; e_hi = 170; w = 7; }
    else if (lane8 == 6) { e_lo = 171; e_hi = 175; w = 7; }
    else { e_lo = 260; e_hi = 260; w = 1; }
    if (SOA) { off_lo = e_lo * T::NPS + grp; off_hi = e_hi * T::NPS + grp; rst = w * T::NPS; }
    else {
      const int nb = lane8 < 5 ? grp * 169 : (lane8 < 7 ? T::OFF_B + grp * 91 - 169 : T::OFF_D + grp * 13 - 260);
      off_lo = nb + e_lo; off_hi = nb + e_hi; rst = w;
    }
  }
  const bool has_hi = lane8 != 7;
#pragma unroll 1
  for (int pass = 0; pass < T::PASSES; ++pass) {
    const int jn = pass * T::NP + grp;                  // node within the tile
    const long long pnode0 = node0 + (long long)pass * T::NP;
    if (pnode0 >= a.n) break;                            // block-uniform
    if (node0 + jn < a.n && (DT || lane8 < 7)) {
      F2 col[13];
      jac_phase2<DT>(recs + jn * L::NODE, lane8, a.normalise != 0, col);
      float* o = img + off_lo;
      float* oh = img + off_hi;
#pragma unroll
      for (int r = 0; r < 13; ++r, o += rst, oh += rst) {
        *o = col[r].lo;
        if (has_hi) *oh = col[r].hi;
      }
    }
    sync();
    // ---- stream the image out
    const long long left = a.n - pnode0;
    const int m = left < T::NP ? (int)left : T::NP;    // valid nodes of this pass
    if (SOA) {
      // element e of nodes pnode0 .. pnode0+m-1: global base[e*n + pnode0 ..], image row e (pitch NPS)
      if (vec_ok && m == T::NP) {                        // n % 4 == 0 and 16-byte aligned bases (host-checked)
        if (a.A) jk_out_soa<169, T::NP, T::NPS>(a.A + pnode0, img, a.n, tid);
        if (a.Bm) jk_out_soa<91, T::NP, T::NPS>(a.Bm + pnode0, imgB, a.n, tid);
        if (DT && a.Dt) jk_out_soa<13, T::NP, T::NPS>(a.Dt + pnode0, imgD, a.n, tid);
      } else {
        auto out = [&](float* g, const float* im, int E) {
          if (!g) return;
          for (int i = tid; i < E * T::NP; i += TPB2) {
            const int e = i / T::NP, j = i - e * T::NP;
            if (j < m) g[(long long)e * a.n + pnode0 + j] = im[e * T::NPS + j];
          }
        };
        out(a.A, img, 169);
        out(a.Bm, imgB, 91);
        if (DT) out(a.Dt, imgD, 13);
      }
    } else {
      if (vec_ok && m == T::NP) {                        // 16-byte aligned bases; pnode0 % 4 == 0 by construction
        if (a.A) jk_out_lin<T::NP * 169 / 4, TPB2>(a.A + pnode0 * 169, img, tid);
        if (a.Bm) jk_out_lin<T::NP * 91 / 4, TPB2>(a.Bm + pnode0 * 91, imgB, tid);
        if (DT && a.Dt) jk_out_lin<T::NP * 13 / 4, TPB2>(a.Dt + pnode0 * 13, imgD, tid);
      } else {
        auto out = [&](float* g, const float* im, int count) {
          if (!g) return;
          for (int i = tid; i < count; i += TPB2) g[i] = im[i];
        };
        out(a.A ? a.A + pnode0 * 169 : nullptr, img, m * 169);
        out(a.Bm ? a.Bm + pnode0 * 91 : nullptr, imgB, m * 91);
        if (DT) out(a.Dt ? a.Dt + pnode0 * 13 : nullptr, imgD, m * 13);
      }
    }
    sync();
  }
}

template <int NT, bool SOA, bool DT>
__global__ void __launch_bounds__(2 * NT, (NT == 128) ? 1 : ((NT == 64) ? 2 : 4))
linearize_tile_kernel(const __grid_constant__ DevConst P, const LinArgs a, const int vec_ok) {
  using T = JacTile<NT, SOA, DT>;
  using L = JRec<DT>;
  extern __shared__ float4 jk_smem4[];
  float* const recs = reinterpret_cast<float*>(jk_smem4);
  float* const img = recs + NT * L::NODE;
  const int tid = threadIdx.x;
  const long long node0 = (long long)blockIdx.x * NT;
  jk_phase1<NT, DT>(P, a, node0, tid, recs, img /* the exchange area aliases the image */, 1);
  __syncthreads();
  jk_phase2<NT, SOA, DT>(a, node0, tid, recs, img, vec_ok, [] { __syncthreads(); });
}

}  // namespace airc

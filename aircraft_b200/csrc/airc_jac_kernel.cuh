// airc_jac_kernel.cuh -- the tile kernel of K2 (mathematics and design notes: airc_jac.cuh).
//
// One block = one tile of NT nodes, 2 NT threads.
//   phase 1: threads 0 .. NT-1, thread t computes node tile*NT + t (primal RK4, surrogate Jacobian, f_x / f_u
//            blocks of the four stage points) and parks the node record (348 floats; 380 with the d/d(dt)
//            column) in shared memory with 128-bit stores; x+ (and the multiple-shooting defect) go straight
//            to global memory.  Phase 1 needs ~250 registers per thread and its records fill the shared
//            memory, so NT = 128 nodes per SM is what fits; the second half of the block only joins for
//   phase 2: 4 passes of NP = NT/4 nodes; in a pass every group of 8 lanes runs the tangent recursion of one
//            node (one packed column pair per lane), drops its 13 x 2 results into a shared-memory IMAGE of
//            the pass's output (A, B, dPhi/ddt in the caller's layout; the constant columns dPhi/dp = [I;0]
//            and dPhi/d(thrust) = 0 are written into the image once per block), and the whole block streams
//            the image out with 128-bit stores: AoS -- NP x 169 (91, 13) contiguous floats; SoA -- one
//            16-byte chunk = four consecutive nodes of one matrix element.
// Shared memory (NT = 128): 174 KB of records + 33..37 KB of image, one block (8 warps) per SM; NT = 64: two
// blocks per SM (small problems: more blocks to spread over the SMs).
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

template <int NT, bool SOA, bool DT>
__global__ void __launch_bounds__(2 * NT, (NT == 128) ? 1 : ((NT == 64) ? 2 : 4))
linearize_tile_kernel(const __grid_constant__ DevConst P, const LinArgs a, const int vec_ok) {
  using T = JacTile<NT, SOA, DT>;
  using L = JRec<DT>;
  extern __shared__ float4 jk_smem4[];
  float* const recs = reinterpret_cast<float*>(jk_smem4);
  float* const img = recs + NT * L::NODE;
  float* const imgB = img + T::OFF_B;
  float* const imgD = img + T::OFF_D;
  const int tid = threadIdx.x;
  const long long node0 = (long long)blockIdx.x * NT;

  // ---- the constant columns of the image, once per block: dPhi/dp = [I;0] (f does not read p) and
  //      dPhi/d(thrust) = 0 (thrust is inert, aircraft.py:320)
  for (int i = tid; i < T::NP * 39; i += T::TPB) {
    const int j = i / 39, rem = i - j * 39, r = rem / 3, cc = rem - r * 3;
    if (SOA) {
      img[(r * 13 + cc) * T::NPS + j] = (r == cc) ? 1.0f : 0.0f;
      imgB[(r * 7 + 3 + cc) * T::NPS + j] = 0.0f;
    } else {
      img[j * 169 + r * 13 + cc] = (r == cc) ? 1.0f : 0.0f;
      imgB[j * 91 + r * 7 + 3 + cc] = 0.0f;
    }
  }

  // ------------------------------------------------------------------ phase 1: one node per thread
  if (tid < NT) {
    const long long node = node0 + tid;
    if (node < a.n) {
      State s;
      load_state(a.x, node, s);
      const Ctl c = load_ctl(a.u.base + node * a.u.sb, a.u.sc);
      const float dt = a.dt_per_node ? __ldg(a.dt_per_node + node) : a.dt;
      jac_phase1<DT>(P, s, c, dt, recs + tid * L::NODE);
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
  }
  __syncthreads();

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
    __syncthreads();
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
          for (int i = tid; i < E * T::NP; i += T::TPB) {
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
        if (a.A) jk_out_lin<T::NP * 169 / 4, T::TPB>(a.A + pnode0 * 169, img, tid);
        if (a.Bm) jk_out_lin<T::NP * 91 / 4, T::TPB>(a.Bm + pnode0 * 91, imgB, tid);
        if (DT && a.Dt) jk_out_lin<T::NP * 13 / 4, T::TPB>(a.Dt + pnode0 * 13, imgD, tid);
      } else {
        auto out = [&](float* g, const float* im, int count) {
          if (!g) return;
          for (int i = tid; i < count; i += T::TPB) g[i] = im[i];
        };
        out(a.A ? a.A + pnode0 * 169 : nullptr, img, m * 169);
        out(a.Bm ? a.Bm + pnode0 * 91 : nullptr, imgB, m * 91);
        if (DT) out(a.Dt ? a.Dt + pnode0 * 13 : nullptr, imgD, m * 13);
      }
    }
    __syncthreads();
  }
}

}  // namespace airc

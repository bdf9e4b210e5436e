// host_emul.cpp -- TEST INFRASTRUCTURE ONLY (lives under oracle/, never linked into the product).
//
// Compiles the product's FP32 device formulas (aircraft_b200/csrc/airc_*.cuh, all AIRC_HD)
// with g++ so that the `-m "not gpu"` tests can check the kernel arithmetic -- the folded
// network, the rotation-matrix form, the forward-mode tangents -- against the float64 oracle
// on a machine without a GPU.  MUFU approximations are replaced by libm here, so this checks
// the algebra, not the last ulp; the GPU parity tests check the real thing.
#include <stdint.h>

#include "../aircraft_b200/csrc/airc_fuse.hpp"
#include "../aircraft_b200/csrc/airc_step.cuh"
#include "../aircraft_b200/csrc/airc_coop.cuh"
#include "../aircraft_b200/csrc/airc_sampler.cuh"
#include "../aircraft_b200/csrc/airc_hess.cuh"
#include "../aircraft_b200/csrc/airc_hess2.cuh"
#include "../aircraft_b200/csrc/airc_jac.cuh"

using namespace airc;

static void load(const float* x, const float* u, State& s, Ctl& c) {
  for (int i = 0; i < 3; ++i) { s.p[i] = x[i]; s.v[i] = x[3 + i]; s.w[i] = x[10 + i]; }
  for (int i = 0; i < 4; ++i) s.q[i] = x[6 + i];
  c.da = u[0]; c.de = u[1]; c.dr = u[2]; c.df = u[6];
}
static void store(const State& s, float* x) {
  for (int i = 0; i < 3; ++i) { x[i] = s.p[i]; x[3 + i] = s.v[i]; x[10 + i] = s.w[i]; }
  for (int i = 0; i < 4; ++i) x[6 + i] = s.q[i];
}

extern "C" {

void emul_fused_net(const AircNNBlob* nn, float* out /* sizeof(FusedNet)/4 floats */) {
  FusedNet f;
  fuse_network(*nn, f);
  memcpy(out, &f, sizeof(f));
}

void emul_surrogate(const AircNNBlob* nn, const float* in5, int64_t n, float* out6) {
  FusedNet f;
  fuse_network(*nn, f);
  for (int64_t i = 0; i < n; ++i) {
    Ctl c{in5[5 * i + 3], in5[5 * i + 4], 0.f, 0.f};
    float J[6][5];
    surrogate<false>(f, in5[5 * i], in5[5 * i + 1], in5[5 * i + 2], c, out6 + 6 * i, J);
  }
}

void emul_state_derivative(const AircParams* p, const AircNNBlob* nn, const float* x,
                           const float* u, int64_t n, float* xdot) {
  DevConst P;
  build_devconst(*p, *nn, P);
  for (int64_t i = 0; i < n; ++i) {
    State s; Ctl c; Stage st; float J[6][5];
    load(x + 13 * i, u + 7 * i, s, c);
    float* o = xdot + 13 * i;
    f_eval<false>(P, s.v, s.q, s.w, c, o + 3, o + 6, o + 10, st, J);
    for (int k = 0; k < 3; ++k) o[k] = s.v[k];
  }
}

void emul_state_update(const AircParams* p, const AircNNBlob* nn, const float* x, const float* u,
                       float dt, int64_t n, int substeps, uint32_t flags, float* xn) {
  DevConst P;
  build_devconst(*p, *nn, P);
  for (int64_t i = 0; i < n; ++i) {
    State s; Ctl c;
    load(x + 13 * i, u + 7 * i, s, c);
    state_update<0>(P, s, c, dt, substeps, flags & AIRC_FLAG_NORMALISE, nullptr);
    store(s, xn + 13 * i);
  }
}

// A[n][13][13], B[n][13][7]
void emul_linearize(const AircParams* p, const AircNNBlob* nn, const float* x, const float* u,
                    float dt, int64_t n, int substeps, uint32_t flags, float* xn, float* A,
                    float* Bm) {
  DevConst P;
  build_devconst(*p, *nn, P);
  for (int64_t i = 0; i < n; ++i) {
    for (int col = 0; col < 20; ++col) {
      State s; Ctl c;
      load(x + 13 * i, u + 7 * i, s, c);
      Tangent T;
      memset(&T, 0, sizeof(T));
      float* tp = &T.p[0];  // p,v,q,w are contiguous: 13 floats
      if (col < 13) tp[col] = 1.0f;
      else {
        const int cu = col - 13;
        if (cu == 0) T.dc.da = 1.0f;
        if (cu == 1) T.dc.de = 1.0f;
        if (cu == 2) T.dc.dr = 1.0f;
        if (cu == 6) T.dc.df = 1.0f;
      }
      state_update<1>(P, s, c, dt, substeps, flags & AIRC_FLAG_NORMALISE, &T);
      if (col == 0) store(s, xn + 13 * i);
      for (int r = 0; r < 13; ++r) {
        if (col < 13) A[(i * 13 + r) * 13 + col] = tp[r];
        else Bm[(i * 13 + r) * 7 + (col - 13)] = tp[r];
      }
    }
  }
}

// The K2 arithmetic: packed (two-column) tangents through rk4_step_lin with the cooperative network
// policy degenerated to one lane (L = 1, no shuffles).  A[n][13][13], B[n][13][7].
static void linearize_packed_impl(const AircParams* p, const AircNNBlob* nn, const float* x, const float* u,
                                  float dt, int64_t n, int substeps, uint32_t flags, float* xn, float* A,
                                  float* Bm, float* Dt) {
  DevConst P;
  build_devconst(*p, *nn, P);
  const NetCoop<1> net{&P.net, 0};
  for (int64_t i = 0; i < n; ++i) {
    for (int pair = 0; pair < (Dt ? 8 : 7); ++pair) {  // slots (0,1) .. (12,13), (14 = dt, 15 idle) as in the kernel
      State s; Ctl c;
      load(x + 13 * i, u + 7 * i, s, c);
      const int s0 = 2 * pair, s1 = 2 * pair + 1;
      auto seed = [&](int slot) { return F2{s0 == slot ? 1.0f : 0.0f, s1 == slot ? 1.0f : 0.0f}; };
      TangentT<F2> D;
      for (int k = 0; k < 3; ++k) { D.p[k] = F2{0.f, 0.f}; D.v[k] = seed(k); D.w[k] = seed(7 + k); }
      for (int k = 0; k < 4; ++k) D.q[k] = seed(3 + k);
      D.dc.da = seed(10); D.dc.de = seed(11); D.dc.dr = seed(12); D.dc.df = seed(13);
      D.ddt = seed(14);
      state_update_lin<F2>(P, net, s, c, dt, substeps, flags & AIRC_FLAG_NORMALISE, D);
      if (pair == 0) store(s, xn + 13 * i);
      F2 col[13];
      for (int k = 0; k < 3; ++k) { col[k] = D.p[k]; col[3 + k] = D.v[k]; col[10 + k] = D.w[k]; }
      for (int k = 0; k < 4; ++k) col[6 + k] = D.q[k];
      static const int bcol[4] = {0, 1, 2, 6};
      for (int r = 0; r < 13; ++r) {
        const float vals[2] = {col[r].lo, col[r].hi};
        for (int hlf = 0; hlf < 2; ++hlf) {
          const int slot = 2 * pair + hlf;
          if (slot < 10) A[(i * 13 + r) * 13 + 3 + slot] = vals[hlf];
          else if (slot < 14) Bm[(i * 13 + r) * 7 + bcol[slot - 10]] = vals[hlf];
          else if (slot == 14) Dt[i * 13 + r] = vals[hlf];
        }
      }
    }
    for (int r = 0; r < 13; ++r) {
      for (int cc = 0; cc < 3; ++cc) A[(i * 13 + r) * 13 + cc] = (r == cc) ? 1.0f : 0.0f;
      for (int cc = 3; cc < 6; ++cc) Bm[(i * 13 + r) * 7 + cc] = 0.0f;
    }
  }
}

void emul_linearize_packed(const AircParams* p, const AircNNBlob* nn, const float* x, const float* u,
                           float dt, int64_t n, int substeps, uint32_t flags, float* xn, float* A,
                           float* Bm) {
  linearize_packed_impl(p, nn, x, u, dt, n, substeps, flags, xn, A, Bm, nullptr);
}
// ... plus the dPhi/d(dt) column, Dt[n][13]
void emul_linearize_packed_dt(const AircParams* p, const AircNNBlob* nn, const float* x, const float* u,
                              float dt, int64_t n, int substeps, uint32_t flags, float* xn, float* A,
                              float* Bm, float* Dt) {
  linearize_packed_impl(p, nn, x, u, dt, n, substeps, flags, xn, A, Bm, Dt);
}

// The tile kernel's arithmetic (airc_jac.cuh): phase 1 parks the f_x / f_u blocks of the four stage points in a
// node record, phase 2 runs the tangent recursion per column pair.  One substep.  A[n][13][13], B[n][13][7],
// Dt[n][13] (may be NULL).
void emul_linearize_tile(const AircParams* p, const AircNNBlob* nn, const float* x, const float* u, float dt,
                         int64_t n, uint32_t flags, float* xn, float* A, float* Bm, float* Dt) {
  DevConst P;
  build_devconst(*p, *nn, P);
  const bool norm = flags & AIRC_FLAG_NORMALISE;
  for (int64_t i = 0; i < n; ++i) {
    alignas(16) float rec[JRec<true>::NODE];
    State s; Ctl c;
    load(x + 13 * i, u + 7 * i, s, c);
    if (Dt) jac_phase1<true>(P, s, c, dt, rec);
    else jac_phase1<false>(P, s, c, dt, rec);
    if (norm) normalise_q<0>(s, nullptr);
    store(s, xn + 13 * i);
    static const int bcol[4] = {0, 1, 2, 6};
    for (int pair = 0; pair < (Dt ? 8 : 7); ++pair) {
      F2 col[13];
      if (Dt) jac_phase2<true>(rec, pair, norm, col);
      else jac_phase2<false>(rec, pair, norm, col);
      for (int r = 0; r < 13; ++r) {
        const float vals[2] = {col[r].lo, col[r].hi};
        for (int hlf = 0; hlf < 2; ++hlf) {
          const int slot = 2 * pair + hlf;
          if (slot < 10) A[(i * 13 + r) * 13 + 3 + slot] = vals[hlf];
          else if (slot < 14) Bm[(i * 13 + r) * 7 + bcol[slot - 10]] = vals[hlf];
          else if (slot == 14) Dt[i * 13 + r] = vals[hlf];
        }
      }
    }
    for (int r = 0; r < 13; ++r) {
      for (int cc = 0; cc < 3; ++cc) A[(i * 13 + r) * 13 + cc] = (r == cc) ? 1.0f : 0.0f;
      for (int cc = 3; cc < 6; ++cc) Bm[(i * 13 + r) * 7 + cc] = 0.0f;
    }
  }
}

// the registry's other coefficient models (airc_alt.cuh) through the product's f / RK4
void emul_model_state_update(const AircParams* p, int model, const float* model_params, const float* x,
                             const float* u, float dt, int64_t n, int substeps, uint32_t flags, float* xn) {
  DevConst P;
  build_devconst_model(*p, model, model_params, P);
  for (int64_t i = 0; i < n; ++i) {
    State s; Ctl c;
    load(x + 13 * i, u + 7 * i, s, c);
    state_update<0, 1>(P, s, c, dt, substeps, flags & AIRC_FLAG_NORMALISE, nullptr);
    store(s, xn + 13 * i);
  }
}
void emul_model_state_derivative(const AircParams* p, int model, const float* model_params, const float* x,
                                 const float* u, int64_t n, float* xdot) {
  DevConst P;
  build_devconst_model(*p, model, model_params, P);
  for (int64_t i = 0; i < n; ++i) {
    State s; Ctl c; Stage st; float J[6][5];
    load(x + 13 * i, u + 7 * i, s, c);
    float* o = xdot + 13 * i;
    f_eval<false, 1>(P, s.v, s.q, s.w, c, o + 3, o + 6, o + 10, st, J);
    for (int k = 0; k < 3; ++k) o[k] = s.v[k];
  }
}

// Jacobian by two-column forward mode on jets (linearize_jet_kernel's arithmetic) for the table models:
// xn[n][13], A[n][13][13], Bm[n][13][7], Dt[n][13]
void emul_model_linearize(const AircParams* p, int model, const float* model_params, const float* x,
                          const float* u, float dt, int64_t n, int substeps, uint32_t flags, float* xn, float* A,
                          float* Bm, float* Dt) {
  DevConst P;
  build_devconst_model(*p, model, model_params, P);
  memset(A, 0, sizeof(float) * 169 * n);
  memset(Bm, 0, sizeof(float) * 91 * n);
  for (int64_t i = 0; i < n; ++i) {
    State s; Ctl c;
    load(x + 13 * i, u + 7 * i, s, c);
    for (int r = 0; r < 3; ++r) A[(i * 13 + r) * 13 + r] = 1.0f;
    for (int lane = 0; lane < 8; ++lane) {
      Jet out[13];
      state_update_dual<1>(P, s, c, dt, substeps, flags & AIRC_FLAG_NORMALISE, 2 * lane, 2 * lane + 1, out);
      for (int half = 0; half < 2; ++half) {
        const int d = 2 * lane + half;
        for (int r = 0; r < 13; ++r) {
          const float val = half ? out[r].g.hi : out[r].g.lo;
          if (d < 10) A[(i * 13 + r) * 13 + 3 + d] = val;
          else if (d < 14) Bm[(i * 13 + r) * 7 + (d == 13 ? 6 : d - 10)] = val;
          else if (d == 14) Dt[i * 13 + r] = val;
        }
      }
      if (lane == 7) for (int r = 0; r < 13; ++r) xn[13 * i + r] = out[r].v;
    }
  }
}

void emul_model_hessian(const AircParams* p, int model, const float* model_params, const float* x, const float* u,
                        const float* lam, float dt, int64_t n, int substeps, uint32_t flags, float* H, float* grad,
                        float* value) {
  DevConst P;
  build_devconst_model(*p, model, model_params, P);
  memset(H, 0, sizeof(float) * NZ * NZ * n);
  for (int64_t i = 0; i < n; ++i) {
    State s; Ctl c;
    load(x + 13 * i, u + 7 * i, s, c);
    float* Hn = H + i * NZ * NZ;
    float* g = grad + i * NZ;
    for (int r = 0; r < 3; ++r) { g[r] = lam[13 * i + r]; g[16 + r] = 0.f; }
    for (int t = 0; t < HESS_LANES; ++t) {
      HessSeed seed;
      hess_task(t, seed.i, seed.j0, seed.j1);
      const Jet L = lagrangian_step_jet<1>(P, s, c, dt, substeps, flags & AIRC_FLAG_NORMALISE, lam + 13 * i, seed);
      const int zi = hess_zindex(seed.i), z0 = hess_zindex(seed.j0);
      Hn[zi * NZ + z0] = Hn[z0 * NZ + zi] = L.og.lo;
      if (seed.j1 < NHACT) {
        const int z1 = hess_zindex(seed.j1);
        Hn[zi * NZ + z1] = Hn[z1 * NZ + zi] = L.og.hi;
      }
      if (seed.j0 == seed.i) g[zi] = L.o;
      if (t == 0) value[i] = L.v;
    }
  }
}

// K3 arithmetic (airc_hess.cuh): the 64 (outer, inner-pair) tasks of a node, one after the other.
// H[n][21][21] (zero-filled here), grad[n][21], value[n]
void emul_hessian(const AircParams* p, const AircNNBlob* nn, const float* x, const float* u, const float* lam,
                  float dt, int64_t n, int substeps, uint32_t flags, float* H, float* grad, float* value) {
  DevConst P;
  build_devconst(*p, *nn, P);
  memset(H, 0, sizeof(float) * NZ * NZ * n);
  for (int64_t i = 0; i < n; ++i) {
    State s; Ctl c;
    load(x + 13 * i, u + 7 * i, s, c);
    float* Hn = H + i * NZ * NZ;
    float* g = grad + i * NZ;
    for (int r = 0; r < 3; ++r) { g[r] = lam[13 * i + r]; g[16 + r] = 0.f; }
    for (int t = 0; t < HESS_LANES; ++t) {
      HessSeed seed;
      hess_task(t, seed.i, seed.j0, seed.j1);
      const Jet L = lagrangian_step_jet(P, s, c, dt, substeps, flags & AIRC_FLAG_NORMALISE, lam + 13 * i, seed);
      const int zi = hess_zindex(seed.i), z0 = hess_zindex(seed.j0);
      Hn[zi * NZ + z0] = Hn[z0 * NZ + zi] = L.og.lo;
      if (seed.j1 < NHACT) {
        const int z1 = hess_zindex(seed.j1);
        Hn[zi * NZ + z1] = Hn[z1 * NZ + zi] = L.og.hi;
      }
      if (seed.j0 == seed.i) g[zi] = L.o;
      if (t == 0) value[i] = L.v;
    }
  }
}

// K4 arithmetic for short horizons (airc_hess.cuh, NetSharedJet): the jets with the network's value-level sums
// shared by the 64 threads of a node; every task adds the 32 units' terms in the device's order.
void emul_hessian_shared(const AircParams* p, const AircNNBlob* nn, const float* x, const float* u, const float* lam,
                         float dt, int64_t n, int substeps, uint32_t flags, float* H, float* grad, float* value) {
  DevConst P;
  build_devconst(*p, *nn, P);
  memset(H, 0, sizeof(float) * NZ * NZ * n);
  NetSharedJet net;
  net.prod = net.sums = nullptr;
  for (int64_t i = 0; i < n; ++i) {
    State s; Ctl c;
    load(x + 13 * i, u + 7 * i, s, c);
    float* Hn = H + i * NZ * NZ;
    float* g = grad + i * NZ;
    for (int r = 0; r < 3; ++r) { g[r] = lam[13 * i + r]; g[16 + r] = 0.f; }
    for (int t = 0; t < HESS_LANES; ++t) {
      HessSeed seed;
      hess_task(t, seed.i, seed.j0, seed.j1);
      net.t = t;
      const Jet L = lagrangian_step_jet<0, NetSharedJet>(P, s, c, dt, substeps, flags & AIRC_FLAG_NORMALISE,
                                                         lam + 13 * i, seed, net);
      const int zi = hess_zindex(seed.i), z0 = hess_zindex(seed.j0);
      Hn[zi * NZ + z0] = Hn[z0 * NZ + zi] = L.og.lo;
      if (seed.j1 < NHACT) {
        const int z1 = hess_zindex(seed.j1);
        Hn[zi * NZ + z1] = Hn[z1 * NZ + zi] = L.og.hi;
      }
      if (seed.j0 == seed.i) g[zi] = L.o;
      if (t == 0) value[i] = L.v;
    }
  }
}

// K4 arithmetic (airc_hess2.cuh, forward over reverse): the eight lanes of a node, one after the other, and the
// kernel's write-out (entries i <= j of column j with their mirror image).
void emul_hessian2(const AircParams* p, const AircNNBlob* nn, const float* x, const float* u, const float* lam,
                   float dt, int64_t n, int substeps, uint32_t flags, float* H, float* grad, float* value) {
  DevConst P;
  build_devconst(*p, *nn, P);
  alignas(16) float tab[NET_TAB_FLOATS];
  for (int i = 0; i < NET_TAB_FLOATS; ++i) net_tab_entry(P, i, tab);
  memset(H, 0, sizeof(float) * NZ * NZ * n);
  for (int64_t i = 0; i < n; ++i) {
    State s; Ctl c;
    load(x + 13 * i, u + 7 * i, s, c);
    float* Hn = H + i * NZ * NZ;
    float* g = grad + i * NZ;
    for (int r = 0; r < 3; ++r) { g[r] = lam[13 * i + r]; g[16 + r] = 0.f; }
    for (int lane = 0; lane < HESS2_LANES; ++lane) {
      Dual gd[NHACT];
      float val;
      hessian2_lane(P, tab, s, c, dt, substeps, (flags & AIRC_FLAG_NORMALISE) != 0, lam + 13 * i, lane, gd, val);
      for (int half = 0; half < H2_DIRS; ++half) {
        const int j = H2_DIRS * lane + half;
        if (j >= NHACT) continue;
        for (int r = 0; r <= j; ++r) {
          const float e = TanOps<Tan>::get(gd[r].t, half);
          Hn[hess_zindex(r) * NZ + hess_zindex(j)] = Hn[hess_zindex(j) * NZ + hess_zindex(r)] = e;
        }
      }
      if (lane == 0) {
        for (int r = 0; r < NHACT; ++r) g[hess_zindex(r)] = gd[r].v;
        value[i] = val;
      }
    }
  }
}

// the product's sampler (airc_sampler.cuh) on the host: U[count][H][7]
void emul_sweep_controls(const AircSweepParams* sp, const float* Unom, int H, int64_t first,
                         int64_t count, float* U) {
  Sampler s;
  for (int c = 0; c < 4; ++c) {
    s.s0[c] = (float)((double)sp->sigma0[c] / 147.80053);
    s.s[c] = (float)((double)sp->sigma[c] / 147.80053);
    s.lo[c] = sp->lo[c]; s.hi[c] = sp->hi[c];
  }
  s.seed_lo_mixed = mix32((unsigned int)(sp->seed & 0xFFFFFFFFull));
  s.seed_hi = (unsigned int)(sp->seed >> 32);
  s.index_offset = sp->index_offset;
  for (int64_t i = 0; i < count; ++i) {
    const unsigned int tk = traj_key(s.seed_lo_mixed, s.seed_hi, s.index_offset + (unsigned long long)(first + i));
    float d[4] = {0.f, 0.f, 0.f, 0.f};
    for (int k = 0; k < H; ++k) {
      float un[4];
      sample_step(s, tk, k, Unom + 7 * k, d, un);
      float* o = U + ((int64_t)i * H + k) * 7;
      for (int c = 0; c < 7; ++c) o[c] = Unom[7 * k + c];
      o[0] = un[0]; o[1] = un[1]; o[2] = un[2]; o[6] = un[3];
    }
  }
}
}  // extern "C"

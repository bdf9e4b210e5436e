// cpu_ref.cpp -- TEST INFRASTRUCTURE / CPU BASELINE ONLY.  Never linked into the product.
//
// A float64 C++ restatement of the reference's hot path *as written* (quaternion products,
// un-fused 5->16->32->tanh->6 network with explicit scaling), one trajectory per OpenMP thread.
// It stands in for the reference's CasADi-VM evaluation, which cannot run on the build or GPU
// machines (casadi / liecasadi / l4casadi are not installed), as the timed
// "reference CPU path" of bench.py (cpu_baseline.kind = "port").  It is validated against the
// float64 torch oracle (oracle/aircraft_oracle.py) in tests/test_cpu_ref.py.
//
// Follows:
//   /root/reference/src/aircraft/dynamics/base.py:147-177,229-241   v_frd_rel, airspeed, alpha, beta, qbar
//   /root/reference/src/aircraft/surrogates/models.py:143-155        ScaledModel.forward
//   /root/reference/src/aircraft/dynamics/coefficient_models.py:99-104  NeuralModel (+ rudder term)
//   /root/reference/src/aircraft/dynamics/aircraft.py:255-330        coefficients, forces, moments
//   /root/reference/src/aircraft/dynamics/base.py:280-297,344-406    forces_ned, q_dot, v_dot, omega_dot, f
//   /root/reference/src/aircraft/dynamics/base.py:408-480            state_step (RK4), state_update (fold)
// liecasadi 0.0.6 quaternion semantics (xyzw, Hamilton product, inverse = conj/|q|^2).
// Jacobians: forward-mode dual numbers with 20 tangent directions (what ca.jacobian delivers).
#include <math.h>
#include <stdint.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "../include/aircraft_b200.h"

namespace {

// ---------------------------------------------------------------- forward-mode scalar
template <int N>
struct Dual {
  double v;
  double d[N];
  Dual() : v(0.0) { for (int i = 0; i < N; ++i) d[i] = 0.0; }
  Dual(double a) : v(a) { for (int i = 0; i < N; ++i) d[i] = 0.0; }
};
#define DUAL_BIN(op, expr_v, expr_d)                                              \
  template <int N> inline Dual<N> operator op(const Dual<N>& a, const Dual<N>& b) { \
    Dual<N> r; r.v = expr_v; for (int i = 0; i < N; ++i) r.d[i] = expr_d; return r; }
DUAL_BIN(+, a.v + b.v, a.d[i] + b.d[i])
DUAL_BIN(-, a.v - b.v, a.d[i] - b.d[i])
DUAL_BIN(*, a.v * b.v, a.d[i] * b.v + a.v * b.d[i])
template <int N> inline Dual<N> operator/(const Dual<N>& a, const Dual<N>& b) {
  Dual<N> r; const double ib = 1.0 / b.v; r.v = a.v * ib;
  for (int i = 0; i < N; ++i) r.d[i] = (a.d[i] - r.v * b.d[i]) * ib;
  return r;
}
template <int N> inline Dual<N> operator-(const Dual<N>& a) {
  Dual<N> r; r.v = -a.v; for (int i = 0; i < N; ++i) r.d[i] = -a.d[i]; return r;
}
template <int N> inline Dual<N> operator+(const Dual<N>& a, double b) { Dual<N> r = a; r.v += b; return r; }
template <int N> inline Dual<N> operator-(const Dual<N>& a, double b) { Dual<N> r = a; r.v -= b; return r; }
template <int N> inline Dual<N> operator*(const Dual<N>& a, double b) {
  Dual<N> r; r.v = a.v * b; for (int i = 0; i < N; ++i) r.d[i] = a.d[i] * b; return r;
}
template <int N> inline Dual<N> operator*(double b, const Dual<N>& a) { return a * b; }
template <int N> inline Dual<N> operator/(const Dual<N>& a, double b) { return a * (1.0 / b); }
template <int N> inline Dual<N> chain(const Dual<N>& a, double fv, double fd) {
  Dual<N> r; r.v = fv; for (int i = 0; i < N; ++i) r.d[i] = fd * a.d[i]; return r;
}
template <int N> inline Dual<N> Sqrt(const Dual<N>& a) { double s = sqrt(a.v); return chain(a, s, 0.5 / s); }
template <int N> inline Dual<N> Tanh(const Dual<N>& a) { double t = tanh(a.v); return chain(a, t, 1.0 - t * t); }
template <int N> inline Dual<N> Asin(const Dual<N>& a) { return chain(a, asin(a.v), 1.0 / sqrt(1.0 - a.v * a.v)); }
template <int N> inline Dual<N> Exp(const Dual<N>& a) { double e = exp(a.v); return chain(a, e, e); }
template <int N> inline Dual<N> Abs(const Dual<N>& a) { return chain(a, fabs(a.v), a.v < 0 ? -1.0 : (a.v > 0 ? 1.0 : 0.0)); }
template <int N> inline Dual<N> Atan2(const Dual<N>& y, const Dual<N>& x) {
  Dual<N> r; r.v = atan2(y.v, x.v); const double den = 1.0 / (x.v * x.v + y.v * y.v);
  for (int i = 0; i < N; ++i) r.d[i] = (x.v * y.d[i] - y.v * x.d[i]) * den;
  return r;
}
template <int N> inline double Val(const Dual<N>& a) { return a.v; }
inline double Sqrt(double a) { return sqrt(a); }
inline double Tanh(double a) { return tanh(a); }
inline double Asin(double a) { return asin(a); }
inline double Exp(double a) { return exp(a); }
inline double Abs(double a) { return fabs(a); }
inline double Atan2(double y, double x) { return atan2(y, x); }
inline double Val(double a) { return a; }
inline double sign_casadi(double x) { return x > 0 ? 1.0 : (x < 0 ? -1.0 : x); }  // derivative 0

// ---------------------------------------------------------------- quaternions (xyzw)
template <class T> inline void qmul(const T a[4], const T b[4], T r[4]) {
  r[0] = a[3] * b[0] + b[3] * a[0] + (a[1] * b[2] - a[2] * b[1]);
  r[1] = a[3] * b[1] + b[3] * a[1] + (a[2] * b[0] - a[0] * b[2]);
  r[2] = a[3] * b[2] + b[3] * a[2] + (a[0] * b[1] - a[1] * b[0]);
  r[3] = a[3] * b[3] - (a[0] * b[0] + a[1] * b[1] + a[2] * b[2]);
}
template <class T> inline void qinv(const T q[4], T r[4]) {
  const T n2 = q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3];
  r[0] = -q[0] / n2; r[1] = -q[1] / n2; r[2] = -q[2] / n2; r[3] = q[3] / n2;
}

struct Model {
  AircParams p;
  double W1[16][5], b1[16], W2[32][16], b2[32], W3[6][32], b3[6];
  double in_mean[5], in_std[5], out_mean[6], out_std[6];
};

void make_model(const AircParams* p, const AircNNBlob* nn, Model& m) {
  m.p = *p;
  for (int i = 0; i < 16; ++i) { m.b1[i] = nn->b1[i]; for (int j = 0; j < 5; ++j) m.W1[i][j] = nn->W1[i * 5 + j]; }
  for (int i = 0; i < 32; ++i) { m.b2[i] = nn->b2[i]; for (int j = 0; j < 16; ++j) m.W2[i][j] = nn->W2[i * 16 + j]; }
  for (int i = 0; i < 6; ++i) { m.b3[i] = nn->b3[i]; for (int j = 0; j < 32; ++j) m.W3[i][j] = nn->W3[i * 32 + j]; }
  for (int i = 0; i < 5; ++i) { m.in_mean[i] = nn->input_mean[i]; m.in_std[i] = nn->input_std[i]; }
  for (int i = 0; i < 6; ++i) { m.out_mean[i] = nn->output_mean[i]; m.out_std[i] = nn->output_std[i]; }
}

// ScaledModel.forward, models.py:143-155 (no activation between the first two layers, :114-123)
template <class T> void surrogate(const Model& m, const T in[5], T out[6]) {
  T z[5], h1[16], h2[32];
  for (int i = 0; i < 5; ++i) z[i] = (in[i] - m.in_mean[i]) / m.in_std[i];
  for (int i = 0; i < 16; ++i) { T s = T(m.b1[i]); for (int j = 0; j < 5; ++j) s = s + z[j] * m.W1[i][j]; h1[i] = s; }
  for (int i = 0; i < 32; ++i) { T s = T(m.b2[i]); for (int j = 0; j < 16; ++j) s = s + h1[j] * m.W2[i][j]; h2[i] = Tanh(s); }
  for (int i = 0; i < 6; ++i) { T s = T(m.b3[i]); for (int j = 0; j < 32; ++j) s = s + h2[j] * m.W3[i][j]; out[i] = s * m.out_std[i] + m.out_mean[i]; }
}

// f(x,u), base.py:385-406
template <class T> void f(const Model& m, const T x[13], const T u[7], T xd[13]) {
  const AircParams& P = m.p;
  const double eps = P.epsilon;
  const T* v = x + 3; const T* q = x + 6; const T* w = x + 10;
  T qi[4], vq[4] = {v[0], v[1], v[2], T(0.0)}, t1[4], t2[4];
  qinv(q, qi);
  qmul(qi, vq, t1); qmul(t1, q, t2);                       // base.py:156-159
  T vb[3] = {t2[0] + eps, t2[1] + eps, t2[2] + eps};
  const T vb2 = vb[0] * vb[0] + vb[1] * vb[1] + vb[2] * vb[2];
  const T V = Sqrt(vb2 + eps);                             // base.py:167
  const T alpha = Atan2(vb[2], vb[0] + eps);               // base.py:176
  const T beta = Asin(vb[1] / V);                          // base.py:234
  const T qbar = 0.5 * 1.225 * vb2;                        // base.py:240
  T in[5] = {qbar, alpha, beta, u[0], u[1]}, C[6];
  surrogate(m, in, C);
  C[5] = C[5] + (-0.1) * 6 * u[2] * M_PI / 180;            // coefficient_models.py:102-103
  if (P.stall_scaling) {                                   // aircraft.py:280-294
    const double a30 = 30.0 * M_PI / 180.0;
    const T sa = T(1.0) / (Exp((Abs(alpha) - a30) * 10.0) + 1.0);
    const T sb = T(1.0) / (Exp((Abs(beta) - a30) * 10.0) + 1.0);
    C[2] = C[2] * sa * sb;
    C[4] = C[4] * sa;
  }
  C[0] = C[0] + (-0.1) * u[6];                             // aircraft.py:297-299
  C[2] = C[2] + (-0.6) * u[6];
  const T qS = qbar * P.reference_area;
  T F[3] = {C[0] * qS * sign_casadi(Val(vb[0])), C[1] * qS, C[2] * qS};   // aircraft.py:313,319
  T M[3] = {C[3] * qS * P.span + (P.com[1] * F[2] - P.com[2] * F[1]),      // aircraft.py:327-328
            C[4] * qS * P.chord + (P.com[2] * F[0] - P.com[0] * F[2]),     // base.py:255,275
            C[5] * qS * P.span + (P.com[0] * F[1] - P.com[1] * F[0])};
  T Fq[4] = {F[0], F[1], F[2], T(0.0)};
  qmul(q, Fq, t1); qmul(t1, qi, t2);                       // base.py:284-286
  xd[0] = v[0]; xd[1] = v[1]; xd[2] = v[2];
  for (int i = 0; i < 3; ++i) xd[3 + i] = t2[i] / P.mass + P.gravity[i];   // base.py:359
  T wq[4] = {w[0], w[1], w[2], T(0.0)};
  qmul(q, wq, t1);                                         // base.py:292-295
  for (int i = 0; i < 4; ++i) xd[6 + i] = t1[i] * 0.5;
  T Iw[3], r[3];
  for (int i = 0; i < 3; ++i) Iw[i] = w[0] * P.inertia[3 * i] + w[1] * P.inertia[3 * i + 1] + w[2] * P.inertia[3 * i + 2];
  r[0] = M[0] - (w[1] * Iw[2] - w[2] * Iw[1]);
  r[1] = M[1] - (w[2] * Iw[0] - w[0] * Iw[2]);
  r[2] = M[2] - (w[0] * Iw[1] - w[1] * Iw[0]);
  for (int i = 0; i < 3; ++i)                              // base.py:376-377
    xd[10 + i] = r[0] * P.inertia_inv[3 * i] + r[1] * P.inertia_inv[3 * i + 1] + r[2] * P.inertia_inv[3 * i + 2];
}

template <class T> void normalise_q(T x[13]) {
  const T n = Sqrt(x[6] * x[6] + x[7] * x[7] + x[8] * x[8] + x[9] * x[9]);
  for (int i = 6; i < 10; ++i) x[i] = x[i] / n;
}

// state_step, base.py:408-446
template <class T> void rk4(const Model& m, T x[13], const T u[7], const T& h, bool normalise) {
  T k1[13], k2[13], k3[13], k4[13], s[13];
  f(m, x, u, k1);
  for (int i = 0; i < 13; ++i) s[i] = x[i] + h * 0.5 * k1[i];
  f(m, s, u, k2);
  for (int i = 0; i < 13; ++i) s[i] = x[i] + h * 0.5 * k2[i];
  f(m, s, u, k3);
  for (int i = 0; i < 13; ++i) s[i] = x[i] + h * k3[i];
  f(m, s, u, k4);
  for (int i = 0; i < 13; ++i) x[i] = x[i] + h / 6.0 * (k1[i] + 2.0 * k2[i] + 2.0 * k3[i] + k4[i]);
  if (normalise) normalise_q(x);
}

// state_update, base.py:450-480
template <class T> void phi(const Model& m, T x[13], const T u[7], const T& dt, int substeps, bool normalise) {
  if (substeps == 1) { rk4(m, x, u, dt, normalise); return; }
  const T h = dt / (double)substeps;
  for (int s = 0; s < substeps; ++s) rk4(m, x, u, h, false);
  if (normalise) normalise_q(x);
}

// ---------------------------------------------------------------- primary MPC loss
// main/control/control.py:35-70 restated for one trajectory X[H+1][13], U[H][7]:
//   1000|p_N[:2]-goal|^2 + 100 sum_{k,i}(1-exp(-du^2/1e-2)) + (z_N-z_0)^2
//   - sum_{k<N} |v_b(x_k)|^2/(100 N) + vel_param*1000*vx_N + 1000(vy_N^2+vz_N^2)
double body_speed2(const Model& m, const double x[13]) {
  const double* v = x + 3; const double* q = x + 6;
  double qi[4], vq[4] = {v[0], v[1], v[2], 0.0}, t1[4], t2[4];
  qinv(q, qi); qmul(qi, vq, t1); qmul(t1, q, t2);          // base.py:156-159
  double acc = 0.0;
  for (int i = 0; i < 3; ++i) { const double c = t2[i] + m.p.epsilon; acc += c * c; }
  return acc;
}

struct CostAcc {
  double sum_du = 0.0, sum_v2 = 0.0, z0 = 0.0;
  double uprev[7];
  void node(const Model& m, const AircCostParams& cp, const double x[13], const double u[7], int k) {
    if (k == 0) z0 = x[2];
    else for (int i = 0; i < 7; ++i) { const double d = u[i] - uprev[i]; sum_du += 1.0 - exp(-d * d / (double)cp.du_eps); }
    for (int i = 0; i < 7; ++i) uprev[i] = u[i];
    sum_v2 += body_speed2(m, x);
  }
  double finish(const AircCostParams& cp, const double xN[13], int H) const {
    const double ex = xN[0] - cp.goal[0], ey = xN[1] - cp.goal[1], ez = xN[2] - z0;
    return cp.w_goal * (ex * ex + ey * ey) + cp.w_du * sum_du + cp.w_height * ez * ez -
           (H > 0 ? cp.w_speed / H : 0.0) * sum_v2 + cp.vel_param * cp.w_vx * xN[3] +
           cp.w_vyz * (xN[4] * xN[4] + xN[5] * xN[5]);
  }
};

// ---------------------------------------------------------------- sweep sampler
// The counter-based generator specified in include/aircraft_b200.h (AircSweepParams), restated.
// Control values are FP32 by specification (the product generates them in FP32 registers).
inline uint32_t mix32(uint32_t x) {
  x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16;
  return x;
}
struct SweepGen {
  float s0[4], s[4], lo[4], hi[4];
  uint32_t seed_lo_mixed, seed_hi;
  explicit SweepGen(const AircSweepParams& sp) {
    for (int c = 0; c < 4; ++c) {
      s0[c] = (float)((double)sp.sigma0[c] / 147.80053);
      s[c] = (float)((double)sp.sigma[c] / 147.80053);
      lo[c] = sp.lo[c]; hi[c] = sp.hi[c];
    }
    seed_lo_mixed = mix32((uint32_t)(sp.seed & 0xFFFFFFFFull));
    seed_hi = (uint32_t)(sp.seed >> 32);
  }
  uint32_t tkey(uint64_t g) const { return mix32((uint32_t)g + seed_lo_mixed) ^ seed_hi; }
  // advances the walk offsets d to step k, writes the 7 controls of that step
  __attribute__((optimize("fp-contract=off")))
  void step(uint32_t tk, int k, const float* unom_k, float d[4], float u[7]) const {
    static const int chan[4] = {0, 1, 2, 6};
    for (int c = 0; c < 7; ++c) u[c] = unom_k[c];
    const uint32_t cnt = tk + (uint32_t)k * 0x9E3779B9U;
    for (int c = 0; c < 4; ++c) {
      const uint32_t w = mix32(cnt + (uint32_t)c * 0x7F4A7C15U);
      const float n = (float)((int)((w & 255U) + ((w >> 8) & 255U) + ((w >> 16) & 255U) + (w >> 24)) - 510);
      d[c] = (k == 0) ? s0[c] * n : fmaf(s[c], n, d[c]);
      const float t = unom_k[chan[c]] + d[c];
      u[chan[c]] = (t != t) ? t : fminf(fmaxf(t, lo[c]), hi[c]);   // NaN in the nominal sequence stays NaN
    }
  }
};

int set_threads(int threads) {
#ifdef _OPENMP
  if (threads > 0) omp_set_num_threads(threads);
  return omp_get_max_threads();
#else
  (void)threads; return 1;
#endif
}

}  // namespace

// A[n][13][13], Bm[n][13][7] by 20-direction forward mode; a 21st direction carries d/d(dt)
// (the dt decision variable of the reference's variable-step transcriptions, control/base.py:255-286)
template <int ND>
static void linearize_nd(const AircParams* p, const AircNNBlob* nn, const double* x, const double* u,
                         double dt, int64_t n, int substeps, uint32_t flags, double* xn, double* A,
                         double* Bm, double* Dt, int threads) {
  Model m; make_model(p, nn, m); set_threads(threads);
  typedef Dual<ND> D;
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < n; ++i) {
    D s[13], c[7];
    for (int k = 0; k < 13; ++k) { s[k] = D(x[13 * i + k]); s[k].d[k] = 1.0; }
    for (int k = 0; k < 7; ++k) { c[k] = D(u[7 * i + k]); c[k].d[13 + k] = 1.0; }
    D h(dt);
    if (ND > 20) h.d[ND - 1] = 1.0;
    phi<D>(m, s, c, h, substeps, flags & AIRC_FLAG_NORMALISE);
    for (int r = 0; r < 13; ++r) {
      if (xn) xn[13 * i + r] = s[r].v;
      if (A) for (int k = 0; k < 13; ++k) A[(i * 13 + r) * 13 + k] = s[r].d[k];
      if (Bm) for (int k = 0; k < 7; ++k) Bm[(i * 13 + r) * 7 + k] = s[r].d[13 + k];
      if (ND > 20 && Dt) Dt[13 * i + r] = s[r].d[ND - 1];
    }
  }
}

extern "C" {

int cpuref_max_threads(void) {
#ifdef _OPENMP
  return omp_get_num_procs();
#else
  return 1;
#endif
}

void cpuref_state_derivative(const AircParams* p, const AircNNBlob* nn, const double* x, const double* u,
                             int64_t n, double* xdot, int threads) {
  Model m; make_model(p, nn, m); set_threads(threads);
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < n; ++i) f<double>(m, x + 13 * i, u + 7 * i, xdot + 13 * i);
}

void cpuref_state_update(const AircParams* p, const AircNNBlob* nn, const double* x, const double* u,
                         double dt, int64_t n, int substeps, uint32_t flags, double* xn, int threads) {
  Model m; make_model(p, nn, m); set_threads(threads);
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < n; ++i) {
    double s[13]; memcpy(s, x + 13 * i, sizeof(s));
    phi<double>(m, s, u + 7 * i, dt, substeps, flags & AIRC_FLAG_NORMALISE);
    memcpy(xn + 13 * i, s, sizeof(s));
  }
}

// x0[B][13], U[B][H][7] (or [H][7] with AIRC_FLAG_SHARED_U), X[B][H+1][13] or NULL, xf[B][13] or NULL
void cpuref_rollout(const AircParams* p, const AircNNBlob* nn, const double* x0, const double* U,
                    double dt, int64_t B, int H, int substeps, uint32_t flags, double* X, double* xf,
                    int threads) {
  Model m; make_model(p, nn, m); set_threads(threads);
#pragma omp parallel for schedule(static)
  for (int64_t b = 0; b < B; ++b) {
    double s[13]; memcpy(s, x0 + 13 * b, sizeof(s));
    const double* Ub = (flags & AIRC_FLAG_SHARED_U) ? U : U + (int64_t)7 * H * b;
    if (X) memcpy(X + (int64_t)13 * (H + 1) * b, s, sizeof(s));
    for (int k = 0; k < H; ++k) {
      phi<double>(m, s, Ub + 7 * k, dt, substeps, flags & AIRC_FLAG_NORMALISE);
      if (X) memcpy(X + ((int64_t)(H + 1) * b + k + 1) * 13, s, sizeof(s));
    }
    if (xf) memcpy(xf + 13 * b, s, sizeof(s));
  }
}

void cpuref_linearize(const AircParams* p, const AircNNBlob* nn, const double* x, const double* u,
                      double dt, int64_t n, int substeps, uint32_t flags, double* xn, double* A,
                      double* Bm, int threads) {
  linearize_nd<20>(p, nn, x, u, dt, n, substeps, flags, xn, A, Bm, nullptr, threads);
}

void cpuref_linearize_dt(const AircParams* p, const AircNNBlob* nn, const double* x, const double* u,
                         double dt, int64_t n, int substeps, uint32_t flags, double* xn, double* A,
                         double* Bm, double* Dt, int threads) {
  linearize_nd<21>(p, nn, x, u, dt, n, substeps, flags, xn, A, Bm, Dt, threads);
}

// costs of explicit control sequences: x0[B][13], U[B][H][7] (or shared [H][7]) -> cost[B]
void cpuref_rollout_cost(const AircParams* p, const AircNNBlob* nn, const double* x0, const double* U,
                         double dt, int64_t B, int H, int substeps, uint32_t flags,
                         const AircCostParams* cp, double* cost, int threads) {
  Model m; make_model(p, nn, m); set_threads(threads);
#pragma omp parallel for schedule(static)
  for (int64_t b = 0; b < B; ++b) {
    double s[13]; memcpy(s, x0 + 13 * b, sizeof(s));
    const double* Ub = (flags & AIRC_FLAG_SHARED_U) ? U : U + (int64_t)7 * H * b;
    CostAcc acc;
    for (int k = 0; k < H; ++k) {
      acc.node(m, *cp, s, Ub + 7 * k, k);
      phi<double>(m, s, Ub + 7 * k, dt, substeps, flags & AIRC_FLAG_NORMALISE);
    }
    if (H == 0) acc.z0 = s[2];
    cost[b] = acc.finish(*cp, s, H);
  }
}

// the sampled control sequences of trajectories [first, first+count): U[count][H][7] FP32
void cpuref_sweep_controls(const AircSweepParams* sp, const float* Unom, int H, int64_t first,
                           int64_t count, float* U, int threads) {
  const SweepGen gen(*sp); set_threads(threads);
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < count; ++i) {
    const uint32_t tk = gen.tkey(sp->index_offset + (uint64_t)(first + i));
    float d[4] = {0.f, 0.f, 0.f, 0.f};
    for (int k = 0; k < H; ++k) gen.step(tk, k, Unom + 7 * k, d, U + ((int64_t)i * H + k) * 7);
  }
}

// The whole sampling-MPC sweep on the CPU: sample, roll out (float64), cost, (min, arg-min).
// x0[13], Unom[H][7] (FP32, as the product receives it); cost[B] optional.  Returns the arg-min
// (local index) through best_idx and the minimum through best_cost.
void cpuref_sweep(const AircParams* p, const AircNNBlob* nn, const double* x0, const float* Unom,
                  double dt, int64_t B, int H, int substeps, uint32_t flags, const AircSweepParams* sp,
                  const AircCostParams* cp, double* cost, double* best_cost, int64_t* best_idx,
                  int threads) {
  Model m; make_model(p, nn, m); set_threads(threads);
  const SweepGen gen(*sp);
  double gbest = INFINITY; int64_t gidx = -1;
#pragma omp parallel
  {
    double lbest = INFINITY; int64_t lidx = -1;
#pragma omp for schedule(static) nowait
    for (int64_t b = 0; b < B; ++b) {
      double s[13]; memcpy(s, x0, sizeof(s));
      const uint32_t tk = gen.tkey(sp->index_offset + (uint64_t)b);
      float d[4] = {0.f, 0.f, 0.f, 0.f}, uf[7];
      double u[7];
      CostAcc acc;
      for (int k = 0; k < H; ++k) {
        gen.step(tk, k, Unom + 7 * k, d, uf);
        for (int c = 0; c < 7; ++c) u[c] = uf[c];
        acc.node(m, *cp, s, u, k);
        phi<double>(m, s, u, dt, substeps, flags & AIRC_FLAG_NORMALISE);
      }
      if (H == 0) acc.z0 = s[2];
      const double J = acc.finish(*cp, s, H);
      if (cost) cost[b] = J;
      if (J < lbest) { lbest = J; lidx = b; }
    }
#pragma omp critical
    if (lidx >= 0 && (lbest < gbest || (lbest == gbest && lidx < gidx))) { gbest = lbest; gidx = lidx; }
  }
  if (best_cost) *best_cost = gbest;
  if (best_idx) *best_idx = gidx;
}

}  // extern "C"

// airc_api.cu -- the C ABI of include/aircraft_b200.h over the kernels in airc_kernels.cuh.
// Built for sm_100a only (see __graft_entry__.build / Makefile).
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <new>

#include "../../include/aircraft_b200.h"
#include "airc_fuse.hpp"
#include "airc_kernels.cuh"
#include "airc_kernels2.cuh"
#include "airc_jac_kernel.cuh"
#include "airc_hess.cuh"
#include "airc_hess2.cuh"
#include "airc_peak.cuh"
#include "airc_sweep.cuh"
#include "airc_sweep2.cuh"
#include "airc_sweep2tc.cuh"

using namespace airc;

struct AircHandle {
  DevConst dc;
  UnitNet un;          // the surrogate in unit-major order for the packed (two-trajectory) kernels
  TcNetImage tcn;      // the surrogate as split-TF32 operand matrices for the tensor-core sweep (airc_tc.cuh)
  int packed = 1;      // 1: sweep on the packed kernels, rollout picks by batch size; 0: table models (scalar kernels only)
  int device;
  int sm_count;
  // used only by the *_host entry points (those are not re-entrant per handle)
  void* scratch = nullptr;
  size_t scratch_bytes = 0;
  cudaStream_t hstream = nullptr;
};

static thread_local char g_err[512] = "";

static int fail(int code, const char* fmt, const char* detail) {
  snprintf(g_err, sizeof(g_err), fmt, detail ? detail : "");
  return code;
}

#define CUDA_TRY(expr)                                                        \
  do {                                                                        \
    cudaError_t e__ = (expr);                                                 \
    if (e__ != cudaSuccess) return fail(AIRC_ECUDA, #expr ": %s", cudaGetErrorString(e__)); \
  } while (0)

static int check_launch() {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail(AIRC_ECUDA, "kernel launch: %s", cudaGetErrorString(e));
  return AIRC_OK;
}

// RAII device switch so that calls work whatever device is current in the caller's thread
struct DeviceGuard {
  int prev = -1;
  bool ok = true;
  explicit DeviceGuard(int dev) {
    if (cudaGetDevice(&prev) != cudaSuccess) { ok = false; return; }
    if (prev != dev && cudaSetDevice(dev) != cudaSuccess) ok = false;
  }
  ~DeviceGuard() {
    int cur;
    if (prev >= 0 && cudaGetDevice(&cur) == cudaSuccess && cur != prev) cudaSetDevice(prev);
  }
};

static VecView view(const float* base, long long n, int dim, int layout) {
  return layout == AIRC_LAYOUT_SOA ? VecView{base, 1, n} : VecView{base, (long long)dim, 1};
}
static VecViewW viewW(float* base, long long n, int dim, int layout) {
  return layout == AIRC_LAYOUT_SOA ? VecViewW{base, 1, n} : VecViewW{base, (long long)dim, 1};
}

extern "C" {

int airc_version(void) { return 100; }

const char* airc_build_info(void) {
  return "aircraft_b200 0.1.0 sm_100a (" __DATE__ " " __TIME__ ") kernels: state_derivative,"
         " state_update, rollout, sweep, linearize, argmin";
}

const char* airc_last_error(const AircHandle*) { return g_err; }

int airc_create(const AircParams* params, const AircNNBlob* nn, int device, AircHandle** out) {
  if (!params || !nn || !out) return fail(AIRC_EINVAL, "airc_create: NULL argument%s", "");
  if (!(params->mass > 0.0) || !(params->reference_area > 0.0))
    return fail(AIRC_EINVAL, "airc_create: mass and reference_area must be positive%s", "");
  for (int i = 0; i < 5; ++i)
    if (!(nn->input_std[i] != 0.0f)) return fail(AIRC_EINVAL, "airc_create: input_std has a zero%s", "");
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0)
    return fail(AIRC_ENODEV, "airc_create: no CUDA device (%s)", cudaGetErrorString(e));
  if (device < 0 || device >= count) return fail(AIRC_EINVAL, "airc_create: bad device index%s", "");
  cudaDeviceProp prop;
  CUDA_TRY(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10)
    return fail(AIRC_ENODEV, "airc_create: device is not sm_100 (%s); this library has no other code path", prop.name);
  AircHandle* h = new (std::nothrow) AircHandle;
  if (!h) return fail(AIRC_ENOMEM, "airc_create: out of host memory%s", "");
  build_devconst(*params, *nn, h->dc);
  make_unit_net(h->dc.net, h->un);
  make_tc_net(h->un, h->tcn);
#ifdef AIRC_AB_ROLLOUT_IMPL        // build-variant A/B only (profiles/build_variants.py): 0 scalar only, 2 packed always
  h->packed = AIRC_AB_ROLLOUT_IMPL;
#endif
  h->device = device;
  h->sm_count = prop.multiProcessorCount;
  *out = h;
  return AIRC_OK;
}

int airc_create_model(const AircParams* params, int model, const float* model_params, int n_model_params,
                      int device, AircHandle** out) {
  if (!params || !out) return fail(AIRC_EINVAL, "airc_create_model: NULL argument%s", "");
  const int want = model == AIRC_MODEL_POLY ? 6 * (NPOLY + 1) + 1 : (model == AIRC_MODEL_LINEAR ? 36 : 0);
  if ((model != AIRC_MODEL_POLY && model != AIRC_MODEL_LINEAR && model != AIRC_MODEL_DEFAULT) ||
      n_model_params != want || (want && !model_params))
    return fail(AIRC_EINVAL, "airc_create_model: unknown model or wrong parameter count%s", "");
  static_assert(AIRC_MODEL_POLY == AIRC_MODEL_POLY_ && AIRC_MODEL_LINEAR == AIRC_MODEL_LINEAR_ &&
                AIRC_MODEL_DEFAULT == AIRC_MODEL_DEFAULT_, "header and kernels disagree on the model ids");
  AircNNBlob* dummy = new (std::nothrow) AircNNBlob;   // airc_create wants a network; its storage is then reused
  if (!dummy) return fail(AIRC_ENOMEM, "airc_create_model: out of host memory%s", "");
  memset(dummy, 0, sizeof(*dummy));
  for (int i = 0; i < 5; ++i) dummy->input_std[i] = 1.0f;
  for (int i = 0; i < 6; ++i) dummy->output_std[i] = 1.0f;
  AircHandle* h = nullptr;
  const int rc = airc_create(params, dummy, device, &h);
  delete dummy;
  if (rc) return rc;
  build_devconst_model(*params, model, model_params, h->dc);
  h->packed = 0;
  *out = h;
  return AIRC_OK;
}

int airc_destroy(AircHandle* h) {
  if (!h) return AIRC_OK;
  {
    DeviceGuard g(h->device);
    if (h->scratch) cudaFree(h->scratch);
    if (h->hstream) cudaStreamDestroy(h->hstream);
  }
  delete h;
  return AIRC_OK;
}

int airc_state_derivative(const AircHandle* h, const float* x, const float* u, int64_t n,
                          int layout, float* xdot, void* stream) {
  if (!h || !x || !u || !xdot || n < 0 || (layout != 0 && layout != 1))
    return fail(AIRC_EINVAL, "airc_state_derivative: bad argument%s", "");
  if (n == 0) return AIRC_OK;
  DeviceGuard g(h->device);
  if (!g.ok) return fail(AIRC_ECUDA, "cudaSetDevice failed%s", "");
  NodeArgs a{view(x, n, 13, layout), view(u, n, 7, layout), viewW(xdot, n, 13, layout), nullptr, 0.f, n, 1, 0};
  const int bs = 128;
  if (h->dc.model) state_derivative_kernel<1><<<(unsigned)((n + bs - 1) / bs), bs, 0, (cudaStream_t)stream>>>(h->dc, a);
  else state_derivative_kernel<0><<<(unsigned)((n + bs - 1) / bs), bs, 0, (cudaStream_t)stream>>>(h->dc, a);
  return check_launch();
}

int airc_diagnostics(const AircHandle* h, const float* x, const float* u, int64_t n, int layout,
                     float* out, void* stream) {
  if (!h || !x || !u || !out || n < 0 || (layout != 0 && layout != 1))
    return fail(AIRC_EINVAL, "airc_diagnostics: bad argument%s", "");
  if (n == 0) return AIRC_OK;
  DeviceGuard g(h->device);
  if (!g.ok) return fail(AIRC_ECUDA, "cudaSetDevice failed%s", "");
  NodeArgs a{view(x, n, 13, layout), view(u, n, 7, layout), viewW(out, n, AIRC_NDIAG, layout), nullptr, 0.f, n, 1, 0};
  static_assert(AIRC_NDIAG == NDIAG, "header and kernel disagree on the diagnostics record");
  const int bs = 128;
  if (h->dc.model) diagnostics_kernel<1><<<(unsigned)((n + bs - 1) / bs), bs, 0, (cudaStream_t)stream>>>(h->dc, a);
  else diagnostics_kernel<0><<<(unsigned)((n + bs - 1) / bs), bs, 0, (cudaStream_t)stream>>>(h->dc, a);
  return check_launch();
}

int airc_state_update(const AircHandle* h, const float* x, const float* u, float dt,
                      const float* dt_per_node, int64_t n, int substeps, uint32_t flags,
                      int layout, float* x_next, void* stream) {
  if (!h || !x || !u || !x_next || n < 0 || substeps < 1 || (layout != 0 && layout != 1) ||
      (flags & ~(uint32_t)AIRC_FLAG_NORMALISE))
    return fail(AIRC_EINVAL, "airc_state_update: bad argument%s", "");
  if (n == 0) return AIRC_OK;
  DeviceGuard g(h->device);
  if (!g.ok) return fail(AIRC_ECUDA, "cudaSetDevice failed%s", "");
  NodeArgs a{view(x, n, 13, layout), view(u, n, 7, layout), viewW(x_next, n, 13, layout),
             dt_per_node, dt, n, substeps, (int)(flags & AIRC_FLAG_NORMALISE)};
  const int bs = (n <= 148 * 64) ? 32 : 128;
  if (h->dc.model) state_update_kernel<1><<<(unsigned)((n + bs - 1) / bs), bs, 0, (cudaStream_t)stream>>>(h->dc, a);
  else state_update_kernel<0><<<(unsigned)((n + bs - 1) / bs), bs, 0, (cudaStream_t)stream>>>(h->dc, a);
  return check_launch();
}

int airc_rollout(const AircHandle* h, const float* x0, const float* U, float dt, int64_t B,
                 int H, int substeps, uint32_t flags, int layout, float* X, float* x_final,
                 float* cost, const AircCostParams* cp, void* stream) {
  if (!h || B < 0 || H < 0 || substeps < 1 || (layout != 0 && layout != 1) ||
      (flags & ~(uint32_t)(AIRC_FLAG_NORMALISE | AIRC_FLAG_SHARED_U | AIRC_FLAG_COOPERATIVE | AIRC_FLAG_PACKED)) || (cost && !cp))
    return fail(AIRC_EINVAL, "airc_rollout: bad argument%s", "");
  if (B == 0) return AIRC_OK;   // empty batch: nothing to read, so NULL data pointers are fine
  if (!x0 || (!U && H > 0)) return fail(AIRC_EINVAL, "airc_rollout: NULL x0 or U%s", "");
  DeviceGuard g(h->device);
  if (!g.ok) return fail(AIRC_ECUDA, "cudaSetDevice failed%s", "");
  RolloutArgs a;
  memset(&a, 0, sizeof(a));
  a.x0 = view(x0, B, 13, layout);
  a.U = U;
  if (flags & AIRC_FLAG_SHARED_U) { a.u_sb = 0; a.u_sk = 7; a.u_sc = 1; }
  else if (layout == AIRC_LAYOUT_SOA) { a.u_sb = 1; a.u_sk = 7 * B; a.u_sc = B; }
  else { a.u_sb = 7LL * H; a.u_sk = 7; a.u_sc = 1; }
  a.X = X;
  if (layout == AIRC_LAYOUT_SOA) { a.X_sb = 1; a.X_sk = 13 * B; a.X_sc = B; }
  else { a.X_sb = 13LL * (H + 1); a.X_sk = 13; a.X_sc = 1; }
  a.xf = viewW(x_final, B, 13, layout);
  a.cost = cost;
  if (cost) {
    a.cc = CostConst{cp->goal[0], cp->goal[1], cp->vel_param, cp->w_goal, cp->w_du,
                     (float)(-1.4426950408889634 / (double)cp->du_eps), cp->w_height,
                     H > 0 ? cp->w_speed / (float)H : 0.f, cp->w_vx, cp->w_vyz};
  }
  a.dt = dt; a.B = B; a.H = H; a.substeps = substeps;
  a.normalise = (int)(flags & AIRC_FLAG_NORMALISE);
  // small batches: one warp per block so that every SM gets work
  const int bs = (B <= (int64_t)h->sm_count * 32 * 2) ? 32 : ((B <= (int64_t)h->sm_count * 64 * 4) ? 64 : AIRC_ROLLOUT_BS);
  const unsigned grid = (unsigned)((B + bs - 1) / bs);
  cudaStream_t st = (cudaStream_t)stream;
  // Explicit-control rollouts: the packed kernel (two trajectories per thread, 12 warps/SM) wins once
  // the batch fills at least two of its waves (2 x 113 664 on a B200); below that the one-trajectory
  // kernel's finer wave granularity wins (profiles/r1_variants_packed.txt).  AIRC_ROLLOUT_IMPL=1|2
  // forces one of them (A/B timing).
  const long long wave2 = (long long)h->sm_count * 3 * AIRC_R2_BS * 2;
  if (h->dc.model) {   // poly / linear / default coefficient models: one-trajectory kernel only
    if (cost && X) rollout_kernel<true, true, 1><<<grid, bs, 0, st>>>(h->dc, a);
    else if (cost) rollout_kernel<true, false, 1><<<grid, bs, 0, st>>>(h->dc, a);
    else if (X) rollout_kernel<false, true, 1><<<grid, bs, 0, st>>>(h->dc, a);
    else rollout_kernel<false, false, 1><<<grid, bs, 0, st>>>(h->dc, a);
    return check_launch();
  }
  // small batches (latency case, configs 1 and 2): 8, 4 or 2 lanes share one trajectory (airc_coop.cuh) -- as many as
  // keep the batch at ONE warp per SM sub-partition (B <= 16 / 32 / 64 x SM count): the launch is bound by each warp's
  // dependency chain, and a second warp on a scheduler costs more than the longer chain of fewer lanes.  The default
  // while the batch leaves sub-partitions idle on the one-trajectory kernel (B <= 32 x SM count);
  // AIRC_FLAG_COOPERATIVE extends it to 64 x SM count (profiles/r2_variants_latency.txt).
  if (h->packed == 1 && !(flags & AIRC_FLAG_PACKED) &&
      (B <= (long long)h->sm_count * 32 || ((flags & AIRC_FLAG_COOPERATIVE) && B <= (long long)h->sm_count * 64))) {
#ifndef AIRC_AB_COOP_LANES          // build-variant A/B only: 0 = lanes by batch size (shipped), 2 / 4 / 8 = fixed
#define AIRC_AB_COOP_LANES 0
#endif
    const long long slots = (long long)h->sm_count * 128;          // threads of one warp per sub-partition
    const int Lc = AIRC_AB_COOP_LANES ? AIRC_AB_COOP_LANES : (B * 8 <= slots ? 8 : (B * 4 <= slots ? 4 : 2));
    const unsigned gridc = (unsigned)((B * Lc + 127) / 128);
#define AIRC_COOP_LAUNCH(LL)                                                                             \
    do {                                                                                                 \
      if (cost && X) rollout_coop_kernel<true, true, LL><<<gridc, 128, 0, st>>>(h->dc, a);              \
      else if (cost) rollout_coop_kernel<true, false, LL><<<gridc, 128, 0, st>>>(h->dc, a);             \
      else if (X) rollout_coop_kernel<false, true, LL><<<gridc, 128, 0, st>>>(h->dc, a);                \
      else rollout_coop_kernel<false, false, LL><<<gridc, 128, 0, st>>>(h->dc, a);                      \
    } while (0)
    if (Lc == 8) AIRC_COOP_LAUNCH(8);
    else if (Lc == 4) AIRC_COOP_LAUNCH(4);
    else AIRC_COOP_LAUNCH(2);
#undef AIRC_COOP_LAUNCH
    return check_launch();
  }
  if (h->packed == 2 || (h->packed == 1 && ((flags & AIRC_FLAG_PACKED) || B >= 2 * wave2))) {
    const unsigned grid2 = (unsigned)((B + 2 * AIRC_R2_BS - 1) / (2 * AIRC_R2_BS));
    if (cost && X) rollout2_kernel<true, true><<<grid2, AIRC_R2_BS, 0, st>>>(h->dc, h->un, a);
    else if (cost) rollout2_kernel<true, false><<<grid2, AIRC_R2_BS, 0, st>>>(h->dc, h->un, a);
    else if (X) rollout2_kernel<false, true><<<grid2, AIRC_R2_BS, 0, st>>>(h->dc, h->un, a);
    else rollout2_kernel<false, false><<<grid2, AIRC_R2_BS, 0, st>>>(h->dc, h->un, a);
    return check_launch();
  }
  if (cost && X) rollout_kernel<true, true><<<grid, bs, 0, st>>>(h->dc, a);
  else if (cost) rollout_kernel<true, false><<<grid, bs, 0, st>>>(h->dc, a);
  else if (X) rollout_kernel<false, true><<<grid, bs, 0, st>>>(h->dc, a);
  else rollout_kernel<false, false><<<grid, bs, 0, st>>>(h->dc, a);
  return check_launch();
}

static int linearize_impl(const AircHandle* h, const float* x, const float* u, float dt,
                          const float* dt_per_node, int64_t n, int substeps, uint32_t flags, int layout,
                          float* x_next, float* A, float* Bm, float* Dt, void* stream,
                          const float* x_target = nullptr, float* defect = nullptr) {
  if (!h || !x || !u || n < 0 || substeps < 1 || (layout != 0 && layout != 1) ||
      (flags & ~(uint32_t)AIRC_FLAG_NORMALISE))
    return fail(AIRC_EINVAL, "airc_linearize: bad argument%s", "");
  if (n == 0) return AIRC_OK;
  DeviceGuard g(h->device);
  if (!g.ok) return fail(AIRC_ECUDA, "cudaSetDevice failed%s", "");
  LinArgs a;
  memset(&a, 0, sizeof(a));
  a.x = view(x, n, 13, layout);
  a.u = view(u, n, 7, layout);
  a.xn = viewW(x_next, n, 13, layout);
  a.A = A; a.Bm = Bm; a.Dt = Dt;
  if (defect) { a.xt = view(x_target, n, 13, layout); a.r = viewW(defect, n, 13, layout); }
  if (layout == AIRC_LAYOUT_SOA) { a.A_sn = 1; a.A_se = n; a.B_sn = 1; a.B_se = n; }
  else { a.A_sn = 169; a.A_se = 1; a.B_sn = 91; a.B_se = 1; }
  a.dt_per_node = dt_per_node; a.dt = dt; a.n = n; a.substeps = substeps;
  a.normalise = (int)(flags & AIRC_FLAG_NORMALISE);
  const int bs = 128;
  if (h->dc.model) {   // poly / linear / default: two-column forward mode on jets (airc_hess.cuh)
    const unsigned gridj = (unsigned)((n * 8 + bs - 1) / bs);
    if (layout == AIRC_LAYOUT_SOA) linearize_jet_kernel<true, 1><<<gridj, bs, 0, (cudaStream_t)stream>>>(h->dc, a);
    else linearize_jet_kernel<false, 1><<<gridj, bs, 0, (cudaStream_t)stream>>>(h->dc, a);
    return check_launch();
  }
  cudaStream_t st = (cudaStream_t)stream;
  if (substeps == 1) {
    // K2, structured form (airc_jac.cuh): one launch configuration per (tile size, layout, dt column)
    const bool soa = layout == AIRC_LAYOUT_SOA, want_dt = Dt != nullptr;
    auto al16 = [](const void* p) { return p == nullptr || ((uintptr_t)p & 15u) == 0; };
    const int vec_ok = (al16(A) && al16(Bm) && al16(Dt) && (!soa || (n & 3) == 0)) ? 1 : 0;
    // Tile size (profiles/r2_variants_linearize.txt): 64 nodes per block, two blocks per SM -- one block's tangent
    // recursion and write-out overlap the other's assembly phase; AoS without the d/d(dt) column runs 32-node tiles
    // (four blocks per SM; in SoA their 32-byte row segments cost more than the extra overlap gives).
#define AIRC_JAC_LAUNCH(NT_, SOA_, DT_)                                                                     \
    do {                                                                                                    \
      using T_ = JacTile<NT_, SOA_, DT_>;                                                                   \
      CUDA_TRY(cudaFuncSetAttribute(linearize_tile_kernel<NT_, SOA_, DT_>,                                 \
                                    cudaFuncAttributeMaxDynamicSharedMemorySize, (int)T_::SMEM));          \
      linearize_tile_kernel<NT_, SOA_, DT_><<<(unsigned)((n + NT_ - 1) / NT_), T_::TPB, T_::SMEM, st>>>(h->dc, a, vec_ok); \
    } while (0)
#define AIRC_JAC_PICK(NT_)                                                                                  \
    do {                                                                                                    \
      if (soa && want_dt) AIRC_JAC_LAUNCH(NT_, true, true);                                                 \
      else if (soa) AIRC_JAC_LAUNCH(NT_, true, false);                                                      \
      else if (want_dt) AIRC_JAC_LAUNCH(NT_, false, true);                                                  \
      else AIRC_JAC_LAUNCH(NT_, false, false);                                                              \
    } while (0)
#if defined(AIRC_JAC_FORCE_NT)       // build-variant A/B only (profiles/build_variants.py); never set in the shipped build
    AIRC_JAC_PICK(AIRC_JAC_FORCE_NT);
#else
    if (!soa && !want_dt) AIRC_JAC_LAUNCH(32, false, false);
    else AIRC_JAC_PICK(64);
#endif
#undef AIRC_JAC_PICK
#undef AIRC_JAC_LAUNCH
    return check_launch();
  }
  // substeps > 1: the cooperative kernel (8 lanes per node, tangents pushed through f; airc_coop.cuh)
  const long long threads = n * 8;
  const unsigned grid = (unsigned)((threads + bs - 1) / bs);
  if (layout == AIRC_LAYOUT_SOA) linearize_coop_kernel<8, true><<<grid, bs, 0, st>>>(h->dc, a);
  else linearize_coop_kernel<8, false><<<grid, bs, 0, st>>>(h->dc, a);
  return check_launch();
}

int airc_linearize(const AircHandle* h, const float* x, const float* u, float dt,
                   const float* dt_per_node, int64_t n, int substeps, uint32_t flags, int layout,
                   float* x_next, float* A, float* Bm, void* stream) {
  return linearize_impl(h, x, u, dt, dt_per_node, n, substeps, flags, layout, x_next, A, Bm, nullptr, stream);
}

int airc_linearize_dt(const AircHandle* h, const float* x, const float* u, float dt,
                      const float* dt_per_node, int64_t n, int substeps, uint32_t flags, int layout,
                      float* x_next, float* A, float* Bm, float* dPhi_ddt, void* stream) {
  return linearize_impl(h, x, u, dt, dt_per_node, n, substeps, flags, layout, x_next, A, Bm, dPhi_ddt, stream);
}

int airc_defects(const AircHandle* h, const float* x, const float* u, const float* x_target, float dt,
                 const float* dt_per_node, int64_t n, int substeps, uint32_t flags, int layout,
                 float* defect, float* A, float* Bm, float* dPhi_ddt, void* stream) {
  if (!x_target || !defect) return fail(AIRC_EINVAL, "airc_defects: bad argument%s", "");
  return linearize_impl(h, x, u, dt, dt_per_node, n, substeps, flags, layout, nullptr, A, Bm, dPhi_ddt, stream,
                        x_target, defect);
}

int airc_hessian(const AircHandle* h, const float* x, const float* u, const float* lam, float dt,
                 const float* dt_per_node, int64_t n, int substeps, uint32_t flags, int layout,
                 float* H, float* grad, float* value, void* stream) {
  if (!h || !x || !u || !lam || !H || n < 0 || substeps < 1 || (layout != 0 && layout != 1) ||
      (flags & ~(uint32_t)AIRC_FLAG_NORMALISE))
    return fail(AIRC_EINVAL, "airc_hessian: bad argument%s", "");
  if (n == 0) return AIRC_OK;
  DeviceGuard g(h->device);
  if (!g.ok) return fail(AIRC_ECUDA, "cudaSetDevice failed%s", "");
  cudaStream_t st = (cudaStream_t)stream;
  // rows/columns of p and thrust are structurally zero; the kernel writes the 15x15 active block
  CUDA_TRY(cudaMemsetAsync(H, 0, (size_t)n * AIRC_NZ * AIRC_NZ * sizeof(float), st));
  HessArgs a;
  memset(&a, 0, sizeof(a));
  a.x = view(x, n, 13, layout);
  a.u = view(u, n, 7, layout);
  a.lam = view(lam, n, 13, layout);
  a.dt_per_node = dt_per_node; a.dt = dt; a.n = n; a.substeps = substeps;
  a.normalise = (int)(flags & AIRC_FLAG_NORMALISE);
  a.H = H; a.grad = grad; a.value = value;
  const int bs = 128;
#ifndef AIRC_AB_HESS_JETS          // build-variant A/B only: 1 keeps the network model on the second-order jets too
#define AIRC_AB_HESS_JETS 0
#endif
#ifndef AIRC_AB_HESS_SMALL         // build-variant A/B only: node count per SM up to which the jets kernel runs
#define AIRC_AB_HESS_SMALL 4
#endif
  if (!h->dc.model && substeps <= HESS2_MAXSUB && !AIRC_AB_HESS_JETS && n > (long long)AIRC_AB_HESS_SMALL * h->sm_count) {
    // the NN surrogate: forward over reverse, eight threads per node (airc_hess2.cuh).  Short horizons (up to
    // four nodes per SM) stay on the jets: 64 threads per node finish one node sooner than eight do.
    // One wave = 32 nodes per SM either way.  A wave of one 256-thread block per SM takes 0.88 of the time of a wave
    // of two 128-thread blocks (the warps share the fetch of the straight-line code), but a straggling 256-thread
    // block costs a whole wave time where 128-thread blocks refill half an SM at a time: the larger block when its
    // whole waves are cheaper than the smaller block's fractional ones (profiles/r2_variants_hessian.txt).
    const double waves = (double)n / (32.0 * h->sm_count);
    const int bs2 = AIRC_AB_H2_BS ? AIRC_AB_H2_BS : (ceil(waves) * 0.88 <= waves + 0.1 ? 256 : 128);
    const unsigned grid = (unsigned)((n * HESS2_LANES + bs2 - 1) / bs2);
#define AIRC_H2_LAUNCH(BS_)                                                                   \
    do {                                                                                      \
      if (layout == AIRC_LAYOUT_SOA) hessian2_kernel<true, BS_><<<grid, BS_, 0, st>>>(h->dc, a); \
      else hessian2_kernel<false, BS_><<<grid, BS_, 0, st>>>(h->dc, a);                       \
    } while (0)
    if (bs2 == 256) AIRC_H2_LAUNCH(256);
#if AIRC_AB_H2_BS && AIRC_AB_H2_BS != 128 && AIRC_AB_H2_BS != 256
    else if (bs2 == AIRC_AB_H2_BS) AIRC_H2_LAUNCH(AIRC_AB_H2_BS);
#endif
    else AIRC_H2_LAUNCH(128);
#undef AIRC_H2_LAUNCH
    return check_launch();
  }
#ifndef AIRC_AB_HESS_SHARED        // build-variant A/B only: 0 keeps short horizons on the unshared jets kernel
#define AIRC_AB_HESS_SHARED 1
#endif
  if (!h->dc.model && AIRC_AB_HESS_SHARED && n <= (long long)AIRC_AB_HESS_SMALL * h->sm_count) {
    // short horizons of the network model: jets, one node per 64-thread block, the network shared by the block
    if (layout == AIRC_LAYOUT_SOA) hessian_kernel<true, 0, true><<<(unsigned)n, HESS_LANES, 0, st>>>(h->dc, a);
    else hessian_kernel<false, 0, true><<<(unsigned)n, HESS_LANES, 0, st>>>(h->dc, a);
    return check_launch();
  }
  // table models, and more substeps than the reverse sweep keeps: second-order jets, 64 threads per node
  const long long threads = n * HESS_LANES;
  const unsigned grid = (unsigned)((threads + bs - 1) / bs);
  if (h->dc.model) {
    if (layout == AIRC_LAYOUT_SOA) hessian_kernel<true, 1><<<grid, bs, 0, st>>>(h->dc, a);
    else hessian_kernel<false, 1><<<grid, bs, 0, st>>>(h->dc, a);
  } else if (layout == AIRC_LAYOUT_SOA) hessian_kernel<true, 0><<<grid, bs, 0, st>>>(h->dc, a);
  else hessian_kernel<false, 0><<<grid, bs, 0, st>>>(h->dc, a);
  return check_launch();
}

int airc_argmin(const AircHandle* h, const float* cost, int64_t B, float* out_min,
                int64_t* out_idx, void* stream) {
  if (!h || !cost || !out_idx || B <= 0 || B > 0xFFFFFFFFLL)
    return fail(AIRC_EINVAL, "airc_argmin: bad argument%s", "");
  DeviceGuard g(h->device);
  if (!g.ok) return fail(AIRC_ECUDA, "cudaSetDevice failed%s", "");
  cudaStream_t st = (cudaStream_t)stream;
  // out_idx doubles as the 64-bit key during the reduction
  CUDA_TRY(cudaMemsetAsync(out_idx, 0xFF, sizeof(int64_t), st));
  long long blocks = (B + 255) / 256;
  if (blocks > (long long)h->sm_count * 8) blocks = (long long)h->sm_count * 8;
  argmin_kernel<<<(unsigned)blocks, 256, 0, st>>>(cost, B, (unsigned long long*)out_idx);
  argmin_finish_kernel<<<1, 1, 0, st>>>((const unsigned long long*)out_idx, out_min, (long long*)out_idx);
  return check_launch();
}

// ------------------------------------------------------------------ sampling-MPC sweep
static const double kSamplerStd = 147.80053;   // sqrt(4*(256^2-1)/12): std of a sum of four bytes

static int ckpt_interval_of(int H, int requested) {
  int C = requested > 0 ? requested : 8;
  if (C > H) C = H > 0 ? H : 1;
  while ((H + C - 1) / C > 1024) ++C;      // one winner thread per segment, one block
  return C;
}

static CostConst make_cost_const(const AircCostParams* cp, int H) {
  return CostConst{cp->goal[0], cp->goal[1], cp->vel_param, cp->w_goal, cp->w_du,
                   (float)(-1.4426950408889634 / (double)cp->du_eps), cp->w_height,
                   H > 0 ? cp->w_speed / (float)H : 0.f, cp->w_vx, cp->w_vyz};
}

static Sampler make_sampler(const AircSweepParams* sp) {
  Sampler s;
  for (int c = 0; c < 4; ++c) {
    s.s0[c] = (float)((double)sp->sigma0[c] / kSamplerStd);
    s.s[c] = (float)((double)sp->sigma[c] / kSamplerStd);
    s.lo[c] = sp->lo[c];
    s.hi[c] = sp->hi[c];
  }
  s.seed_lo_mixed = mix32((unsigned int)(sp->seed & 0xFFFFFFFFull));
  s.seed_hi = (unsigned int)(sp->seed >> 32);
  s.index_offset = sp->index_offset;
  return s;
}

static size_t al256(size_t b) { return (b + 255) / 256 * 256; }

int64_t airc_sweep_record_floats(int H) {
  return H < 0 ? 0 : (int64_t)AIRC_SWEEP_RECORD_HEADER + 13LL * (H + 1) + 7LL * H;
}

struct SweepLayout {
  size_t off_ckpt, off_slice, off_flag, total;
  int C, NS;
  long long G64;
};
// workspace: [0,256) key | thrust_du | task counter ; checkpoints ; slice records ; group flags
static SweepLayout sweep_layout(int64_t B, int H, int ckpt_interval) {
  SweepLayout L;
  L.C = ckpt_interval_of(H, ckpt_interval);
  L.NS = H > 0 ? (H + L.C - 1) / L.C : 1;
  L.G64 = (B + 31) / 32;                     // flags are sized for the 32-trajectory grouping
  L.off_ckpt = 256;
  L.off_slice = L.off_ckpt + al256((size_t)(L.NS - 1) * 17 * (size_t)B * sizeof(float));
  L.off_flag = L.off_slice + al256((size_t)23 * (size_t)B * sizeof(float));
  L.total = L.off_flag + al256((size_t)L.G64 * sizeof(int));
  return L;
}

size_t airc_sweep_workspace_bytes(int64_t B, int H, int ckpt_interval) {
  if (B < 0 || H < 0) return 0;
  return sweep_layout(B, H, ckpt_interval).total;
}

// segments that minimise ceil(G S / W)/S (+ a small per-segment hand-over cost)
static int choose_segments(long long G, long long W, int H) {
  if (G <= W || H < 2) return 1;
  int best = 1;
  double be = 1e30;
  for (int S = 1; S <= 32 && S <= H; ++S) {
    const double e = (double)((G * S + W - 1) / W) / S + 0.003 * S;
    if (e < be - 1e-12) { be = e; best = S; }
  }
  return best;
}

int airc_sweep(const AircHandle* h, const float* x0, const float* U_nom, float dt, int64_t B, int H,
               int substeps, uint32_t flags, const AircSweepParams* sp, const AircCostParams* cp,
               void* workspace, size_t workspace_bytes, float* cost, float* record, void* stream) {
  if (!h || !x0 || (!U_nom && H > 0) || !sp || !cp || !record || B < 0 || B > 0xFFFFFFFFLL || H < 0 ||
      substeps < 1 || (flags & ~(uint32_t)AIRC_FLAG_NORMALISE))
    return fail(AIRC_EINVAL, "airc_sweep: bad argument%s", "");
  const int C = ckpt_interval_of(H, sp->ckpt_interval);
  const int S = H > 0 ? (H + C - 1) / C : 1;
  const size_t need_min = 256;
  if (!workspace || workspace_bytes < need_min)
    return fail(AIRC_EINVAL, "airc_sweep: workspace too small (see airc_sweep_workspace_bytes)%s", "");
  // checkpoints are used when the workspace has room for them; otherwise the winner is re-rolled
  // sequentially (same result, H steps of latency instead of C)
  const bool use_ckpt = S > 1 && workspace_bytes >= airc_sweep_workspace_bytes(B, H, C);
  const size_t ckpt_off = sweep_layout(B, H, sp->ckpt_interval).off_ckpt;
  DeviceGuard g(h->device);
  if (!g.ok) return fail(AIRC_ECUDA, "cudaSetDevice failed%s", "");
  cudaStream_t st = (cudaStream_t)stream;
  char* ws = (char*)workspace;
  unsigned long long* key = (unsigned long long*)ws;
  float* thrust_du = (float*)(ws + 8);
  float* ckpt = use_ckpt ? (float*)(ws + ckpt_off) : nullptr;
  const CostConst cc = make_cost_const(cp, H);
  const Sampler smp = make_sampler(sp);
  if (h->packed) {
    const SweepLayout L = sweep_layout(B, H, sp->ckpt_interval);
    const bool full = workspace_bytes >= L.total;
    const bool ck = full && L.NS > 1;
    int* flag = full ? (int*)(ws + L.off_flag) : nullptr;
    int per_sm = 0;
    if (ck) CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, sweep2_kernel<true>, AIRC_S2_BS, 0));
    else CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, sweep2_kernel<false>, AIRC_S2_BS, 0));
    const long long slots = (long long)h->sm_count * (per_sm > 0 ? per_sm : 1);    // resident blocks
    const long long W = slots * (AIRC_S2_BS / 32);                                 // resident warps
    Sweep2Args a;
    memset(&a, 0, sizeof(a));
    // small sweeps: 32 trajectories per warp (hi halves idle) so that twice as many warps share the
    // work, and one warp per block so that it spreads over all SMs
    a.gs = (B <= W * 32) ? 32 : 64;
    a.G = (int)((B + a.gs - 1) / a.gs);
    const int bs = (a.G <= slots) ? 32 : AIRC_S2_BS;
    a.GB = (a.G + bs / 32 - 1) / (bs / 32);
    a.S = full ? choose_segments(a.GB, bs == 32 ? slots * (AIRC_S2_BS / 32) : slots, H) : 1;
#ifdef AIRC_AB_SWEEP_SEGMENTS      // build-variant A/B only
    if (full) a.S = AIRC_AB_SWEEP_SEGMENTS;
#endif
    a.L = a.S > 1 ? (H + a.S - 1) / a.S : (H > 0 ? H : 1);
    a.S = H > 0 ? (H + a.L - 1) / a.L : 1;
    a.x0 = x0; a.Unom = U_nom; a.cost = cost; a.key = key; a.thrust_du = thrust_du;
    a.ckpt = ck ? (float*)(ws + L.off_ckpt) : nullptr;
    a.slice = full ? (float*)(ws + L.off_slice) : nullptr;
    a.flag = flag;
    a.cc = cc; a.smp = smp; a.dt = dt; a.B = B; a.H = H; a.C = L.C; a.substeps = substeps;
    a.normalise = (int)(flags & AIRC_FLAG_NORMALISE);
    // Large sweeps (more 128-thread tasks per segment than the machine holds at once): the surrogate's layers run
    // on the tensor cores, one 384-thread block = three 128-thread groups per SM (airc_sweep2tc.cuh).
#ifndef AIRC_AB_SWEEP_TC           // build-variant A/B only: 0 keeps every sweep on the FFMA2 kernel
#define AIRC_AB_SWEEP_TC 1
#endif
    const bool use_tc = AIRC_AB_SWEEP_TC && full && a.gs == 64 && bs == AIRC_S2_BS && B >= (long long)h->sm_count * 3 * 256;
    if (use_tc) {
      const long long slots_tc = (long long)h->sm_count * 3;
      a.GB = (a.G + 3) / 4;
      a.S = choose_segments(a.GB, slots_tc, H);
#ifdef AIRC_AB_SWEEP_SEGMENTS
      a.S = AIRC_AB_SWEEP_SEGMENTS;
#endif
      a.L = a.S > 1 ? (H + a.S - 1) / a.S : (H > 0 ? H : 1);
      a.S = H > 0 ? (H + a.L - 1) / a.L : 1;
    }
    sweep2_prep_kernel<<<1, 256, 0, st>>>(U_nom, H, cc.neg_log2e_over_eps, key, thrust_du, flag, full ? a.G : 0);
    if (B > 0 && use_tc) {
      const unsigned grid = (unsigned)(((long long)a.GB * a.S + 2) / 3);
      if (ck) sweep2tc_kernel<true><<<grid, AIRC_S2TC_BS, 0, st>>>(h->dc, h->tcn, a);
      else sweep2tc_kernel<false><<<grid, AIRC_S2TC_BS, 0, st>>>(h->dc, h->tcn, a);
    } else if (B > 0) {
      const unsigned grid = (unsigned)((long long)a.GB * a.S);
      if (ck) sweep2_kernel<true><<<grid, bs, 0, st>>>(h->dc, h->un, a);
      else sweep2_kernel<false><<<grid, bs, 0, st>>>(h->dc, h->un, a);
    }
    WinnerArgs w;
    memset(&w, 0, sizeof(w));
    w.x0 = x0; w.Unom = U_nom; w.ckpt = a.ckpt; w.key = key; w.smp = smp; w.record = record; w.dt = dt;
    w.B = B; w.H = H; w.C = ck ? L.C : (H > 0 ? H : 1); w.substeps = substeps;
    w.normalise = (int)(flags & AIRC_FLAG_NORMALISE);
    const int wthreads = ck ? L.NS : 1;
    sweep_winner2_kernel<<<(wthreads + 31) / 32, 32, 0, st>>>(h->dc, h->un, w);
    return check_launch();
  }
  sweep_prep_kernel<<<1, 32, 0, st>>>(U_nom, H, cc.neg_log2e_over_eps, key, thrust_du);
  if (B > 0) {
    SweepArgs a;
    memset(&a, 0, sizeof(a));
    a.x0 = x0; a.Unom = U_nom; a.cost = cost; a.ckpt = ckpt; a.key = key; a.thrust_du = thrust_du;
    a.cc = cc; a.smp = smp; a.dt = dt; a.B = B; a.H = H; a.C = C; a.substeps = substeps;
    a.normalise = (int)(flags & AIRC_FLAG_NORMALISE);
    const int bs = (B <= (int64_t)h->sm_count * 32 * 2) ? 32 : ((B <= (int64_t)h->sm_count * 64 * 4) ? 64 : AIRC_ROLLOUT_BS);
    const unsigned grid = (unsigned)((B + bs - 1) / bs);
    if (h->dc.model) {
      if (use_ckpt) sweep_kernel<true, 1><<<grid, bs, 0, st>>>(h->dc, a);
      else sweep_kernel<false, 1><<<grid, bs, 0, st>>>(h->dc, a);
    } else if (use_ckpt) sweep_kernel<true, 0><<<grid, bs, 0, st>>>(h->dc, a);
    else sweep_kernel<false, 0><<<grid, bs, 0, st>>>(h->dc, a);
  }
  WinnerArgs w;
  memset(&w, 0, sizeof(w));
  w.x0 = x0; w.Unom = U_nom; w.ckpt = ckpt; w.key = key; w.smp = smp; w.record = record; w.dt = dt;
  w.B = B; w.H = H; w.C = use_ckpt ? C : (H > 0 ? H : 1); w.substeps = substeps;
  w.normalise = (int)(flags & AIRC_FLAG_NORMALISE);
  const int wthreads = use_ckpt ? S : 1;
  if (h->dc.model) sweep_winner_kernel<1><<<(wthreads + 31) / 32, 32, 0, st>>>(h->dc, w);
  else sweep_winner_kernel<0><<<(wthreads + 31) / 32, 32, 0, st>>>(h->dc, w);
  return check_launch();
}

// ---- airc_sweep_static: the sweep on a workspace that lives in the library image (no allocation at call time).
// In this translation unit's HOST code the name g_sweep_static_ws is only a host-side stand-in for the device
// array: its device address comes from cudaGetSymbolAddress, once per device with that device current, and that
// address is what the kernels receive.
static __device__ __align__(256) unsigned char g_sweep_static_ws[AIRC_SWEEP_STATIC_WORKSPACE_BYTES];

static int sweep_static_workspace(int device, void** out) {
  enum { MAX_DEV = 64 };
  static void* ptrs[MAX_DEV];            // per-device address, filled by the first call for that device
  static int state[MAX_DEV];             // 0 unresolved, 1 resolved
  if (device < 0 || device >= MAX_DEV) return fail(AIRC_EINVAL, "airc_sweep_static: device index out of range%s", "");
  if (__atomic_load_n(&state[device], __ATOMIC_ACQUIRE) != 1) {
    DeviceGuard g(device);
    if (!g.ok) return fail(AIRC_ECUDA, "cudaSetDevice failed%s", "");
    void* p = nullptr;
    CUDA_TRY(cudaGetSymbolAddress(&p, g_sweep_static_ws));
    if (!p) return fail(AIRC_ECUDA, "airc_sweep_static: cudaGetSymbolAddress returned NULL%s", "");
    ptrs[device] = p;                    // threads racing here all resolve the same address
    __atomic_store_n(&state[device], 1, __ATOMIC_RELEASE);
  }
  *out = ptrs[device];
  return AIRC_OK;
}

int airc_sweep_static(const AircHandle* h, const float* x0, const float* U_nom, float dt, int64_t B, int H,
                      int substeps, uint32_t flags, const AircSweepParams* sp, const AircCostParams* cp,
                      float* cost, float* record, void* stream) {
  if (!h || !sp || B < 0 || H < 0) return fail(AIRC_EINVAL, "airc_sweep_static: bad argument%s", "");
  const size_t need = airc_sweep_workspace_bytes(B, H, sp->ckpt_interval);
  if (need > (size_t)AIRC_SWEEP_STATIC_WORKSPACE_BYTES) {
    snprintf(g_err, sizeof(g_err),
             "airc_sweep_static: B=%lld x H=%d needs %zu bytes of workspace, the static array holds %zu "
             "(AIRC_SWEEP_STATIC_WORKSPACE_BYTES); use airc_sweep with a caller-owned workspace",
             (long long)B, H, need, (size_t)AIRC_SWEEP_STATIC_WORKSPACE_BYTES);
    return AIRC_EINVAL;
  }
  void* ws = nullptr;
  const int rc = sweep_static_workspace(h->device, &ws);
  if (rc) return rc;
  // airc_sweep's prep kernel resets the arg-min key, the thrust term and the hand-over flags on the stream at the
  // start of every call, and every checkpoint / slice slot is written before it is read: nothing that the array
  // still holds from the previous call is consumed.
  return airc_sweep(h, x0, U_nom, dt, B, H, substeps, flags, sp, cp, ws, (size_t)AIRC_SWEEP_STATIC_WORKSPACE_BYTES,
                    cost, record, stream);
}

int airc_select_record(const AircHandle* h, const float* records, int n, int64_t record_floats,
                       float* out_record, void* stream) {
  if (!h || !records || !out_record || n < 1 || record_floats < AIRC_SWEEP_RECORD_HEADER)
    return fail(AIRC_EINVAL, "airc_select_record: bad argument%s", "");
  DeviceGuard g(h->device);
  if (!g.ok) return fail(AIRC_ECUDA, "cudaSetDevice failed%s", "");
  long long blocks = (record_floats + 255) / 256;
  if (blocks > 64) blocks = 64;
  select_record_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(records, n, record_floats, out_record);
  return check_launch();
}

int airc_sweep_controls(const AircHandle* h, const float* U_nom, int H, const AircSweepParams* sp,
                        int64_t first, int64_t count, int layout, float* U, void* stream) {
  if (!h || (!U_nom && H > 0) || !sp || H < 0 || first < 0 || count < 0 || (layout != 0 && layout != 1) || (!U && count > 0 && H > 0))
    return fail(AIRC_EINVAL, "airc_sweep_controls: bad argument%s", "");
  if (count == 0 || H == 0) return AIRC_OK;
  DeviceGuard g(h->device);
  if (!g.ok) return fail(AIRC_ECUDA, "cudaSetDevice failed%s", "");
  sweep_controls_kernel<<<(unsigned)((count + 127) / 128), 128, 0, (cudaStream_t)stream>>>(
      U_nom, H, make_sampler(sp), first, count, layout == AIRC_LAYOUT_SOA ? 1 : 0, U);
  return check_launch();
}

int airc_measure_fp32_peak(const AircHandle* h, int repeats, double* tflops, double* ms) {
  if (!h || !tflops || repeats < 1) return fail(AIRC_EINVAL, "airc_measure_fp32_peak: bad argument%s", "");
  DeviceGuard g(h->device);
  if (!g.ok) return fail(AIRC_ECUDA, "cudaSetDevice failed%s", "");
  float* d = nullptr;
  CUDA_TRY(cudaMalloc(&d, 256));
  cudaEvent_t e0, e1;
  CUDA_TRY(cudaEventCreate(&e0));
  CUDA_TRY(cudaEventCreate(&e1));
  const int iters = 4096, blocks = h->sm_count * 8, threads = 256;
  ffma_peak_kernel<<<blocks, threads>>>(d, iters, 0.999f, 1e-4f);   // warm-up
  double best = 1e30;
  for (int r = 0; r < repeats; ++r) {
    cudaEventRecord(e0);
    ffma_peak_kernel<<<blocks, threads>>>(d, iters, 0.999f, 1e-4f);
    cudaEventRecord(e1);
    CUDA_TRY(cudaEventSynchronize(e1));
    float t = 0.f;
    cudaEventElapsedTime(&t, e0, e1);
    if (t < best) best = t;
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(d);
  const double flops = 2.0 * 64.0 * (double)iters * (double)blocks * (double)threads;
  *tflops = flops / (best * 1e-3) * 1e-12;
  if (ms) *ms = best;
  return check_launch();
}

// ------------------------------------------------------------------ host-buffer entry points
static int ensure_scratch(AircHandle* h, size_t bytes) {
  if (!h->hstream) CUDA_TRY(cudaStreamCreateWithFlags(&h->hstream, cudaStreamNonBlocking));
  if (bytes <= h->scratch_bytes) return AIRC_OK;
  if (h->scratch) { cudaFree(h->scratch); h->scratch = nullptr; h->scratch_bytes = 0; }
  size_t want = bytes + bytes / 4 + 4096;
  cudaError_t e = cudaMalloc(&h->scratch, want);
  if (e != cudaSuccess) { cudaGetLastError(); return fail(AIRC_ENOMEM, "cudaMalloc scratch: %s", cudaGetErrorString(e)); }
  h->scratch_bytes = want;
  return AIRC_OK;
}

static size_t al(size_t n_floats) { return ((n_floats * sizeof(float) + 255) / 256) * 256; }

int airc_state_derivative_host(AircHandle* h, const float* x, const float* u, int64_t n, int layout,
                               float* xdot) {
  if (!h || !x || !u || !xdot || n < 0) return fail(AIRC_EINVAL, "airc_state_derivative_host: bad argument%s", "");
  if (n == 0) return AIRC_OK;
  DeviceGuard g(h->device);
  if (!g.ok) return fail(AIRC_ECUDA, "cudaSetDevice failed%s", "");
  const size_t bx = al(13 * n), bu = al(7 * n);
  int rc = ensure_scratch(h, 2 * bx + bu);
  if (rc) return rc;
  char* base = (char*)h->scratch;
  float *dx = (float*)base, *du = (float*)(base + bx), *dn = (float*)(base + bx + bu);
  CUDA_TRY(cudaMemcpyAsync(dx, x, 13 * n * sizeof(float), cudaMemcpyHostToDevice, h->hstream));
  CUDA_TRY(cudaMemcpyAsync(du, u, 7 * n * sizeof(float), cudaMemcpyHostToDevice, h->hstream));
  rc = airc_state_derivative(h, dx, du, n, layout, dn, h->hstream);
  if (rc) return rc;
  CUDA_TRY(cudaMemcpyAsync(xdot, dn, 13 * n * sizeof(float), cudaMemcpyDeviceToHost, h->hstream));
  CUDA_TRY(cudaStreamSynchronize(h->hstream));
  return AIRC_OK;
}

int airc_diagnostics_host(AircHandle* h, const float* x, const float* u, int64_t n, int layout, float* out) {
  if (!h || !x || !u || !out || n < 0) return fail(AIRC_EINVAL, "airc_diagnostics_host: bad argument%s", "");
  if (n == 0) return AIRC_OK;
  DeviceGuard g(h->device);
  if (!g.ok) return fail(AIRC_ECUDA, "cudaSetDevice failed%s", "");
  const size_t bx = al(13 * n), bu = al(7 * n), bo = al((size_t)AIRC_NDIAG * n);
  int rc = ensure_scratch(h, bx + bu + bo);
  if (rc) return rc;
  char* base = (char*)h->scratch;
  float *dx = (float*)base, *du = (float*)(base + bx), *dn = (float*)(base + bx + bu);
  CUDA_TRY(cudaMemcpyAsync(dx, x, 13 * n * sizeof(float), cudaMemcpyHostToDevice, h->hstream));
  CUDA_TRY(cudaMemcpyAsync(du, u, 7 * n * sizeof(float), cudaMemcpyHostToDevice, h->hstream));
  rc = airc_diagnostics(h, dx, du, n, layout, dn, h->hstream);
  if (rc) return rc;
  CUDA_TRY(cudaMemcpyAsync(out, dn, (size_t)AIRC_NDIAG * n * sizeof(float), cudaMemcpyDeviceToHost, h->hstream));
  CUDA_TRY(cudaStreamSynchronize(h->hstream));
  return AIRC_OK;
}

int airc_state_update_host(AircHandle* h, const float* x, const float* u, float dt,
                           const float* dt_per_node, int64_t n, int substeps, uint32_t flags,
                           int layout, float* x_next) {
  if (!h || !x || !u || !x_next || n < 0) return fail(AIRC_EINVAL, "airc_state_update_host: bad argument%s", "");
  if (n == 0) return AIRC_OK;
  DeviceGuard g(h->device);
  if (!g.ok) return fail(AIRC_ECUDA, "cudaSetDevice failed%s", "");
  const size_t bx = al(13 * n), bu = al(7 * n), bd = dt_per_node ? al(n) : 0;
  int rc = ensure_scratch(h, 2 * bx + bu + bd);
  if (rc) return rc;
  char* base = (char*)h->scratch;
  float *dx = (float*)base, *du = (float*)(base + bx), *dn = (float*)(base + bx + bu);
  float* ddt = dt_per_node ? (float*)(base + 2 * bx + bu) : nullptr;
  CUDA_TRY(cudaMemcpyAsync(dx, x, 13 * n * sizeof(float), cudaMemcpyHostToDevice, h->hstream));
  CUDA_TRY(cudaMemcpyAsync(du, u, 7 * n * sizeof(float), cudaMemcpyHostToDevice, h->hstream));
  if (ddt) CUDA_TRY(cudaMemcpyAsync(ddt, dt_per_node, n * sizeof(float), cudaMemcpyHostToDevice, h->hstream));
  rc = airc_state_update(h, dx, du, dt, ddt, n, substeps, flags, layout, dn, h->hstream);
  if (rc) return rc;
  CUDA_TRY(cudaMemcpyAsync(x_next, dn, 13 * n * sizeof(float), cudaMemcpyDeviceToHost, h->hstream));
  CUDA_TRY(cudaStreamSynchronize(h->hstream));
  return AIRC_OK;
}

int airc_rollout_host(AircHandle* h, const float* x0, const float* U, float dt, int64_t B, int H,
                      int substeps, uint32_t flags, int layout, float* X, float* x_final,
                      float* cost, const AircCostParams* cp) {
  if (!h || B < 0 || H < 0) return fail(AIRC_EINVAL, "airc_rollout_host: bad argument%s", "");
  if (B == 0) return AIRC_OK;
  if (!x0 || (!U && H > 0)) return fail(AIRC_EINVAL, "airc_rollout_host: NULL x0 or U%s", "");
  DeviceGuard g(h->device);
  if (!g.ok) return fail(AIRC_ECUDA, "cudaSetDevice failed%s", "");
  const size_t nU = (flags & AIRC_FLAG_SHARED_U) ? (size_t)7 * H : (size_t)7 * H * B;
  const size_t nX = X ? (size_t)13 * (H + 1) * B : 0;
  const size_t bx = al(13 * B), bU = al(nU), bX = al(nX), bc = al(B);
  int rc = ensure_scratch(h, 2 * bx + bU + bX + bc);
  if (rc) return rc;
  char* base = (char*)h->scratch;
  float* dx0 = (float*)base;
  float* dU = (float*)(base + bx);
  float* dX = X ? (float*)(base + bx + bU) : nullptr;
  float* dxf = x_final ? (float*)(base + bx + bU + bX) : nullptr;
  float* dc = cost ? (float*)(base + 2 * bx + bU + bX) : nullptr;
  CUDA_TRY(cudaMemcpyAsync(dx0, x0, 13 * B * sizeof(float), cudaMemcpyHostToDevice, h->hstream));
  if (nU) CUDA_TRY(cudaMemcpyAsync(dU, U, nU * sizeof(float), cudaMemcpyHostToDevice, h->hstream));
  rc = airc_rollout(h, dx0, dU, dt, B, H, substeps, flags, layout, dX, dxf, dc, cp, h->hstream);
  if (rc) return rc;
  if (X) CUDA_TRY(cudaMemcpyAsync(X, dX, nX * sizeof(float), cudaMemcpyDeviceToHost, h->hstream));
  if (x_final) CUDA_TRY(cudaMemcpyAsync(x_final, dxf, 13 * B * sizeof(float), cudaMemcpyDeviceToHost, h->hstream));
  if (cost) CUDA_TRY(cudaMemcpyAsync(cost, dc, B * sizeof(float), cudaMemcpyDeviceToHost, h->hstream));
  CUDA_TRY(cudaStreamSynchronize(h->hstream));
  return AIRC_OK;
}

static int linearize_host_impl(AircHandle* h, const float* x, const float* u, float dt,
                               const float* dt_per_node, int64_t n, int substeps, uint32_t flags,
                               int layout, float* x_next, float* A, float* Bm, float* Dt) {
  if (!h || !x || !u || n < 0) return fail(AIRC_EINVAL, "airc_linearize_host: bad argument%s", "");
  if (n == 0) return AIRC_OK;
  DeviceGuard g(h->device);
  if (!g.ok) return fail(AIRC_ECUDA, "cudaSetDevice failed%s", "");
  const size_t bx = al(13 * n), bu = al(7 * n), bA = A ? al(169 * n) : 0, bB = Bm ? al(91 * n) : 0;
  const size_t bd = dt_per_node ? al(n) : 0, bT = Dt ? al(13 * n) : 0;
  int rc = ensure_scratch(h, 2 * bx + bu + bA + bB + bd + bT);
  if (rc) return rc;
  char* base = (char*)h->scratch;
  float* dx = (float*)base;
  float* du = (float*)(base + bx);
  float* dn = x_next ? (float*)(base + bx + bu) : nullptr;
  float* dA = A ? (float*)(base + 2 * bx + bu) : nullptr;
  float* dB = Bm ? (float*)(base + 2 * bx + bu + bA) : nullptr;
  float* ddt = dt_per_node ? (float*)(base + 2 * bx + bu + bA + bB) : nullptr;
  float* dT = Dt ? (float*)(base + 2 * bx + bu + bA + bB + bd) : nullptr;
  if (ddt) CUDA_TRY(cudaMemcpyAsync(ddt, dt_per_node, n * sizeof(float), cudaMemcpyHostToDevice, h->hstream));
  CUDA_TRY(cudaMemcpyAsync(dx, x, 13 * n * sizeof(float), cudaMemcpyHostToDevice, h->hstream));
  CUDA_TRY(cudaMemcpyAsync(du, u, 7 * n * sizeof(float), cudaMemcpyHostToDevice, h->hstream));
  rc = airc_linearize_dt(h, dx, du, dt, ddt, n, substeps, flags, layout, dn, dA, dB, dT, h->hstream);
  if (rc) return rc;
  if (x_next) CUDA_TRY(cudaMemcpyAsync(x_next, dn, 13 * n * sizeof(float), cudaMemcpyDeviceToHost, h->hstream));
  if (A) CUDA_TRY(cudaMemcpyAsync(A, dA, 169 * n * sizeof(float), cudaMemcpyDeviceToHost, h->hstream));
  if (Bm) CUDA_TRY(cudaMemcpyAsync(Bm, dB, 91 * n * sizeof(float), cudaMemcpyDeviceToHost, h->hstream));
  if (Dt) CUDA_TRY(cudaMemcpyAsync(Dt, dT, 13 * n * sizeof(float), cudaMemcpyDeviceToHost, h->hstream));
  CUDA_TRY(cudaStreamSynchronize(h->hstream));
  return AIRC_OK;
}

int airc_linearize_host(AircHandle* h, const float* x, const float* u, float dt,
                        const float* dt_per_node, int64_t n, int substeps, uint32_t flags,
                        int layout, float* x_next, float* A, float* Bm) {
  return linearize_host_impl(h, x, u, dt, dt_per_node, n, substeps, flags, layout, x_next, A, Bm, nullptr);
}

int airc_linearize_dt_host(AircHandle* h, const float* x, const float* u, float dt,
                           const float* dt_per_node, int64_t n, int substeps, uint32_t flags,
                           int layout, float* x_next, float* A, float* Bm, float* dPhi_ddt) {
  return linearize_host_impl(h, x, u, dt, dt_per_node, n, substeps, flags, layout, x_next, A, Bm, dPhi_ddt);
}

int airc_hessian_host(AircHandle* h, const float* x, const float* u, const float* lam, float dt,
                      const float* dt_per_node, int64_t n, int substeps, uint32_t flags, int layout,
                      float* H, float* grad, float* value) {
  if (!h || !x || !u || !lam || !H || n < 0) return fail(AIRC_EINVAL, "airc_hessian_host: bad argument%s", "");
  if (n == 0) return AIRC_OK;
  DeviceGuard g(h->device);
  if (!g.ok) return fail(AIRC_ECUDA, "cudaSetDevice failed%s", "");
  const size_t bx = al(13 * n), bu = al(7 * n), bH = al((size_t)AIRC_NZ * AIRC_NZ * n);
  const size_t bg = grad ? al((size_t)AIRC_NZ * n) : 0, bv = value ? al(n) : 0, bd = dt_per_node ? al(n) : 0;
  int rc = ensure_scratch(h, 2 * bx + bu + bH + bg + bv + bd);
  if (rc) return rc;
  char* base = (char*)h->scratch;
  float* dx = (float*)base;
  float* dl = (float*)(base + bx);
  float* du = (float*)(base + 2 * bx);
  float* dH = (float*)(base + 2 * bx + bu);
  float* dg = grad ? (float*)(base + 2 * bx + bu + bH) : nullptr;
  float* dv = value ? (float*)(base + 2 * bx + bu + bH + bg) : nullptr;
  float* ddt = dt_per_node ? (float*)(base + 2 * bx + bu + bH + bg + bv) : nullptr;
  if (ddt) CUDA_TRY(cudaMemcpyAsync(ddt, dt_per_node, n * sizeof(float), cudaMemcpyHostToDevice, h->hstream));
  CUDA_TRY(cudaMemcpyAsync(dx, x, 13 * n * sizeof(float), cudaMemcpyHostToDevice, h->hstream));
  CUDA_TRY(cudaMemcpyAsync(dl, lam, 13 * n * sizeof(float), cudaMemcpyHostToDevice, h->hstream));
  CUDA_TRY(cudaMemcpyAsync(du, u, 7 * n * sizeof(float), cudaMemcpyHostToDevice, h->hstream));
  rc = airc_hessian(h, dx, du, dl, dt, ddt, n, substeps, flags, layout, dH, dg, dv, h->hstream);
  if (rc) return rc;
  CUDA_TRY(cudaMemcpyAsync(H, dH, (size_t)AIRC_NZ * AIRC_NZ * n * sizeof(float), cudaMemcpyDeviceToHost, h->hstream));
  if (grad) CUDA_TRY(cudaMemcpyAsync(grad, dg, (size_t)AIRC_NZ * n * sizeof(float), cudaMemcpyDeviceToHost, h->hstream));
  if (value) CUDA_TRY(cudaMemcpyAsync(value, dv, n * sizeof(float), cudaMemcpyDeviceToHost, h->hstream));
  CUDA_TRY(cudaStreamSynchronize(h->hstream));
  return AIRC_OK;
}

int airc_sweep_host(AircHandle* h, const float* x0, const float* U_nom, float dt, int64_t B, int H,
                    int substeps, uint32_t flags, const AircSweepParams* sp, const AircCostParams* cp,
                    float* cost, float* record) {
  if (!h || !x0 || (!U_nom && H > 0) || !sp || !cp || !record || B < 0 || H < 0)
    return fail(AIRC_EINVAL, "airc_sweep_host: bad argument%s", "");
  DeviceGuard g(h->device);
  if (!g.ok) return fail(AIRC_ECUDA, "cudaSetDevice failed%s", "");
  const size_t wsb = airc_sweep_workspace_bytes(B, H, sp->ckpt_interval);
  const size_t R = (size_t)airc_sweep_record_floats(H);
  const size_t bx = al(13), bU = al((size_t)7 * H), bR = al(R), bc = cost ? al((size_t)B) : 0;
  int rc = ensure_scratch(h, bx + bU + bR + bc + wsb);
  if (rc) return rc;
  char* base = (char*)h->scratch;
  float* dx0 = (float*)base;
  float* dU = (float*)(base + bx);
  float* dR = (float*)(base + bx + bU);
  float* dc = cost ? (float*)(base + bx + bU + bR) : nullptr;
  void* ws = base + bx + bU + bR + bc;
  CUDA_TRY(cudaMemcpyAsync(dx0, x0, 13 * sizeof(float), cudaMemcpyHostToDevice, h->hstream));
  if (H > 0) CUDA_TRY(cudaMemcpyAsync(dU, U_nom, (size_t)7 * H * sizeof(float), cudaMemcpyHostToDevice, h->hstream));
  rc = airc_sweep(h, dx0, dU, dt, B, H, substeps, flags, sp, cp, ws, wsb, dc, dR, h->hstream);
  if (rc) return rc;
  CUDA_TRY(cudaMemcpyAsync(record, dR, R * sizeof(float), cudaMemcpyDeviceToHost, h->hstream));
  if (cost && B > 0) CUDA_TRY(cudaMemcpyAsync(cost, dc, (size_t)B * sizeof(float), cudaMemcpyDeviceToHost, h->hstream));
  CUDA_TRY(cudaStreamSynchronize(h->hstream));
  return AIRC_OK;
}

}  // extern "C"

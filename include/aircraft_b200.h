/* aircraft_b200.h -- C ABI of the B200-native AIrcraft hot path.
 *
 * One shared library (libaircraft_b200.so), plain pointers and sizes, no torch types.
 * It replaces, for numeric evaluation, the CasADi Functions the reference builds in
 *   /root/reference/src/aircraft/dynamics/base.py:385-406   state_derivative  f(x,u)
 *   /root/reference/src/aircraft/dynamics/base.py:408-446   state_step        (RK4)
 *   /root/reference/src/aircraft/dynamics/base.py:450-480   state_update      Phi(x,u,dt)
 * with the NN coefficient model of
 *   /root/reference/src/aircraft/dynamics/coefficient_models.py:91-104 (NeuralModel)
 *   /root/reference/src/aircraft/surrogates/models.py:101-155          (ScaledModel)
 * and the derivative evaluations the reference obtains from ca.jacobian
 *   /root/reference/src/aircraft/control/aircraft.py:85-95, base.py:51-52,
 *   /root/reference/main/stability/stability.py:48-54.
 *
 * State  x[13] = [p_ned(3), v_ned(3), q_frd_ned xyzw(4), omega_frd(3)]   (base.py:94-99)
 * Control u[7] = [aileron, elevator, rudder, thrust(3), flaps]           (aircraft.py:152-158)
 *
 * All data pointers are CALLER-OWNED DEVICE buffers of FP32 unless a function name ends in
 * _host.  Kernels are enqueued on `stream` (a cudaStream_t passed as void*; NULL = legacy
 * default stream) and the call returns without synchronising.  The handle owns only the
 * constant block; it is immutable after create and may be used from several threads/streams.
 * Every function returns 0 on success or a negative AIRC_E* code; none throws.
 * NaN/Inf propagate through the arithmetic unchanged, as in the reference
 * (main/dynamics/dynamics.py:108-110 checks them on the caller's side).
 */
#ifndef AIRCRAFT_B200_H
#define AIRCRAFT_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define AIRC_NX 13
#define AIRC_NU 7
#define AIRC_NN_BLOB_FLOATS 860 /* see AircNNBlob below */

/* error codes */
#define AIRC_OK 0
#define AIRC_EINVAL (-1)   /* bad argument (NULL, negative size, unknown flag/layout) */
#define AIRC_ECUDA (-2)    /* CUDA runtime error; text via airc_last_error() */
#define AIRC_ENODEV (-3)   /* no usable sm_100 device */
#define AIRC_ENOMEM (-4)
#define AIRC_EUNSUPPORTED (-5) /* reserved: entry point not built for this handle's coefficient model (none today) */

/* layout of batched vectors */
#define AIRC_LAYOUT_SOA 0  /* [13][B] / [H][7][B]: batch index fastest -- the native layout; it is
                              also the reference's numpy layout state[13,N], control[7,N]
                              (main/control/control.py:77-90)                                   */
#define AIRC_LAYOUT_AOS 1  /* [B][13] / [B][H][7] */

/* flags */
#define AIRC_FLAG_NORMALISE 1u   /* q <- q/|q| as base.py:443-444 / :473-474 (opts['quaternion']=='integration') */
#define AIRC_FLAG_SHARED_U  2u   /* rollout: U is one [H][7] sequence shared by all trajectories */
#define AIRC_FLAG_COOPERATIVE 4u /* rollout, 'nn' model: latency mode -- 8, 4 or 2 lanes share one trajectory (the network's
                                  * hidden units are split among them; as many lanes as keep the batch at one warp per
                                  * SM sub-partition: B <= 16 / 32 / 64 x SM count).  It is the DEFAULT for B <= 32 x SM
                                  * count (the batch would leave sub-partitions idle on the one-trajectory kernel);
                                  * this flag extends it to B <= 64 x SM count.  Kernel instantiations differ in the
                                  * last bits (summation order, contraction): a sharded batch is bit-identical to the
                                  * whole one only while both sides run the same instantiation. */

#define AIRC_FLAG_PACKED 8u      /* rollout, 'nn' model: run the two-trajectories-per-thread kernel (the sweep's arithmetic)
                                  * whatever the batch size; by default it serves batches of two waves and more, where it
                                  * is the faster one.  A trajectory's bits do not depend on the batch under this flag. */

/* Aircraft constants: AircraftConfiguration (utils.py:201-215), SixDOFOpts (base.py:9-14),
 * AircraftOpts (aircraft.py:22-38).  inertia / inertia_inv are about the shifted CoM
 * (aircraft.py:168-187, base.py:139-144) and are computed by the host side in float64. */
typedef struct AircParams {
  double mass;
  double span;
  double chord;
  double reference_area;
  double com[3];          /* aero_centre_offset */
  double inertia[9];      /* row-major 3x3 */
  double inertia_inv[9];  /* row-major 3x3 */
  double gravity[3];      /* NED, default {0,0,9.81} */
  double epsilon;         /* default 1e-6 */
  int32_t stall_scaling;  /* aircraft.py:280-294, default 0 */
  int32_t reserved;
} AircParams;

/* The surrogate checkpoint, verbatim FP32 tensors of data/networks/model-dynamics.pth in this
 * order (train_nn_surrogate.py:246-252): W1[16][5] b1[16] W2[32][16] b2[32] W3[6][32] b3[6]
 * input_mean[5] input_std[5] output_mean[6] output_std[6]  = 860 floats.
 * The library folds L1.L2 and the scalers in float64 at create time. */
typedef struct AircNNBlob {
  float W1[16 * 5];
  float b1[16];
  float W2[32 * 16];
  float b2[32];
  float W3[6 * 32];
  float b3[6];
  float input_mean[5];
  float input_std[5];
  float output_mean[6];
  float output_std[6];
} AircNNBlob;

/* Primary MPC cost of main/control/control.py:35-70:
 *  w_goal*|p_N[:2]-goal|^2 + w_du*sum_{k,i}(1-exp(-du^2/du_eps)) + w_height*(z_N-z_0)^2
 *  - w_speed*sum_k |v_b(x_k)|^2/H + vel_param*w_vx*vx_N + w_vyz*(vy_N^2+vz_N^2)          */
typedef struct AircCostParams {
  float goal[2];
  float vel_param;
  float w_goal;    /* 1000 */
  float w_du;      /* 100  */
  float du_eps;    /* 1e-2 */
  float w_height;  /* 1    */
  float w_speed;   /* 1/100 */
  float w_vx;      /* 1000 */
  float w_vyz;     /* 1000 */
} AircCostParams;

typedef struct AircHandle AircHandle;

/* version / introspection */
int airc_version(void);
const char* airc_build_info(void);

/* Create a handle for CUDA device `device`.  `nn` may be a HOST pointer only. */
int airc_create(const AircParams* params, const AircNNBlob* nn, int device, AircHandle** out);

/* The registry's other coefficient models (src/aircraft/dynamics/coefficient_models.py:32-37), for
 * every entry point: the primal ones and the sampling sweep run the one-node / one-trajectory-per-thread
 * kernels (the packed two-trajectory kernels are NN-only), airc_linearize(_dt) / airc_defects / airc_hessian
 * run forward mode on jets (airc_hess.cuh).
 *   AIRC_MODEL_POLY     PolynomialModel (coefficient_models.py:106-135; the model of main/control/control.py:166,
 *                       main/mhe/mhtt.py:49, main/rl/rl.py:36).  model_params: float[6][35] = per coefficient
 *                       (CX CY CZ Cl Cm Cn) the 34 weights of the degree-3 monomials of (alpha, beta, aileron,
 *                       elevator) in sklearn PolynomialFeatures order, then the intercept
 *                       (main/surrogates/polynomial_model.py:76-104); then rudder_moment_arm (utils.py:215).  n = 211.
 *   AIRC_MODEL_LINEAR   LinearModel (coefficient_models.py:80-89).  model_params: float[6][6], rows of
 *                       data/networks/linearised.csv (q, alpha, beta, aileron, elevator, intercept).  n = 36.
 *   AIRC_MODEL_DEFAULT  DefaultModel (coefficient_models.py:41-78), constants as written there.  n = 0.   */
#define AIRC_MODEL_POLY 1
#define AIRC_MODEL_LINEAR 2
#define AIRC_MODEL_DEFAULT 3
int airc_create_model(const AircParams* params, int model, const float* model_params, int n_model_params,
                      int device, AircHandle** out);
int airc_destroy(AircHandle* h);
const char* airc_last_error(const AircHandle* h);

/* f(x,u): replaces ca.Function 'dynamics' (base.py:402-406).  n nodes. */
int airc_state_derivative(const AircHandle* h, const float* x, const float* u, int64_t n,
                          int layout, float* xdot, void* stream);

/* The side functions of the dynamics object, all at once, for n nodes: replaces the ca.Functions
 * v_frd_rel, airspeed, alpha, beta, qbar (base.py:147-177,229-241), coefficients (aircraft.py:255-307),
 * forces_frd, moments_frd (aircraft.py:309-330, base.py:253-278) and phi, theta, psi (base.py:179-197) that the
 * reference evaluates on every node for envelope constraints, losses, logging and plots
 * (control/aircraft.py:44-59,140-142; main/control/control.py:63-64).
 * out: [22][n] (SOA) or [n][22] (AOS):  0:3 v_frd_rel | 3 airspeed | 4 alpha | 5 beta | 6 qbar |
 *      7:13 CX CY CZ Cl Cm Cn | 13:16 forces_frd | 16:19 moments_frd | 19:22 phi theta psi */
#define AIRC_NDIAG 22
int airc_diagnostics(const AircHandle* h, const float* x, const float* u, int64_t n, int layout,
                     float* out, void* stream);

/* Phi(x,u,dt): replaces ca.Function 'state_update' (base.py:476-480) for n independent
 * (x,u,dt) nodes.  dt_per_node may be NULL (then dt is used for all nodes). */
int airc_state_update(const AircHandle* h, const float* x, const float* u, float dt,
                      const float* dt_per_node, int64_t n, int substeps, uint32_t flags,
                      int layout, float* x_next, void* stream);

/* x[k+1] = Phi(x[k],u[k],dt), k<H, for B independent trajectories, each kept in registers for
 * the whole horizon: replaces the Python loops of main/control/control.py:84-90,
 * control/moving_horizon.py:208-213, main/dynamics/dynamics.py:105-121.
 *   x0      [13][B] (SOA) or [B][13] (AOS)
 *   U       [H][7][B] (SOA) / [B][H][7] (AOS) / [H][7] with AIRC_FLAG_SHARED_U
 *   X       optional full trajectory [H+1][13][B] (SOA) / [B][H+1][13] (AOS), or NULL
 *   x_final optional [13][B] / [B][13], or NULL
 *   cost    optional [B] (needs cost_params), or NULL                                     */
int airc_rollout(const AircHandle* h, const float* x0, const float* U, float dt, int64_t B,
                 int H, int substeps, uint32_t flags, int layout, float* X, float* x_final,
                 float* cost, const AircCostParams* cost_params, void* stream);

/* (x+, A = dPhi/dx [13][13], Bm = dPhi/du [13][7]) at n nodes by forward-mode tangents:
 * replaces ca.jacobian(state_update, x|u) (control/aircraft.py:88, base.py:51-52).
 *   SOA: x[13][n] u[7][n] x_next[13][n] A[169][n] Bm[91][n]   (row-major 13x13 / 13x7 index first)
 *   AOS: x[n][13] u[n][7] x_next[n][13] A[n][13][13] Bm[n][13][7]
 * x_next, A, Bm may each be NULL.                                                          */
int airc_linearize(const AircHandle* h, const float* x, const float* u, float dt,
                   const float* dt_per_node, int64_t n, int substeps, uint32_t flags, int layout,
                   float* x_next, float* A, float* Bm, void* stream);

/* airc_linearize plus the step-size column dPhi/d(dt): the derivative the reference's time='progress'
 * / variable-dt transcriptions take through ca.jacobian of the defect x[k+1] - Phi(x[k],u[k],dt[k])
 * w.r.t. the dt decision variable (control/base.py:255-286, control/minimal.py time handling).
 * It rides in the 15th tangent slot of the same kernel launch (no extra passes).
 *   dPhi_ddt: SOA [13][n] / AOS [n][13], or NULL (then identical to airc_linearize).
 * With substeps s the derivative is w.r.t. the whole-node dt (each substep sees d(dt/s) = 1/s).  */
int airc_linearize_dt(const AircHandle* h, const float* x, const float* u, float dt,
                      const float* dt_per_node, int64_t n, int substeps, uint32_t flags, int layout,
                      float* x_next, float* A, float* Bm, float* dPhi_ddt, void* stream);

/* Multiple-shooting defects with their Jacobian blocks, one node per (window, step) pair:
 *   defect = x_target - Phi(x, u, dt)          (control/base.py:275-279: next.state - dynamics(...) == 0)
 * together with A = dPhi/dx, Bm = dPhi/du and (optionally) dPhi/d(dt) from the same launch of the
 * Jacobian kernel.  The constraint Jacobian of the reference's NLP is block-bidiagonal with blocks
 * d(defect_k)/d(x_k, u_k, x_{k+1}, dt_k) = (-A_k, -Bm_k, I, -dPhi_ddt_k); these are its blocks, in the form
 * a Riccati / condensing sweep consumes (MHTT.initialise / update_parameters,
 * control/moving_horizon.py:203-246).  Layouts as airc_linearize; x_target and defect like x.
 * A, Bm, dPhi_ddt may each be NULL.                                                                 */
int airc_defects(const AircHandle* h, const float* x, const float* u, const float* x_target, float dt,
                 const float* dt_per_node, int64_t n, int substeps, uint32_t flags, int layout,
                 float* defect, float* A, float* Bm, float* dPhi_ddt, void* stream);

/* Second-order products for the solver: H = d2(lambda^T Phi)/dz2 with z = (x[13], u[7], dt), the
 * per-node block of the Lagrangian Hessian that IPOPT obtains from CasADi with
 * hessian_approximation 'exact' (src/aircraft/config.py:29; constraints control/base.py:275-286; the
 * dt row/column serves time='progress', control/base.py:276,306-307).  15 directions are active; rows
 * and columns of p and thrust are exactly zero; H is symmetric to the bit.  The network model runs
 * forward over reverse (a hand-written adjoint sweep of the RK4 step on dual numbers, 8 threads per
 * node); short horizons (up to 4 nodes per SM; there the 64 threads share the network's sums), the
 * table models and more than 8 substeps run second-order forward mode on 64 threads per node.  The
 * formulations agree to 7e-7 of an entry's scale, so a node's block depends on the batch it is
 * evaluated in only at that level.
 *   lam   multipliers of the 13 rows of Phi, laid out like x
 *   H     SOA [21*21][n] / AOS [n][21][21], full symmetric matrix (required)
 *   grad  optional d(lambda^T Phi)/dz, SOA [21][n] / AOS [n][21]   (= [A B dPhi/ddt]^T lambda)
 *   value optional lambda^T Phi, [n]                                                          */
#define AIRC_NZ 21
int airc_hessian(const AircHandle* h, const float* x, const float* u, const float* lam, float dt,
                 const float* dt_per_node, int64_t n, int substeps, uint32_t flags, int layout,
                 float* H, float* grad, float* value, void* stream);

/* (min, argmin) of cost[B] by warp-shuffle + 64-bit atomicMin; ties go to the lower index.
 * out_min: 1 float, out_idx: 1 int64 (device). */
int airc_argmin(const AircHandle* h, const float* cost, int64_t B, float* out_min,
                int64_t* out_idx, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Sampling-MPC sweep (BASELINE config 4): B control sequences are SAMPLED IN THE KERNEL around one
 * nominal sequence, rolled out from one shared initial state with the fused cost, reduced to
 * (min, arg-min) by warp shuffles + one 64-bit atomicMin per warp, and the winner's trajectory
 * and controls are reproduced from sparse state checkpoints -- no control tensor ever exists in
 * memory and nothing is synchronised with the host.  It replaces, for many candidate sequences
 * at once, the reference's rollout loop + loss of main/control/control.py:35-93.
 *
 * Sampler (counter-based, a pure function of (seed, global trajectory index g, step k, slot c);
 * slots c = 0..3 are the four controls the NN path reads: aileron, elevator, rudder, flaps;
 * thrust rows are inert in the dynamics (aircraft.py:320) and stay at their nominal values):
 *   mix32(x): x ^= x>>16; x *= 0x7feb352d; x ^= x>>15; x *= 0x846ca68b; x ^= x>>16   (uint32)
 *   tkey   = mix32((uint32)g + mix32((uint32)seed)) ^ (uint32)(seed>>32)
 *   w(k,c) = mix32(tkey + k*0x9E3779B9 + c*0x7F4A7C15)
 *   n(k,c) = byte0(w)+byte1(w)+byte2(w)+byte3(w) - 510          (integer in [-510,510], ~N(0,147.8^2))
 *   d_0    = s0_c*n(0,c);  d_k = fmaf(s_c, n(k,c), d_{k-1})     (FP32; s = (float)(sigma/147.80053))
 *   u_k[c] = min(max(U_nom[k][c] + d_k, lo_c), hi_c)        (a NaN in U_nom stays NaN)
 * oracle/cpu_ref.cpp restates the same generator, so every sampled sequence can be regenerated
 * bit-identically on the host. */
typedef struct AircSweepParams {
  uint64_t seed;
  uint64_t index_offset;  /* global index of this shard's trajectory 0 (sampler and arg-min use it) */
  float sigma0[4];        /* std of the initial offset d_0: aileron, elevator, rudder, flaps */
  float sigma[4];         /* std of the random-walk increment per step */
  float lo[4];
  float hi[4];
  int32_t ckpt_interval;  /* state checkpoint every this many steps; 0 = default (8) */
  int32_t reserved;
} AircSweepParams;

/* The winner's record, FP32 words: [0] min cost, [1] its orderable bit pattern (uint32),
 * [2..3] global arg-min (int64), then trajectory [H+1][13], then controls [H][7]. */
#define AIRC_SWEEP_RECORD_HEADER 4
int64_t airc_sweep_record_floats(int H);
/* bytes of device workspace airc_sweep needs (checkpoints dominate: 17 floats x B per checkpoint) */
size_t airc_sweep_workspace_bytes(int64_t B, int H, int ckpt_interval);

/* x0 [13], U_nom [H][7], workspace, cost (optional, [B]) and record are DEVICE pointers.
 * Sweeps of B >= 3 x 256 x SM count trajectories ('nn' model, full workspace) evaluate the network layers on the
 * tensor cores (TF32 operands split in three, FP32 accumulation: same results to ~1e-6 relative); the record's
 * trajectory is then re-rolled from the sweep's checkpoints with the FP32 FMA arithmetic of airc_rollout. */
int airc_sweep(const AircHandle* h, const float* x0, const float* U_nom, float dt, int64_t B, int H,
               int substeps, uint32_t flags, const AircSweepParams* sp, const AircCostParams* cp,
               void* workspace, size_t workspace_bytes, float* cost, float* record, void* stream);

/* airc_sweep WITHOUT a caller workspace and without any allocation at call time.  No counterpart in the reference:
 * added for callers that cannot call cudaMalloc (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid it).  The workspace is a file-scope `static __device__`
 * array of AIRC_SWEEP_STATIC_WORKSPACE_BYTES inside the library image: every GPU that loads the library holds
 * it until the process exits.  x0, U_nom, cost (optional) and record are the caller's DEVICE buffers exactly as
 * for airc_sweep; results are those of airc_sweep with a workspace of airc_sweep_workspace_bytes(B, H, ckpt).
 *  - limit: airc_sweep_workspace_bytes(B, H, sp->ckpt_interval) <= AIRC_SWEEP_STATIC_WORKSPACE_BYTES (32 MiB);
 *    B = 16384 x H = 200 at the default checkpoint interval needs 26.9 MiB.  Larger sweeps return AIRC_EINVAL
 *    with the two sizes in airc_last_error() -- use airc_sweep with your own workspace for those.
 *  - calls no cudaMalloc / cudaFree / allocating library routine.  The first call on a device resolves the
 *    array's device address (cudaGetSymbolAddress, once per device): make that call before a graph capture starts.
 *  - every call on a device shares the one array; it keeps the last call's contents, and each call resets on its
 *    stream what must start clean (arg-min key, hand-over flags).  Calls that overlap on the GPU (two streams,
 *    two handles, two threads) RACE on it: serialise them, or give all but one a workspace through airc_sweep. */
#define AIRC_SWEEP_STATIC_WORKSPACE_BYTES (32u << 20)
int airc_sweep_static(const AircHandle* h, const float* x0, const float* U_nom, float dt, int64_t B, int H,
                      int substeps, uint32_t flags, const AircSweepParams* sp, const AircCostParams* cp,
                      float* cost, float* record, void* stream);

/* Picks the best of n records (lowest cost, ties to the lower global index, NaN last) laid out
 * with a stride of record_floats -- the step after an all-gather of per-rank records. */
int airc_select_record(const AircHandle* h, const float* records, int n, int64_t record_floats,
                       float* out_record, void* stream);

/* The controls the sweep samples for trajectories [first, first+count): U [count][H][7] (AoS) or
 * [H][7][count] (SoA), device memory -- for inspection and for the parity tests. */
int airc_sweep_controls(const AircHandle* h, const float* U_nom, int H, const AircSweepParams* sp,
                        int64_t first, int64_t count, int layout, float* U, void* stream);

/* Measures the device's FP32 FFMA throughput (independent fused multiply-adds, all SMs, best of
 * `repeats` launches, CUDA events): the roofline denominator of the rollout kernel, which no
 * vendor library reports.  Synchronises.  ms (optional) receives the best launch time. */
int airc_measure_fp32_peak(const AircHandle* h, int repeats, double* tflops, double* ms);

/* Convenience for FFI callers that hold HOST buffers (the cgo/ctypes style binding a maintainer
 * of the reference would use from numpy): allocates device scratch, copies in, runs, copies out,
 * synchronises.  Same semantics as the device-pointer versions. */
int airc_state_derivative_host(AircHandle* h, const float* x, const float* u, int64_t n, int layout,
                               float* xdot);
int airc_diagnostics_host(AircHandle* h, const float* x, const float* u, int64_t n, int layout, float* out);
int airc_state_update_host(AircHandle* h, const float* x, const float* u, float dt,
                           const float* dt_per_node, int64_t n, int substeps, uint32_t flags,
                           int layout, float* x_next);
int airc_rollout_host(AircHandle* h, const float* x0, const float* U, float dt, int64_t B, int H,
                      int substeps, uint32_t flags, int layout, float* X, float* x_final,
                      float* cost, const AircCostParams* cost_params);
int airc_linearize_host(AircHandle* h, const float* x, const float* u, float dt,
                        const float* dt_per_node, int64_t n, int substeps, uint32_t flags,
                        int layout, float* x_next, float* A, float* Bm);
int airc_hessian_host(AircHandle* h, const float* x, const float* u, const float* lam, float dt,
                      const float* dt_per_node, int64_t n, int substeps, uint32_t flags, int layout,
                      float* H, float* grad, float* value);
int airc_linearize_dt_host(AircHandle* h, const float* x, const float* u, float dt,
                           const float* dt_per_node, int64_t n, int substeps, uint32_t flags,
                           int layout, float* x_next, float* A, float* Bm, float* dPhi_ddt);

/* airc_sweep with HOST x0 [13], U_nom [H][7], cost (optional, [B]) and record: copies in, runs the three
 * kernels on the handle's stream with handle-owned workspace, copies the record out, synchronises. */
int airc_sweep_host(AircHandle* h, const float* x0, const float* U_nom, float dt, int64_t B, int H,
                    int substeps, uint32_t flags, const AircSweepParams* sp, const AircCostParams* cp,
                    float* cost, float* record);

#ifdef __cplusplus
}
#endif
#endif /* AIRCRAFT_B200_H */

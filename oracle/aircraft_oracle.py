"""CPU oracle for the AIrcraft hot path -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline legs may
import this module.  The product path (``aircraft_b200``) never does; it fails loudly
when the CUDA library is missing.

This is a float64 torch restatement of the reference's CasADi graph for

    state_derivative f(x,u)        /root/reference/src/aircraft/dynamics/base.py:385-406
    state_step (RK4)               /root/reference/src/aircraft/dynamics/base.py:408-446
    state_update (substeps/fold)   /root/reference/src/aircraft/dynamics/base.py:450-480

with the NN-surrogate coefficient model (``coefficient_models.py:91-104``,
``surrogates/models.py:101-155``) and, as a fixture only, ``DefaultModel``
(``coefficient_models.py:41-78``), which is what the reference's printed known answers
in ``todo.md`` were produced with.

Third-party arithmetic that is not under /root/reference (restated from its published
semantics, verified numerically against todo.md -- see tests/test_oracle_golden.py):
  * liecasadi==0.0.6 ``Quaternion``: xyzw coefficients, Hamilton product,
    ``inverse() = conj / ||q||^2``, ``normalize() = q / ||q||``.
  * CasADi 3.6.7 ``atan2, asin, sqrt, sign, cross, inv`` and forward AD (``sign'`` = 0).

Parity status: PINNED, in three pieces that together cover the path.
  1. The quaternion/force conventions by the three todo.md known-answer vectors (DefaultModel).
  2. The NN block by importing the reference's own ``ScaledModel`` (fixtures in tests/golden/).
  3. The composition f -> RK4 -> substeps fold (v_frd_rel, alpha/beta/qbar, forces, moments with the com
     offset, shifted inertia, forces_ned, v_dot, q_dot, omega_dot, state_step, state_update) by a
     trajectory the reference itself produced and stored: data/trajectories/simulation.h5
     (main/dynamics/dynamics.py, poly model, 10 substeps, dt 0.1), reproduced to < 1e-12 over all 39
     steps with ``coeff_model='poly'`` (PolynomialModel restated below from the reference's pickle;
     tests/test_oracle_golden.py::test_state_update_reproduces_reference_stored_trajectory).
What no reference artefact exercises is the *combination* of piece 2 inside piece 3 (no stored
trajectory uses the NN model); the two meet only through the 6-vector of coefficients
(aircraft.py:269-277), which is the same code for every model.  The quaternion normalisation
switch and the Jacobians are checked by properties / central differences only.

All functions are batched over leading dimensions: x[...,13], u[...,7].
State order  x = [p_ned(3), v_ned(3), q_frd_ned xyzw(4), omega_frd(3)]   (base.py:94-99)
Control order u = [aileron, elevator, rudder, thrust(3), flaps]          (aircraft.py:152-158)
"""
from __future__ import annotations

import json
import math
from dataclasses import dataclass, field
from pathlib import Path
from typing import Optional, Sequence

import torch

F64 = torch.float64
_HERE = Path(__file__).resolve().parent
_PKG_DATA = _HERE.parent / "aircraft_b200" / "data"
DEFAULT_WEIGHTS = _PKG_DATA / "networks" / "model-dynamics.pth"
DEFAULT_PROBLEM = _PKG_DATA / "glider" / "problem_definition.json"

# the centre-of-mass offset the MPC/MHTT scripts assign before first use
# (/root/reference/main/control/control.py:169-172, main/mhe/mhtt.py:51-56)
TRIM_COM = (0.0131991, -1.78875e-08, 0.00313384)


# --------------------------------------------------------------------------- quaternions
def qmul(a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    """Hamilton product, xyzw (liecasadi Quaternion.__mul__; call sites base.py:158,286,295)."""
    av, aw = a[..., :3], a[..., 3:4]
    bv, bw = b[..., :3], b[..., 3:4]
    vec = aw * bv + bw * av + torch.linalg.cross(av, bv, dim=-1)
    sc = aw * bw - (av * bv).sum(-1, keepdim=True)
    return torch.cat([vec, sc], dim=-1)


def qinv(q: torch.Tensor) -> torch.Tensor:
    """liecasadi Quaternion.inverse(): conjugate divided by the squared norm."""
    conj = torch.cat([-q[..., :3], q[..., 3:4]], dim=-1)
    return conj / (q * q).sum(-1, keepdim=True)


def _pure(v: torch.Tensor) -> torch.Tensor:
    return torch.cat([v, torch.zeros_like(v[..., :1])], dim=-1)


# --------------------------------------------------------------------------- surrogate
class SurrogateF64:
    """ScaledModel.forward in float64 from the FP32 checkpoint (models.py:143-155;
    loader utils.py:30-40).  Layers: Linear(5,16) -> Linear(16,32) -> Tanh -> Linear(32,6),
    no activation between the first two (models.py:114-123).

    dtype=torch.float32 is the L4CasADi reading of the reference (coefficient_models.py:93-100: the traced torch module
    runs in the checkpoint's own precision inside the float64 CasADi graph): inputs are rounded to float32, the
    network runs in float32 and its output is widened back.  Parity tests quote the GPU's error against both."""

    def __init__(self, path: Path | str = DEFAULT_WEIGHTS, dtype=F64):
        ck = torch.load(str(path), map_location="cpu", weights_only=True)
        sd = ck["model_state_dict"]
        c = lambda t: t.detach().to(dtype)
        self.W1, self.b1 = c(sd["core_layers.0.weight"]), c(sd["core_layers.0.bias"])
        self.W2, self.b2 = c(sd["core_layers.1.weight"]), c(sd["core_layers.1.bias"])
        self.W3, self.b3 = c(sd["core_layers.3.weight"]), c(sd["core_layers.3.bias"])
        self.in_mean, self.in_std = c(ck["input_mean"]), c(ck["input_std"])
        self.out_mean, self.out_std = c(ck["output_mean"]), c(ck["output_std"])
        self.dtype = dtype

    def __call__(self, inp: torch.Tensor) -> torch.Tensor:
        z = (inp.to(self.dtype) - self.in_mean) / self.in_std
        h = z @ self.W1.T + self.b1
        h = h @ self.W2.T + self.b2
        h = torch.tanh(h)
        y = h @ self.W3.T + self.b3
        return (y * self.out_std + self.out_mean).to(inp.dtype)


# --------------------------------------------------------------------------- constants
@dataclass
class OracleConfig:
    """AircraftConfiguration defaults (utils.py:201-215) + SixDOFOpts (base.py:9-14)."""
    mass: float = 1.0
    span: float = 1.0
    chord: float = 1.0
    reference_area: float = 0.238
    com: Sequence[float] = (0.133, 0.0, 0.003)
    Ixx: float = 0.155
    Iyy: float = 0.114
    Izz: float = 0.262
    Ixz: float = 0.01
    epsilon: float = 1e-6
    gravity: Sequence[float] = (0.0, 0.0, 9.81)
    substeps: int = 1                  # physical_integration_substeps (reference default 10)
    normalise: bool = False            # base.py:33 ; control/base.py:182-185
    stall_scaling: bool = False        # aircraft.py:30
    coeff_model: str = "nn"            # 'nn' | 'default' (fixture) | 'poly' (fixture: pins Phi on the reference's stored trajectory)
    rudder_moment_arm: float = 0.5     # utils.py:215 (AircraftConfiguration default; absent from the JSON)
    linear_path: Path | str = field(default=Path(__file__).resolve().parent.parent / "tests" / "golden" / "linear_model.npz")
    poly_path: Path | str = field(default=Path(__file__).resolve().parent.parent / "tests" / "golden" / "poly_model.npz")
    legacy_thrust: bool = False        # todo.md era: forces += thrust (now commented, aircraft.py:320)
    weights: Path | str = field(default=DEFAULT_WEIGHTS)

    @classmethod
    def from_problem(cls, path: Path | str = DEFAULT_PROBLEM, **kw) -> "OracleConfig":
        a = json.load(open(path))["aircraft"]
        base = dict(mass=a["mass"], span=a["span"], chord=a["chord"],
                    reference_area=a["reference_area"], com=tuple(a["aero_centre_offset"]),
                    Ixx=a["Ixx"], Iyy=a["Iyy"], Izz=a["Izz"], Ixz=a["Ixz"])
        base.update(kw)
        return cls(**base)


class AircraftOracle:
    def __init__(self, cfg: Optional[OracleConfig] = None):
        self.cfg = cfg or OracleConfig.from_problem()
        c = self.cfg
        # aircraft.py:168-187 (inertia about the shifted CoM) ; base.py:139-144 (ca.inv)
        x, y, z = (float(t) for t in c.com)
        static = torch.tensor([[c.Ixx, 0.0, c.Ixz], [0.0, c.Iyy, 0.0], [c.Ixz, 0.0, c.Izz]], dtype=F64)
        com_term = torch.tensor([[y * y + z * z, -x * y, -x * z],
                                 [-y * x, x * x + z * z, -y * z],
                                 [-z * x, -z * y, x * x + y * y]], dtype=F64)
        self.I = static + c.mass * com_term
        self.Iinv = torch.linalg.inv(self.I)
        self.com = torch.tensor([x, y, z], dtype=F64)
        self.g = torch.tensor(list(c.gravity), dtype=F64)
        self.nn = SurrogateF64(c.weights) if c.coeff_model == "nn" else None
        if c.coeff_model == "linear":  # the six rows of data/networks/linearised.csv (tests/golden/make_poly_golden.py)
            import numpy as _np
            self.lin = torch.tensor(_np.load(c.linear_path)["table"], dtype=F64)
        if c.coeff_model == "poly":   # tests/golden/make_poly_golden.py
            import numpy as _np
            d = _np.load(c.poly_path)
            self.poly_powers = torch.tensor(d["powers"], dtype=torch.int64)      # [34, 4] over (alpha, beta, aileron, elevator)
            self.poly_coef = torch.tensor(d["coef"], dtype=F64)                  # [6, 34]
            self.poly_icpt = torch.tensor(d["intercept"], dtype=F64)             # [6]

    # ---- aero-angle helpers (base.py:147-177, 229-241)
    def v_frd_rel(self, x: torch.Tensor) -> torch.Tensor:
        q = x[..., 6:10]
        v = _pure(x[..., 3:6])
        return qmul(qmul(qinv(q), v), q)[..., :3] + self.cfg.epsilon

    def aero_inputs(self, x: torch.Tensor):
        eps = self.cfg.epsilon
        vb = self.v_frd_rel(x)
        vb2 = (vb * vb).sum(-1)
        airspeed = torch.sqrt(vb2 + eps)
        alpha = torch.atan2(vb[..., 2], vb[..., 0] + eps)
        beta = torch.asin(vb[..., 1] / airspeed)
        qbar = 0.5 * 1.225 * vb2
        return vb, airspeed, alpha, beta, qbar

    # ---- Euler angles from the xyzw quaternion (base.py:179-197)
    @staticmethod
    def euler(x: torch.Tensor) -> torch.Tensor:
        qx, qy, qz, qw = x[..., 6], x[..., 7], x[..., 8], x[..., 9]
        phi = torch.atan2(2 * (qw * qx + qy * qz), 1 - 2 * (qx ** 2 + qy ** 2))
        theta = torch.asin(2 * (qw * qy - qz * qx))
        psi = torch.atan2(2 * (qw * qz + qx * qy), 1 - 2 * (qy ** 2 + qz ** 2))
        return torch.stack([phi, theta, psi], dim=-1)

    # ---- coefficients (aircraft.py:255-307)
    def coefficients(self, x: torch.Tensor, u: torch.Tensor) -> torch.Tensor:
        vb, _, alpha, beta, qbar = self.aero_inputs(x)
        ail, ele, rud, flaps = u[..., 0], u[..., 1], u[..., 2], u[..., 6]
        if self.cfg.coeff_model == "nn":
            inp = torch.stack([qbar, alpha, beta, ail, ele], dim=-1)
            C = self.nn(inp)
            # coefficient_models.py:102-103
            cn = C[..., 5] + (-0.1) * 6 * rud * math.pi / 180
            C = torch.cat([C[..., :5], cn[..., None]], dim=-1)
        elif self.cfg.coeff_model == "poly":   # PolynomialModel, coefficient_models.py:106-135
            eps, b, arm = self.cfg.epsilon, self.cfg.span, self.cfg.rudder_moment_arm
            uu, vv, ww = vb[..., 0], vb[..., 1], vb[..., 2]
            p, qq, r = x[..., 10], x[..., 11], x[..., 12]
            zero = torch.zeros_like(alpha)
            a_el = torch.atan2(ww + arm * qq, uu + eps)                      # aircraft.py:189-199
            a_lw = torch.atan2(ww - b * p / 4, uu + eps)                     # aircraft.py:202-211
            a_rw = torch.atan2(ww + b * p / 4, uu + eps)                     # aircraft.py:213-222
            vr = vv - arm * r                                               # aircraft.py:224-233
            b_rud = torch.asin(vr / torch.sqrt(uu * uu + vr * vr + ww * ww + eps))
            out = self._poly(alpha, beta, ail, ele)
            lw, rw = self._poly(a_lw, zero, zero, zero), self._poly(a_rw, zero, zero, zero)
            cl = out[..., 3] + b / 4 * (rw[..., 2] / 2 - lw[..., 2] / 2)
            cm = self._poly(a_el, beta, ail, ele)[..., 4]
            cn = self._poly(alpha, b_rud, ail, ele)[..., 5] + 0.01 * 6 * rud * math.pi / 180
            C = torch.stack([out[..., 0], out[..., 1], out[..., 2], cl, cm, cn], dim=-1)
        elif self.cfg.coeff_model == "linear":   # LinearModel, coefficient_models.py:80-89
            inp = torch.stack([qbar, alpha, beta, ail, ele, torch.ones_like(alpha)], dim=-1)
            C = inp @ self.lin.T
            cn = C[..., 5] + (-0.1) * 6 * rud * math.pi / 180
            C = torch.cat([C[..., :5], cn[..., None]], dim=-1)
        else:  # DefaultModel, coefficient_models.py:41-78
            p, qq, r = x[..., 10], x[..., 11], x[..., 12]
            CD = 0.02 + 0.3 * alpha ** 2
            CL = 0.0 + 5.0 * alpha
            CY = -0.98 * beta
            Cl = 0.08 * 4 * ail * math.pi / 180 + (-0.05) * p
            Cm = -1.2 * 5 * ele * math.pi / 180 + (-0.5) * qq
            Cn = -0.1 * 6 * rud * math.pi / 180 + (-0.05) * r
            C = torch.stack([-CD, CY, -CL, Cl, Cm, Cn], dim=-1)
        c = [C[..., i] for i in range(6)]
        if self.cfg.stall_scaling:  # aircraft.py:280-294
            sa = 1 / (1 + torch.exp(10 * (alpha.abs() - math.radians(30))))
            sb = 1 / (1 + torch.exp(10 * (beta.abs() - math.radians(30))))
            c[2] = c[2] * sa * sb
            c[4] = c[4] * sa
        c[0] = c[0] + (-0.1) * flaps      # aircraft.py:297-299
        c[2] = c[2] + (-0.6) * flaps
        return torch.stack(c, dim=-1)

    def _poly(self, alpha, beta, ail, ele):
        """intercept + coef . monomials(alpha, beta, aileron, elevator): the CasADi functions that
        main/surrogates/polynomial_model.py:76-104 generates from the stored sklearn fit."""
        z = torch.stack([alpha, beta, ail, ele], dim=-1)                       # [..., 4]
        mono = (z[..., None, :] ** self.poly_powers).prod(-1)                   # [..., 34]
        return mono @ self.poly_coef.T + self.poly_icpt

    # ---- forces and moments (aircraft.py:309-330 ; base.py:253-278)
    def forces_frd(self, x, u):
        vb, _, _, _, qbar = self.aero_inputs(x)
        C = self.coefficients(x, u)
        F = C[..., :3] * (qbar * self.cfg.reference_area)[..., None]
        fx = F[..., 0] * torch.sign(vb[..., 0])
        F = torch.stack([fx, F[..., 1], F[..., 2]], dim=-1)
        if self.cfg.legacy_thrust:
            F = F + u[..., 3:6]
        return F

    def moments_frd(self, x, u):
        c = self.cfg
        _, _, _, _, qbar = self.aero_inputs(x)
        C = self.coefficients(x, u)
        arms = torch.tensor([c.span, c.chord, c.span], dtype=F64)
        M = C[..., 3:6] * (qbar * c.reference_area)[..., None] * arms
        F = self.forces_frd(x, u)
        return M + torch.linalg.cross(self.com.expand_as(F), F, dim=-1)

    # ---- f(x,u)  (base.py:280-297, 344-406)
    def state_derivative(self, x: torch.Tensor, u: torch.Tensor) -> torch.Tensor:
        x = x.to(F64)
        u = u.to(F64)
        q = x[..., 6:10]
        w = x[..., 10:13]
        F = self.forces_frd(x, u)
        Fn = qmul(qmul(q, _pure(F)), qinv(q))[..., :3]
        vdot = Fn / self.cfg.mass + self.g
        qdot = (0.5 * qmul(q, _pure(w)))
        M = self.moments_frd(x, u)
        Iw = w @ self.I.T
        wdot = (M - torch.linalg.cross(w, Iw, dim=-1)) @ self.Iinv.T
        return torch.cat([x[..., 3:6], vdot, qdot, wdot], dim=-1)

    # ---- RK4 (base.py:408-446)
    def state_step(self, x, u, h, normalise: bool = False):
        f = self.state_derivative
        k1 = f(x, u)
        k2 = f(x + (h / 2) * k1, u)
        k3 = f(x + (h / 2) * k2, u)
        k4 = f(x + h * k3, u)
        xn = x + (h / 6) * (k1 + 2 * k2 + 2 * k3 + k4)
        if normalise:
            xn = self._normalise(xn)
        return xn

    @staticmethod
    def _normalise(x):
        q = x[..., 6:10]
        q = q / torch.sqrt((q * q).sum(-1, keepdim=True))
        return torch.cat([x[..., :6], q, x[..., 10:]], dim=-1)

    # ---- Phi (base.py:450-480)
    def state_update(self, x, u, dt, substeps: Optional[int] = None, normalise: Optional[bool] = None):
        n = self.cfg.substeps if substeps is None else substeps
        norm = self.cfg.normalise if normalise is None else normalise
        x = torch.as_tensor(x, dtype=F64)
        u = torch.as_tensor(u, dtype=F64)
        dt = torch.as_tensor(dt, dtype=F64)
        if dt.ndim > 0:
            dt = dt[..., None]
        if n == 1:
            return self.state_step(x, u, dt, normalise=norm)
        h = dt / n
        for _ in range(n):
            x = self.state_step(x, u, h, normalise=False)
        return self._normalise(x) if norm else x

    # ---- batched rollout: x0[B,13], U[B,H,7] (or [H,7]) -> X[B,H+1,13]
    def rollout(self, x0, U, dt, substeps=None, normalise=None):
        x = torch.as_tensor(x0, dtype=F64)
        U = torch.as_tensor(U, dtype=F64)
        if U.ndim == 2:
            U = U.expand(x.shape[0], *U.shape)
        out = [x]
        for k in range(U.shape[1]):
            x = self.state_update(x, U[:, k], dt, substeps, normalise)
            out.append(x)
        return torch.stack(out, dim=1)

    # ---- A = dPhi/dx, B = dPhi/du by forward AD (reference: ca.jacobian, control/aircraft.py:88,
    #      main/stability/stability.py:51, base.py:51-52)
    def linearize(self, x, u, dt, substeps=None, normalise=None, want_dt=False):
        x = torch.as_tensor(x, dtype=F64).reshape(-1, 13)
        u = torch.as_tensor(u, dtype=F64).reshape(-1, 7)
        dtt = torch.as_tensor(dt, dtype=F64).expand(x.shape[0]) if torch.as_tensor(dt).ndim == 0 \
            else torch.as_tensor(dt, dtype=F64).reshape(-1)

        def phi(xi, ui, di):
            return self.state_update(xi, ui, di, substeps, normalise)

        if want_dt:   # + dPhi/d(dt), the variable-step column (control/base.py:255-286)
            jac = torch.func.vmap(torch.func.jacfwd(phi, argnums=(0, 1, 2)))
            A, Bm, Dt = jac(x, u, dtt)
            return phi(x, u, dtt), A, Bm, Dt
        jac = torch.func.vmap(torch.func.jacfwd(phi, argnums=(0, 1)))
        A, Bm = jac(x, u, dtt)
        return phi(x, u, dtt), A, Bm

    # ---- H = d2(lam^T Phi)/dz2, z = (x, u, dt): the per-node block of IPOPT's exact Lagrangian Hessian
    #      (config.py:29 hessian_approximation 'exact'; constraints control/base.py:275-286)
    def lagrangian_hessian(self, x, u, lam, dt, substeps=None, normalise=None):
        x = torch.as_tensor(x, dtype=F64).reshape(-1, 13)
        u = torch.as_tensor(u, dtype=F64).reshape(-1, 7)
        lam = torch.as_tensor(lam, dtype=F64).reshape(-1, 13)
        dtt = torch.as_tensor(dt, dtype=F64).expand(x.shape[0]) if torch.as_tensor(dt).ndim == 0 \
            else torch.as_tensor(dt, dtype=F64).reshape(-1)
        z = torch.cat([x, u, dtt[:, None]], dim=1)

        def L(zi, li):
            return (li * self.state_update(zi[:13], zi[13:20], zi[20], substeps, normalise)).sum()

        H = torch.func.vmap(torch.func.jacfwd(torch.func.jacfwd(L)))(z, lam)
        g = torch.func.vmap(torch.func.jacfwd(L))(z, lam)
        return H, g, torch.func.vmap(L)(z, lam)

    # ---- primary MPC cost (main/control/control.py:35-70) on a rollout X[B,H+1,13], U[B,H,7]
    def cost(self, X, U, goal=(150.0, 0.0), vel_param=1.0):
        X = torch.as_tensor(X, dtype=F64)
        U = torch.as_tensor(U, dtype=F64)
        if U.ndim == 2:
            U = U.expand(X.shape[0], *U.shape)
        H = U.shape[1]
        g = torch.tensor(goal, dtype=F64)
        goal_loss = 1000 * ((X[:, -1, :2] - g) ** 2).sum(-1)
        du = U[:, 1:] - U[:, :-1]
        control_loss = 100 * (1 - torch.exp(-du ** 2 / 1e-2)).sum((-1, -2))
        height_loss = (X[:, -1, 2] - X[:, 0, 2]) ** 2
        vb = self.v_frd_rel(X[:, :-1])
        speed_loss = -(vb * vb).sum((-1, -2)) / 100 / H
        fv = vel_param * 1000 * X[:, -1, 3] + 1000 * X[:, -1, 4] ** 2 + 1000 * X[:, -1, 5] ** 2
        return goal_loss + control_loss + height_loss + speed_loss + fv

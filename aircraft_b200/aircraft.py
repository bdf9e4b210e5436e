"""Host-side mirror of the reference's dynamics interface for the NN-surrogate hot path.

Mirrors (names, argument meaning, orderings, error behaviour -- not code):
  Aircraft(opts=AircraftOpts(...))        /root/reference/src/aircraft/dynamics/aircraft.py:105-121
  .state_update(x, u, dt) -> x_next       /root/reference/src/aircraft/dynamics/base.py:450-480
  .state_derivative(x, u) -> x_dot        /root/reference/src/aircraft/dynamics/base.py:385-406
  .num_states / .num_controls             base.py:101 ; aircraft.py:162
  .com / .normalise attributes set by the scripts after construction
                                          main/control/control.py:168-172 ; control/base.py:182-185
  COEFF_MODEL_REGISTRY / NeuralModel      /root/reference/src/aircraft/dynamics/coefficient_models.py:20-37,91-104

Like a CasADi Function, state_update / state_derivative accept one column ([13], [7]) or a
column-batched matrix ([13,N], [7,N]; CasADi maps the Function over columns) and return the
same layout.  [13,N] row-major IS the kernels' native SoA layout, so no transpose happens.

Everything numeric runs in the CUDA library through the C ABI (include/aircraft_b200.h).  There
is no CPU fallback: without libaircraft_b200.so or without an sm_100 device this module raises.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from pathlib import Path
from typing import Any, Callable, Dict, Optional, Sequence

import numpy as np

from . import _cabi
from ._cabi import (FLAG_COOPERATIVE, FLAG_NORMALISE, FLAG_PACKED, FLAG_SHARED_U, LAYOUT_AOS, LAYOUT_SOA, NU, NX, NZ)
from .config import (AircraftConfiguration, AircraftOpts, DEFAULT_MODEL_PATH,
                     DEFAULT_PROBLEM_PATH)
from .dm import DM, Function, as_dm, as_dm_tensor
from .params import (load_linear_model, load_nn_blob, load_poly_model, make_cost_params, make_params,
                     make_sweep_params)

try:  # torch is plumbing (device memory, streams); the numpy/host path does not need it
    import torch
except Exception:  # pragma: no cover
    torch = None


class CoefficientModel:
    """Plugin base, as coefficient_models.py:20-22."""

    def __call__(self, inputs):
        raise NotImplementedError


class NeuralModel(CoefficientModel):
    """The 'nn' registry entry (coefficient_models.py:91-104).  Holds the verbatim checkpoint;
    the network itself is evaluated inside the fused kernels, so calling the model on its own
    evaluates C = coefficients([qbar, alpha, beta, aileron, elevator]) through the library's
    coefficient kernel."""

    def __init__(self, path: Path | str, aircraft: "Aircraft", realtime: bool = False):
        self.path = Path(path) if path else DEFAULT_MODEL_PATH
        self.blob = load_nn_blob(self.path)
        self.aircraft = aircraft


class _TableModel(CoefficientModel):
    """The registry's other models (coefficient_models.py:41-135) as parameter tables for
    airc_create_model.  Every entry point is served: the primal ones and the sampling sweep by the
    one-node / one-trajectory-per-thread kernels, linearize / defects / lagrangian_hessian by forward mode on jets."""
    blob = None
    model_id = 0

    def __init__(self, table: Optional[np.ndarray], aircraft: "Aircraft"):
        self.table = table
        self.aircraft = aircraft


class PolynomialModel(_TableModel):
    """'poly' (coefficient_models.py:106-135): what the reference's MPC / MHTT / RL / simulation scripts
    run with.  Reads the reference's own pickle without CasADi (params.load_poly_model)."""
    model_id = _cabi.MODEL_POLY

    def __init__(self, path, aircraft):
        arm = getattr(aircraft.opts.aircraft_config, "rudder_moment_arm", 0.5)
        super().__init__(load_poly_model(path, arm), aircraft)


class LinearModel(_TableModel):
    """'linear' (coefficient_models.py:80-89) from data/networks/linearised.csv."""
    model_id = _cabi.MODEL_LINEAR

    def __init__(self, path, aircraft):
        super().__init__(load_linear_model(path), aircraft)


class DefaultModel(_TableModel):
    """'default' (coefficient_models.py:41-78): constants as written in the reference."""
    model_id = _cabi.MODEL_DEFAULT

    def __init__(self, aircraft):
        super().__init__(None, aircraft)


# same keys as the reference's registry.  Unknown keys fall back to 'default' there (aircraft.py:37);
# this build refuses them loudly instead of silently changing the model.
COEFF_MODEL_REGISTRY: Dict[str, Callable] = {
    "nn": lambda path, aircraft, **kw: NeuralModel(path, aircraft, realtime=kw.get("realtime", False)),
    "linear": lambda path, aircraft, **kw: LinearModel(path, aircraft),
    "poly": lambda path, aircraft, **kw: PolynomialModel(path, aircraft),
    "default": lambda path, aircraft, **kw: DefaultModel(aircraft),
}


@dataclass
class RolloutResult:
    final: Any = None          # [B,13] (aos) or [13,B] (soa)
    trajectory: Any = None     # [B,H+1,13] (aos) or [H+1,13,B] (soa)
    cost: Any = None           # [B]

    @property
    def finite(self):
        """Per-trajectory mask: final state (and cost) free of NaN/Inf.  NaNs propagate through the kernels
        unchanged, as through CasADi; the reference's callers check for them (main/dynamics/dynamics.py:108)."""
        ref = self.final if self.final is not None else self.trajectory
        if ref is None:
            ref = self.cost
        is_t = _is_torch(ref)
        fin = torch.isfinite if is_t else np.isfinite
        B = self.cost.shape[0] if self.cost is not None else None
        def per_traj(a):
            if a is None:
                return None
            f = fin(a)
            if f.ndim == 1:
                return f
            b_first = B is None or a.shape[0] == B          # aos: batch first; soa: batch last
            axes = tuple(range(1, f.ndim)) if b_first else tuple(range(f.ndim - 1))
            return f.all(dim=axes) if is_t else f.all(axis=axes)
        parts = [m for m in (per_traj(self.final), per_traj(self.trajectory), per_traj(self.cost)) if m is not None]
        out = parts[0]
        for m in parts[1:]:
            out = out & m
        return out


class SweepRecord:
    """The winner of a sampling sweep: views into the library's FP32 record
    [min cost, orderable cost bits, global arg-min (2 words), trajectory [H+1,13], controls [H,7]].
    Nothing is computed or synchronised until a field is read."""

    def __init__(self, raw, H: int, cost=None, workspace=None):
        self.raw = raw                  # the undecoded record (what travels through the all-gather)
        self.H = int(H)
        self.cost = cost                # [B] when requested
        self.workspace = workspace      # device scratch, reusable by the next sweep of the same size

    @property
    def min_cost(self):
        return self.raw[0] if _is_torch(self.raw) else float(self.raw[0])

    @property
    def argmin(self):
        """global trajectory index (0-dim int64 tensor, or int for host records)"""
        if _is_torch(self.raw):
            idx = self.raw[2:4].view(torch.int32).to(torch.int64)
            return (idx[0] & 0xFFFFFFFF) | (idx[1] << 32)
        idx = self.raw[2:4].view(np.uint32)
        return int(idx[0]) | (int(idx[1]) << 32)

    @property
    def trajectory(self):
        hdr, nt = _cabi.SWEEP_RECORD_HEADER, 13 * (self.H + 1)
        return self.raw[hdr:hdr + nt].reshape(self.H + 1, 13)

    @property
    def controls(self):
        hdr, nt = _cabi.SWEEP_RECORD_HEADER, 13 * (self.H + 1)
        return self.raw[hdr + nt:hdr + nt + 7 * self.H].reshape(self.H, 7)


def decode_record(raw, H: int, cost=None) -> SweepRecord:
    """Views into a record [4 + 13(H+1) + 7H] (torch tensor on any device, or numpy)."""
    return SweepRecord(raw, H, cost)


def _check(rc: int, lib, what: str):
    if rc != 0:
        msg = lib.airc_last_error(None)
        raise RuntimeError(f"{what} failed (code {rc}): {msg.decode() if msg else ''}")


def _is_torch(t) -> bool:
    return torch is not None and isinstance(t, torch.Tensor)


class Aircraft:
    """B200 drop-in for the reference's ``Aircraft`` on the NN-surrogate path.

    Differences a caller can see, all forced by the hardware path:
      * arithmetic is FP32 on the GPU (results match the float64 reference to ~1e-6 per step);
      * ``state_update`` / ``state_derivative`` are plain callables (no CasADi graph); they return
        DM-compatible arrays (``.full()``, ``.toarray()``: dm.py), so the reference's numeric call sites run
        unmodified.  Symbolic embedding into ``ca.Opti`` goes through ``casadi_adapter.py`` (a
        ``casadi.Callback`` over the C ABI: eval, Jacobian blocks, Lagrangian-Hessian blocks) where CasADi is
        installed; the same blocks are served directly by :meth:`linearize` / :meth:`lagrangian_hessian`;
      * ``com`` and ``normalise`` may be assigned after construction exactly as the reference's
        scripts do; the constant block is rebuilt on the next call.
    """

    num_states = NX
    num_controls = NU

    def __init__(self, opts: Optional[AircraftOpts] = None, *, com: Optional[Sequence[float]] = None,
                 device: int | str | None = None, normalise: bool = False):
        self.opts = opts if opts is not None else AircraftOpts(coeff_model_type="nn")
        if self.opts.coeff_model_type not in COEFF_MODEL_REGISTRY:
            raise NotImplementedError(f"unknown coeff_model_type {self.opts.coeff_model_type!r}")
        cfg = self.opts.aircraft_config
        self.S, self.b, self.c = cfg.reference_area, cfg.span, cfg.chord
        self.mass = cfg.mass
        self.grav = tuple(self.opts.gravity)
        self.epsilon = self.opts.epsilon
        self.physical_integration_substeps = int(self.opts.physical_integration_substeps)
        self._com = tuple(float(c) for c in (cfg.aero_centre_offset if com is None else com))
        self._stall_scaling = bool(self.opts.stall_scaling)
        self.normalise = bool(normalise)
        self.coefficient_model = COEFF_MODEL_REGISTRY[self.opts.coeff_model_type](
            self.opts.coeff_model_path, self, realtime=self.opts.realtime)
        self._device = self._parse_device(device)
        self._lib = _cabi.load_library()      # raises ImportError when the .so is missing
        self._handle = None

    @classmethod
    def from_problem(cls, path: Path | str = DEFAULT_PROBLEM_PATH, *, com=None, device=None,
                     normalise: bool = False, coeff_model_path: Path | str = "",
                     coeff_model_type: str = "nn", **opts_kw):
        """Aircraft from the ``"aircraft"`` block of problem_definition.json (default: the 'nn' model),
        as the reference's scripts build it (main/control/control.py:159-167)."""
        opts = AircraftOpts(coeff_model_type=coeff_model_type, coeff_model_path=coeff_model_path,
                            aircraft_config=AircraftConfiguration.from_json(path), **opts_kw)
        return cls(opts, com=com, device=device, normalise=normalise)

    # ---- attributes the reference's scripts assign after construction
    @property
    def com(self):
        return self._com

    @com.setter
    def com(self, value):
        self._com = tuple(float(c) for c in np.asarray(value, np.float64).reshape(-1))
        self._drop_handle()

    @property
    def stall_scaling(self):
        return self._stall_scaling

    @stall_scaling.setter
    def stall_scaling(self, value):
        self._stall_scaling = bool(value)
        self._drop_handle()

    @property
    def device(self) -> int:
        return self._device

    @staticmethod
    def _parse_device(device) -> int:
        if device is None:
            if torch is not None and torch.cuda.is_available():
                return torch.cuda.current_device()
            return 0
        if isinstance(device, int):
            return device
        s = str(device)
        if s == "cuda":
            return 0
        if s.startswith("cuda:"):
            return int(s.split(":")[1])
        raise ValueError(f"aircraft_b200 runs on CUDA devices only, got {device!r}")

    # ---- handle management
    def _drop_handle(self):
        if getattr(self, "_handle", None):
            self._lib.airc_destroy(self._handle)
        self._handle = None

    def __del__(self):
        try:
            self._drop_handle()
        except Exception:
            pass

    @property
    def handle(self):
        if self._handle is None:
            params = make_params(self.opts.aircraft_config, com=self._com, gravity=self.grav,
                                 epsilon=self.epsilon, stall_scaling=self._stall_scaling)
            params.mass = float(self.mass)
            h = C.c_void_p()
            cm = self.coefficient_model
            if cm.blob is not None:
                rc = self._lib.airc_create(C.byref(params), C.byref(cm.blob), self._device, C.byref(h))
            else:
                tab = cm.table
                rc = self._lib.airc_create_model(C.byref(params), cm.model_id,
                                                 None if tab is None else tab.ctypes.data,
                                                 0 if tab is None else int(tab.size), self._device, C.byref(h))
            _check(rc, self._lib, "airc_create")
            self._handle = h
        return self._handle

    def _flags(self, normalise: Optional[bool]) -> int:
        n = self.normalise if normalise is None else normalise
        return FLAG_NORMALISE if n else 0

    def _substeps(self, substeps: Optional[int]) -> int:
        n = self.physical_integration_substeps if substeps is None else int(substeps)
        if n < 1:
            raise RuntimeError("physical_integration_substeps must be >= 1")
        return n

    # ---- tensor plumbing
    def _stream(self) -> int:
        return torch.cuda.current_stream(self._device).cuda_stream

    def _dev32(self, t, name: str):
        """float32 contiguous CUDA tensor on this aircraft's device (no copy when already so)."""
        if t.device.type != "cuda" or t.device.index != self._device:
            raise RuntimeError(f"{name} is on {t.device}, expected cuda:{self._device}")
        return t.to(torch.float32).contiguous()

    @staticmethod
    def _host32(a, name: str) -> np.ndarray:
        if _is_torch(a):
            a = a.detach().cpu().numpy()
        return np.ascontiguousarray(np.asarray(a), dtype=np.float32)

    @staticmethod
    def _ptr(a) -> int:
        if a is None:
            return None
        return a.data_ptr() if _is_torch(a) else a.ctypes.data

    def _on_device(self, *arrs) -> bool:
        return any(_is_torch(a) and a.device.type == "cuda" for a in arrs)

    # ------------------------------------------------------------------ CasADi-style callables
    def _columns(self, x, u):
        """Validate CasADi-style column arguments; returns (x2[13,N], u2[7,N], N, squeeze)."""
        if _is_torch(x) or _is_torch(u):
            as2 = lambda a, rows: (a if _is_torch(a) else torch.as_tensor(np.asarray(a)))
            x2, u2 = as2(x, NX), as2(u, NU)
            if x2.device != u2.device:
                dev = x2.device if x2.device.type == "cuda" else u2.device
                x2, u2 = x2.to(dev), u2.to(dev)
        else:
            x2, u2 = np.asarray(x), np.asarray(u)
        squeeze = x2.ndim == 1 and u2.ndim == 1
        if x2.ndim == 1:
            x2 = x2.reshape(NX, 1) if x2.shape[0] == NX else x2
        if u2.ndim == 1:
            u2 = u2.reshape(NU, 1) if u2.shape[0] == NU else u2
        if x2.ndim != 2 or x2.shape[0] != NX:
            raise RuntimeError(f"state must have shape [13] or [13,N], got {tuple(x2.shape)}")
        if u2.ndim != 2 or u2.shape[0] != NU:
            raise RuntimeError(f"control must have shape [7] or [7,N], got {tuple(u2.shape)}")
        n = max(x2.shape[1], u2.shape[1])
        for name, a in (("state", x2), ("control", u2)):
            if a.shape[1] not in (1, n):
                raise RuntimeError(f"{name} has {a.shape[1]} columns, expected 1 or {n}")
        if x2.shape[1] != n:   # CasADi's map repeats single-column arguments
            x2 = x2.expand(NX, n) if _is_torch(x2) else np.broadcast_to(x2, (NX, n))
        if u2.shape[1] != n:
            u2 = u2.expand(NU, n) if _is_torch(u2) else np.broadcast_to(u2, (NU, n))
        return x2, u2, n, squeeze

    def _dt_args(self, dt, n: int, on_device: bool):
        """(scalar dt, per-node array or None)."""
        if _is_torch(dt):
            if dt.numel() == 1:
                return float(dt.item()), None
            d = dt.reshape(-1)
        else:
            d = np.asarray(dt, np.float64).reshape(-1)
            if d.size == 1:
                return float(d[0]), None
        if d.shape[0] != n:
            raise RuntimeError(f"dt has {d.shape[0]} entries, expected 1 or {n}")
        if on_device:
            d = d if _is_torch(d) else torch.as_tensor(d)
            return 0.0, d.to(device=f"cuda:{self._device}", dtype=torch.float32).contiguous()
        return 0.0, self._host32(d, "dt")

    def _state_derivative(self, x, u):
        """f(x,u): the reference's ca.Function 'dynamics' (base.py:402-406)."""
        x2, u2, n, squeeze = self._columns(x, u)
        lib, h = self._lib, self.handle
        if self._on_device(x2, u2):
            xd, ud = self._dev32(x2, "x"), self._dev32(u2, "u")
            out = torch.empty((NX, n), dtype=torch.float32, device=xd.device)
            _check(lib.airc_state_derivative(h, xd.data_ptr(), ud.data_ptr(), n, LAYOUT_SOA,
                                             out.data_ptr(), self._stream()), lib, "airc_state_derivative")
            out = as_dm_tensor(out)
        else:
            xh, uh = self._host32(x2, "x"), self._host32(u2, "u")
            out = np.empty((NX, n), np.float32)
            _check(lib.airc_state_derivative_host(h, xh.ctypes.data, uh.ctypes.data, n, LAYOUT_SOA,
                                                  out.ctypes.data), lib, "airc_state_derivative_host")
            out = self._like(out, x)
        return out[:, 0] if squeeze else out

    def _state_update(self, x, u, dt, *, substeps: Optional[int] = None,
                      normalise: Optional[bool] = None):
        """Phi(x,u,dt): the reference's ca.Function 'state_update' (base.py:476-480).
        n = physical_integration_substeps RK4 steps of dt/n; quaternion normalised once at the end
        when ``normalise`` (base.py:466-474)."""
        x2, u2, n, squeeze = self._columns(x, u)
        lib, h = self._lib, self.handle
        on_dev = self._on_device(x2, u2)
        dts, dtn = self._dt_args(dt, n, on_dev)
        ns, fl = self._substeps(substeps), self._flags(normalise)
        if on_dev:
            xd, ud = self._dev32(x2, "x"), self._dev32(u2, "u")
            out = torch.empty((NX, n), dtype=torch.float32, device=xd.device)
            _check(lib.airc_state_update(h, xd.data_ptr(), ud.data_ptr(), dts, self._ptr(dtn), n, ns,
                                         fl, LAYOUT_SOA, out.data_ptr(), self._stream()),
                   lib, "airc_state_update")
            out = as_dm_tensor(out)
        else:
            xh, uh = self._host32(x2, "x"), self._host32(u2, "u")
            out = np.empty((NX, n), np.float32)
            _check(lib.airc_state_update_host(h, xh.ctypes.data, uh.ctypes.data, dts, self._ptr(dtn),
                                              n, ns, fl, LAYOUT_SOA, out.ctypes.data),
                   lib, "airc_state_update_host")
            out = self._like(out, x)
        return out[:, 0] if squeeze else out

    @staticmethod
    def _like(out: np.ndarray, ref):
        """Host results come back in the caller's float dtype as a DM-compatible numpy array (dm.py: the
        reference hands back float64 casadi.DM values and its call sites use .full() / .toarray() on them;
        FP32 results are widened, not recomputed)."""
        if _is_torch(ref):
            t = torch.from_numpy(out)
            return t.to(ref.dtype) if ref.dtype.is_floating_point else t
        dt = getattr(ref, "dtype", None)
        if dt is not None and np.issubdtype(dt, np.floating):
            return as_dm(out.astype(dt, copy=False))
        return as_dm(out.astype(np.float64))

    # ------------------------------------------------------------------ batched-native entry points
    def rollout(self, x0, U, dt: float, *, substeps: Optional[int] = None,
                normalise: Optional[bool] = None, layout: str = "aos",
                return_trajectory: bool = False, return_final: bool = True,
                cost=None, cooperative: bool = False, packed: bool = False) -> RolloutResult:
        """x[k+1] = Phi(x[k], u[k], dt) for k < H over B independent trajectories in ONE kernel:
        the reference's Python loops (main/control/control.py:84-90,
        control/moving_horizon.py:208-213, main/dynamics/dynamics.py:105-121).

        layout 'aos': x0[B,13], U[B,H,7] (or one shared sequence U[H,7]);
                      trajectory [B,H+1,13], final [B,13]
        layout 'soa': x0[13,B], U[H,7,B] (or U[H,7]); trajectory [H+1,13,B], final [13,B]
                      (batch index fastest: the kernels' native, fully coalesced layout)
        cost: None, True (default weights), a dict of make_cost_params kwargs or an
              AircCostParams -> per-trajectory primary MPC loss (main/control/control.py:35-70).
        cooperative: latency mode ('nn' model): 8 / 4 / 2 lanes share one trajectory (by batch size).  The library picks
              it by itself for B <= 32 x SM count; this flag (AIRC_FLAG_COOPERATIVE) extends it to B <= 64 x SM
              count.  Last-bit differences between the kernel instantiations.
        packed: run the two-trajectories-per-thread kernel (the sweep's arithmetic, AIRC_FLAG_PACKED) whatever the
              batch size; by default it serves batches of two waves and more.
        """
        if layout not in ("aos", "soa"):
            raise RuntimeError(f"layout must be 'aos' or 'soa', got {layout!r}")
        soa = layout == "soa"
        on_dev = self._on_device(x0, U)
        if on_dev:
            x0d = self._dev32(x0 if _is_torch(x0) else torch.as_tensor(x0).cuda(self._device), "x0")
            Ud = self._dev32(U if _is_torch(U) else torch.as_tensor(U).cuda(self._device), "U")
        else:
            x0d, Ud = self._host32(x0, "x0"), self._host32(U, "U")
        if x0d.ndim != 2 or x0d.shape[0 if soa else 1] != NX:
            raise RuntimeError(f"x0 must be {'[13,B]' if soa else '[B,13]'}, got {tuple(x0d.shape)}")
        B = x0d.shape[1 if soa else 0]
        flags = self._flags(normalise) | (FLAG_COOPERATIVE if cooperative else 0) | (FLAG_PACKED if packed else 0)
        if Ud.ndim == 2:
            if Ud.shape[1] != NU:
                raise RuntimeError(f"shared control sequence must be [H,7], got {tuple(Ud.shape)}")
            H = Ud.shape[0]
            flags |= FLAG_SHARED_U
        elif Ud.ndim == 3:
            ok = (Ud.shape[1] == NU and Ud.shape[2] == B) if soa else (Ud.shape[0] == B and Ud.shape[2] == NU)
            if not ok:
                raise RuntimeError(f"U must be {'[H,7,B]' if soa else '[B,H,7]'}, got {tuple(Ud.shape)}")
            H = Ud.shape[0] if soa else Ud.shape[1]
        else:
            raise RuntimeError(f"U must have 2 or 3 dimensions, got {tuple(Ud.shape)}")
        cp = None
        if cost is not None and cost is not False:
            cp = cost if isinstance(cost, _cabi.AircCostParams) else \
                make_cost_params(**(cost if isinstance(cost, dict) else {}))
        shp_X = (H + 1, NX, B) if soa else (B, H + 1, NX)
        shp_f = (NX, B) if soa else (B, NX)
        if on_dev:
            mk = lambda s: torch.empty(s, dtype=torch.float32, device=x0d.device)
        else:
            mk = lambda s: np.empty(s, np.float32)
        X = mk(shp_X) if return_trajectory else None
        xf = mk(shp_f) if return_final else None
        cst = mk((B,)) if cp is not None else None
        lib, h = self._lib, self.handle
        cpp = C.byref(cp) if cp is not None else None
        lay = LAYOUT_SOA if soa else LAYOUT_AOS
        ns = self._substeps(substeps)
        if on_dev:
            _check(lib.airc_rollout(h, x0d.data_ptr(), Ud.data_ptr(), float(dt), B, H, ns, flags, lay,
                                    self._ptr(X), self._ptr(xf), self._ptr(cst), cpp, self._stream()),
                   lib, "airc_rollout")
        else:
            _check(lib.airc_rollout_host(h, x0d.ctypes.data, Ud.ctypes.data, float(dt), B, H, ns, flags,
                                         lay, self._ptr(X), self._ptr(xf), self._ptr(cst), cpp),
                   lib, "airc_rollout_host")
            X, xf, cst = (None if a is None else self._like(a, x0) for a in (X, xf, cst))
        return RolloutResult(final=xf, trajectory=X, cost=cst)

    def defects(self, X, U, dt, *, substeps: Optional[int] = None, normalise: Optional[bool] = None,
                want_dt: bool = False):
        """Multiple-shooting defects r_k = x_{k+1} - Phi(x_k, u_k, dt_k) with their Jacobian blocks for
        any number of windows, one launch of the Jacobian kernel (control/base.py:275-279; the warm
        start / re-linearisation step of MHTT, control/moving_horizon.py:203-246).

        X[..., N+1, 13], U[..., N, 7] (U may carry N+1 rows like the reference's guess array; the last
        is ignored), dt scalar or [..., N] -> r[..., N, 13], A[..., N, 13, 13], B[..., N, 13, 7]
        (+ dPhi/d(dt) [..., N, 13] with want_dt).  The constraint Jacobian of the reference's NLP has the
        blocks d r_k / d(x_k, u_k, x_{k+1}, dt_k) = (-A_k, -B_k, I, -dPhi/d(dt)_k)."""
        host = not self._on_device(X, U)
        dev = f"cuda:{self._device}"
        Xd = torch.as_tensor(np.asarray(X) if not _is_torch(X) else X).to(device=dev, dtype=torch.float32)
        Ud = torch.as_tensor(np.asarray(U) if not _is_torch(U) else U).to(device=dev, dtype=torch.float32)
        if Xd.ndim < 2 or Xd.shape[-1] != NX or Ud.ndim != Xd.ndim or Ud.shape[-1] != NU or \
                Ud.shape[:-2] != Xd.shape[:-2] or Ud.shape[-2] not in (Xd.shape[-2] - 1, Xd.shape[-2]) or \
                Xd.shape[-2] < 2:
            raise RuntimeError(f"defects wants X[...,N+1,13], U[...,N,7]; got {tuple(Xd.shape)}, {tuple(Ud.shape)}")
        N = Xd.shape[-2] - 1
        lead = tuple(Xd.shape[:-2]) + (N,)
        n = int(np.prod(lead))
        x = Xd[..., :-1, :].reshape(n, NX).contiguous()
        xt = Xd[..., 1:, :].reshape(n, NX).contiguous()
        u = Ud[..., :N, :].reshape(n, NU).contiguous()
        if _is_torch(dt) or np.ndim(dt) > 0:
            dtt = torch.as_tensor(np.asarray(dt) if not _is_torch(dt) else dt)
            dtt = dtt.to(device=dev, dtype=torch.float32).expand(lead).reshape(n) if dtt.numel() > 1 else float(dtt)
        else:
            dtt = float(dt)
        dts, dtn = self._dt_args(dtt, n, True)
        mk = lambda *shp: torch.empty(lead + shp, dtype=torch.float32, device=dev)  # noqa: E731
        r, A, Bm = mk(NX), mk(NX, NX), mk(NX, NU)
        Dt = mk(NX) if want_dt else None
        lib = self._lib
        _check(lib.airc_defects(self.handle, x.data_ptr(), u.data_ptr(), xt.data_ptr(), dts, self._ptr(dtn), n,
                                self._substeps(substeps), self._flags(normalise), LAYOUT_AOS, r.data_ptr(),
                                A.data_ptr(), Bm.data_ptr(), self._ptr(Dt), self._stream()),
               lib, "airc_defects")
        out = (r, A, Bm, Dt) if want_dt else (r, A, Bm)
        if host:
            out = tuple(self._like(o.cpu().numpy(), X) for o in out)
        return out

    def defects_from_guess(self, guess, dt, **kw):
        """defects() on the reference's column-per-node guess array [13+7(+extra rows), N+1]
        (main/control/control.py:72-93, MHTT.set_initial_from_array moving_horizon.py:108-114)."""
        g = guess if _is_torch(guess) else np.asarray(guess)
        if g.ndim != 2 or g.shape[0] < NX + NU:
            raise RuntimeError(f"guess must be [>=20, N+1], got {tuple(g.shape)}")
        return self.defects(g[:NX].T, g[NX:NX + NU].T, dt, **kw)

    def lagrangian_hessian(self, x, u, lam, dt, *, substeps: Optional[int] = None,
                           normalise: Optional[bool] = None, layout: str = "aos",
                           want_grad: bool = True, want_value: bool = False):
        """H = d2(lam^T Phi)/dz2 with z = (x, u, dt) at every node: the per-node block of the exact
        Lagrangian Hessian that the reference's IPOPT setup requests from CasADi
        (config.py:29 hessian_approximation 'exact'; constraints control/base.py:275-286; the dt
        row/column serves time='progress', control/base.py:276,306-307).

        layout 'aos': x[...,13], u[...,7], lam[...,13] -> H[...,21,21], grad[...,21], value[...]
        layout 'soa': x[13,N],  u[7,N],  lam[13,N]   -> H[21,21,N],   grad[21,N],   value[N]
        Returns (H, grad, value) with None for the parts not requested.  Long batches of the network model are
        differentiated forward over reverse, short horizons and the table models by second-order forward mode
        (include/aircraft_b200.h, airc_hessian); both are exact derivatives, equal to 5e-7 of an entry's scale."""
        if layout not in ("aos", "soa"):
            raise RuntimeError(f"layout must be 'aos' or 'soa', got {layout!r}")
        soa = layout == "soa"
        on_dev = self._on_device(x, u)
        if on_dev:
            xd, ud, ld = (self._dev32(a if _is_torch(a) else torch.as_tensor(a).cuda(self._device), nm)
                          for a, nm in ((x, "x"), (u, "u"), (lam, "lam")))
        else:
            xd, ud, ld = self._host32(x, "x"), self._host32(u, "u"), self._host32(lam, "lam")
        if soa:
            if xd.ndim != 2 or xd.shape[0] != NX or ud.ndim != 2 or ud.shape[0] != NU or \
                    xd.shape[1] != ud.shape[1] or tuple(ld.shape) != tuple(xd.shape):
                raise RuntimeError(f"soa lagrangian_hessian wants x[13,N], u[7,N], lam[13,N]; got "
                                   f"{tuple(xd.shape)}, {tuple(ud.shape)}, {tuple(ld.shape)}")
            n = xd.shape[1]
            shp_H, shp_g, shp_v = (NZ, NZ, n), (NZ, n), (n,)
        else:
            if xd.shape[-1] != NX or ud.shape[-1] != NU or xd.shape[:-1] != ud.shape[:-1] or \
                    tuple(ld.shape) != tuple(xd.shape):
                raise RuntimeError(f"aos lagrangian_hessian wants x[...,13], u[...,7], lam[...,13]; got "
                                   f"{tuple(xd.shape)}, {tuple(ud.shape)}, {tuple(ld.shape)}")
            lead = tuple(xd.shape[:-1])
            n = int(np.prod(lead)) if lead else 1
            shp_H, shp_g, shp_v = lead + (NZ, NZ), lead + (NZ,), lead
        dts, dtn = self._dt_args(dt, n, on_dev)
        mk = (lambda s: torch.empty(s, dtype=torch.float32, device=xd.device)) if on_dev else \
            (lambda s: np.empty(s, np.float32))
        H = mk(shp_H)
        g = mk(shp_g) if want_grad else None
        v = mk(shp_v) if want_value else None
        lib, h = self._lib, self.handle
        lay = LAYOUT_SOA if soa else LAYOUT_AOS
        ns, fl = self._substeps(substeps), self._flags(normalise)
        if on_dev:
            _check(lib.airc_hessian(h, xd.data_ptr(), ud.data_ptr(), ld.data_ptr(), dts, self._ptr(dtn), n, ns,
                                    fl, lay, H.data_ptr(), self._ptr(g), self._ptr(v), self._stream()),
                   lib, "airc_hessian")
        else:
            _check(lib.airc_hessian_host(h, xd.ctypes.data, ud.ctypes.data, ld.ctypes.data, dts, self._ptr(dtn),
                                         n, ns, fl, lay, H.ctypes.data, self._ptr(g), self._ptr(v)),
                   lib, "airc_hessian_host")
            H, g, v = (None if a is None else self._like(a, x) for a in (H, g, v))
        return H, g, v

    def linearize(self, x, u, dt, *, substeps: Optional[int] = None,
                  normalise: Optional[bool] = None, layout: str = "aos",
                  want_next: bool = True, want_dt: bool = False):
        """(x_next, A = dPhi/dx, B = dPhi/du) at every node by forward-mode tangents: what the
        reference obtains from ca.jacobian(state_update, x|u) (control/aircraft.py:85-95,
        base.py:51-52, main/stability/stability.py:48-54) and IPOPT from the defect constraints
        (control/base.py:275-286).

        layout 'aos': x[...,13], u[...,7] -> x_next[...,13], A[...,13,13], B[...,13,7]
        layout 'soa': x[13,N],  u[7,N]   -> x_next[13,N],  A[13,13,N],  B[13,7,N]

        want_dt=True appends dPhi/d(dt) ([...,13] / [13,N]) to the returned tuple: the step-size
        column of the variable-dt / time='progress' transcriptions (control/base.py:255-286),
        carried in a spare tangent slot of the same kernel launch.
        """
        if layout not in ("aos", "soa"):
            raise RuntimeError(f"layout must be 'aos' or 'soa', got {layout!r}")
        soa = layout == "soa"
        on_dev = self._on_device(x, u)
        if on_dev:
            xd = self._dev32(x if _is_torch(x) else torch.as_tensor(x).cuda(self._device), "x")
            ud = self._dev32(u if _is_torch(u) else torch.as_tensor(u).cuda(self._device), "u")
        else:
            xd, ud = self._host32(x, "x"), self._host32(u, "u")
        if soa:
            if xd.ndim != 2 or xd.shape[0] != NX or ud.ndim != 2 or ud.shape[0] != NU or \
                    xd.shape[1] != ud.shape[1]:
                raise RuntimeError(f"soa linearize wants x[13,N], u[7,N]; got {tuple(xd.shape)}, {tuple(ud.shape)}")
            lead, n = (xd.shape[1],), xd.shape[1]
            shp_x, shp_A, shp_B = (NX, n), (NX, NX, n), (NX, NU, n)
        else:
            if xd.shape[-1] != NX or ud.shape[-1] != NU or xd.shape[:-1] != ud.shape[:-1]:
                raise RuntimeError(f"aos linearize wants x[...,13], u[...,7]; got {tuple(xd.shape)}, {tuple(ud.shape)}")
            lead = tuple(xd.shape[:-1])
            n = int(np.prod(lead)) if lead else 1
            shp_x, shp_A, shp_B = lead + (NX,), lead + (NX, NX), lead + (NX, NU)
        dts, dtn = self._dt_args(dt, n, on_dev)
        mk = (lambda s: torch.empty(s, dtype=torch.float32, device=xd.device)) if on_dev else \
            (lambda s: np.empty(s, np.float32))
        xn = mk(shp_x) if want_next else None
        A, Bm = mk(shp_A), mk(shp_B)
        Dt = mk(shp_x) if want_dt else None
        lib, h = self._lib, self.handle
        lay = LAYOUT_SOA if soa else LAYOUT_AOS
        ns, fl = self._substeps(substeps), self._flags(normalise)
        if on_dev:
            _check(lib.airc_linearize_dt(h, xd.data_ptr(), ud.data_ptr(), dts, self._ptr(dtn), n, ns, fl,
                                         lay, self._ptr(xn), A.data_ptr(), Bm.data_ptr(), self._ptr(Dt),
                                         self._stream()),
                   lib, "airc_linearize_dt")
        else:
            _check(lib.airc_linearize_dt_host(h, xd.ctypes.data, ud.ctypes.data, dts, self._ptr(dtn), n,
                                              ns, fl, lay, self._ptr(xn), A.ctypes.data, Bm.ctypes.data,
                                              self._ptr(Dt)),
                   lib, "airc_linearize_dt_host")
            xn, A, Bm, Dt = (None if a is None else self._like(a, x) for a in (xn, A, Bm, Dt))
        return (xn, A, Bm, Dt) if want_dt else (xn, A, Bm)

    # ------------------------------------------------------------------ side functions (diagnostics)
    def diagnostics(self, x, u):
        """All side functions of the reference's dynamics object in one pass per node (column call
        shape like state_update): dict of v_frd_rel[3,N], airspeed[N], alpha[N], beta[N], qbar[N],
        coefficients[6,N], forces_frd[3,N], moments_frd[3,N], phi[N], theta[N], psi[N]
        (base.py:147-241,253-278; aircraft.py:255-330)."""
        x2, u2, n, squeeze = self._columns(x, u)
        lib, h = self._lib, self.handle
        if self._on_device(x2, u2):
            xd, ud = self._dev32(x2, "x"), self._dev32(u2, "u")
            out = torch.empty((_cabi.NDIAG, n), dtype=torch.float32, device=xd.device)
            _check(lib.airc_diagnostics(h, xd.data_ptr(), ud.data_ptr(), n, LAYOUT_SOA, out.data_ptr(),
                                        self._stream()), lib, "airc_diagnostics")
            out = as_dm_tensor(out)
        else:
            xh, uh = self._host32(x2, "x"), self._host32(u2, "u")
            out = np.empty((_cabi.NDIAG, n), np.float32)
            _check(lib.airc_diagnostics_host(h, xh.ctypes.data, uh.ctypes.data, n, LAYOUT_SOA,
                                             out.ctypes.data), lib, "airc_diagnostics_host")
            out = self._like(out, x)
        res = {}
        for name, sl in _cabi.DIAG_SLICES.items():
            v = out[sl]
            res[name] = v[..., 0] if squeeze else v
        return res

    def _diag(self, name, x, u=None):
        if u is None:      # phi/theta/psi are functions of the state alone (base.py:179-197)
            xs = x if hasattr(x, "shape") else np.asarray(x)
            n = xs.shape[1] if len(xs.shape) == 2 else None
            u = torch.zeros((NU,) if n is None else (NU, n), device=x.device) if _is_torch(x) else \
                np.zeros((NU,) if n is None else (NU, n))
        return self.diagnostics(x, u)[name]

    # The reference exposes all of these as PROPERTIES that return a ca.Function (base.py:147-241,385-480;
    # aircraft.py:255-330): `aircraft.state_update(x, u, dt)` calls the returned function and
    # `aircraft.state_update.size1_in(0)` introspects it.  Same here (dm.Function over the CUDA entry points).
    _FUNCTIONS = {
        # attribute: (Function name, inputs, input sizes, output name, output size)
        "state_update": ("state_update", ("x", "u", "dt"), (NX, NU, 1), "x_next", NX),
        "state_derivative": ("dynamics", ("x", "u"), (NX, NU), "x_dot", NX),
        "v_frd_rel": ("v_frd_rel", ("x", "u"), (NX, NU), "v_frd_rel", 3),
        "airspeed": ("airspeed", ("x", "u"), (NX, NU), "airspeed", 1),
        "alpha": ("alpha", ("x", "u"), (NX, NU), "alpha", 1),
        "beta": ("beta", ("x", "u"), (NX, NU), "beta", 1),
        "qbar": ("qbar", ("x", "u"), (NX, NU), "qbar", 1),
        "coefficients": ("coefficients", ("x", "u"), (NX, NU), "coefficients", 6),
        "forces_frd": ("forces_frd", ("x", "u"), (NX, NU), "forces_frd", 3),
        "moments_frd": ("moments_frd", ("x", "u"), (NX, NU), "moments_frd", 3),
        "phi": ("phi", ("x",), (NX,), "phi", 1),
        "theta": ("theta", ("x",), (NX,), "theta", 1),
        "psi": ("psi", ("x",), (NX,), "psi", 1),
    }

    def __getattr__(self, attr):
        spec = Aircraft._FUNCTIONS.get(attr)
        if spec is None:
            raise AttributeError(attr)
        name, ins, sizes, out, osize = spec
        if attr == "state_update":
            fn = self._state_update
        elif attr == "state_derivative":
            fn = self._state_derivative
        elif len(ins) == 1:
            fn = (lambda x, _a=attr: self._diag(_a, x))
        else:
            fn = (lambda x, u, _a=attr: self._diag(_a, x, u))
        f = Function(name, fn, ins, sizes, (out,), (osize,))
        self.__dict__[attr] = f           # cached: later look-ups do not come through __getattr__
        return f

    # ------------------------------------------------------------------ sampling-MPC sweep
    def sweep(self, x0, U_nom, dt: float, B: int, *, sampler=None, cost=True,
              substeps: Optional[int] = None, normalise: Optional[bool] = None,
              return_costs: bool = False, workspace=None) -> SweepRecord:
        """B control sequences sampled IN THE KERNEL around ``U_nom[H,7]`` (counter-based
        generator, see include/aircraft_b200.h), rolled out from the shared state ``x0[13]`` with
        the fused primary MPC loss (main/control/control.py:35-93 for many candidates at once),
        reduced to (min, arg-min) on the device; the winner's trajectory and controls come back
        in one record.  Nothing is synchronised with the host on the device path.

        sampler: None, a dict of make_sweep_params kwargs, or an AircSweepParams.
        Host (numpy) inputs go through airc_sweep_host and return numpy results."""
        sp = sampler if isinstance(sampler, _cabi.AircSweepParams) else make_sweep_params(**(sampler or {}))
        cp = cost if isinstance(cost, _cabi.AircCostParams) else \
            make_cost_params(**(cost if isinstance(cost, dict) else {}))
        lib, h = self._lib, self.handle
        ns, fl = self._substeps(substeps), self._flags(normalise)
        B = int(B)
        if B < 0:
            raise RuntimeError("B must be >= 0")
        on_dev = self._on_device(x0, U_nom)
        if on_dev:
            xd = self._dev32(x0 if _is_torch(x0) else torch.as_tensor(np.asarray(x0)).cuda(self._device), "x0").reshape(-1)
            Ud = self._dev32(U_nom if _is_torch(U_nom) else torch.as_tensor(np.asarray(U_nom)).cuda(self._device), "U_nom")
        else:
            xd, Ud = self._host32(x0, "x0").reshape(-1), self._host32(U_nom, "U_nom")
        if xd.shape[0] != NX:
            raise RuntimeError(f"x0 must be one state [13], got {tuple(xd.shape)}")
        if Ud.ndim != 2 or Ud.shape[1] != NU:
            raise RuntimeError(f"U_nom must be [H,7], got {tuple(Ud.shape)}")
        H = int(Ud.shape[0])
        R = int(lib.airc_sweep_record_floats(H))
        if on_dev:
            dev = xd.device
            need = int(lib.airc_sweep_workspace_bytes(B, H, sp.ckpt_interval))
            if workspace is None or workspace.numel() < need:
                workspace = torch.empty(need, dtype=torch.uint8, device=dev)
            rec = torch.empty(R, dtype=torch.float32, device=dev)
            cst = torch.empty(B, dtype=torch.float32, device=dev) if return_costs else None
            _check(lib.airc_sweep(h, xd.data_ptr(), Ud.data_ptr(), float(dt), B, H, ns, fl, C.byref(sp),
                                  C.byref(cp), workspace.data_ptr(), workspace.numel(), self._ptr(cst),
                                  rec.data_ptr(), self._stream()), lib, "airc_sweep")
            out = decode_record(rec, H, cst)
            out.workspace = workspace
            return out
        rec = np.empty(R, np.float32)
        cst = np.empty(B, np.float32) if return_costs else None
        _check(lib.airc_sweep_host(h, xd.ctypes.data, Ud.ctypes.data, float(dt), B, H, ns, fl, C.byref(sp),
                                   C.byref(cp), self._ptr(cst), rec.ctypes.data), lib, "airc_sweep_host")
        return decode_record(rec, H, cst)

    def sweep_controls(self, U_nom, sampler, first: int, count: int, layout: str = "aos"):
        """The control sequences the sweep samples for trajectories [first, first+count):
        [count,H,7] ('aos') or [H,7,count] ('soa'), on the device."""
        sp = sampler if isinstance(sampler, _cabi.AircSweepParams) else make_sweep_params(**(sampler or {}))
        Ud = self._dev32(U_nom if _is_torch(U_nom) else torch.as_tensor(np.asarray(U_nom)).cuda(self._device), "U_nom")
        H = int(Ud.shape[0])
        soa = layout == "soa"
        U = torch.empty((H, 7, count) if soa else (count, H, 7), dtype=torch.float32, device=Ud.device)
        _check(self._lib.airc_sweep_controls(self.handle, Ud.data_ptr(), H, C.byref(sp), int(first), int(count),
                                             LAYOUT_SOA if soa else LAYOUT_AOS, U.data_ptr(), self._stream()),
               self._lib, "airc_sweep_controls")
        return U

    def select_record(self, records):
        """Best of n stacked records [n,R] (after an all-gather across ranks)."""
        r = self._dev32(records, "records")
        out = torch.empty(r.shape[1], dtype=torch.float32, device=r.device)
        _check(self._lib.airc_select_record(self.handle, r.data_ptr(), r.shape[0], r.shape[1], out.data_ptr(),
                                            self._stream()), self._lib, "airc_select_record")
        return out

    def argmin(self, cost):
        """(min, argmin) of a device cost vector; ties resolve to the lower index, NaNs sort last."""
        c = self._dev32(cost, "cost").reshape(-1)
        out_min = torch.empty(1, dtype=torch.float32, device=c.device)
        out_idx = torch.empty(1, dtype=torch.int64, device=c.device)
        _check(self._lib.airc_argmin(self.handle, c.data_ptr(), c.numel(), out_min.data_ptr(),
                                     out_idx.data_ptr(), self._stream()), self._lib, "airc_argmin")
        return out_min, out_idx

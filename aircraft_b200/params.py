"""Packs the reference's artefacts into the two C-ABI structs.

  checkpoint  data/networks/model-dynamics.pth, read exactly as the reference's loader does
              (/root/reference/src/aircraft/utils.py:30-40: torch.load(weights_only=True),
              keys model_state_dict + input/output mean/std)      -> AircNNBlob (verbatim FP32)
  constants   AircraftConfiguration + SixDOFOpts                   -> AircParams (float64)
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path
from typing import Sequence

import numpy as np
import torch

from ._cabi import AircNNBlob, AircParams, AircCostParams, AircSweepParams
from .config import AircraftConfiguration, DEFAULT_MODEL_PATH, inertia_about_com

_BLOB_FIELDS = [
    ("W1", "core_layers.0.weight", (16, 5)), ("b1", "core_layers.0.bias", (16,)),
    ("W2", "core_layers.1.weight", (32, 16)), ("b2", "core_layers.1.bias", (32,)),
    ("W3", "core_layers.3.weight", (6, 32)), ("b3", "core_layers.3.bias", (6,)),
]
_SCALER_FIELDS = [("input_mean", 5), ("input_std", 5), ("output_mean", 6), ("output_std", 6)]


def load_nn_blob(path: Path | str = DEFAULT_MODEL_PATH) -> AircNNBlob:
    path = Path(path) if path else DEFAULT_MODEL_PATH
    ck = torch.load(str(path), map_location="cpu", weights_only=True)
    sd = ck["model_state_dict"]
    blob = AircNNBlob()
    for fld, key, shape in _BLOB_FIELDS:
        t = sd[key].detach().to(torch.float32).contiguous()
        if tuple(t.shape) != shape:
            raise RuntimeError(f"{path}: {key} has shape {tuple(t.shape)}, expected {shape} "
                               "(ScaledModel 5->16->32->tanh->6)")
        arr = getattr(blob, fld)
        arr[:] = t.reshape(-1).tolist()
    for fld, n in _SCALER_FIELDS:
        t = ck[fld].detach().to(torch.float32).reshape(-1)
        if t.numel() != n:
            raise RuntimeError(f"{path}: {fld} has {t.numel()} entries, expected {n}")
        getattr(blob, fld)[:] = t.tolist()
    return blob


def nn_blob_from_arrays(W1, b1, W2, b2, W3, b3, in_mean, in_std, out_mean, out_std) -> AircNNBlob:
    """For tests: a blob from arbitrary weights of the same architecture."""
    blob = AircNNBlob()
    for fld, val in zip(["W1", "b1", "W2", "b2", "W3", "b3", "input_mean", "input_std",
                         "output_mean", "output_std"],
                        [W1, b1, W2, b2, W3, b3, in_mean, in_std, out_mean, out_std]):
        a = np.ascontiguousarray(np.asarray(val, np.float32)).reshape(-1)
        getattr(blob, fld)[:] = a.tolist()
    return blob


def make_params(cfg: AircraftConfiguration, com: Sequence[float] | None = None,
                gravity: Sequence[float] = (0.0, 0.0, 9.81), epsilon: float = 1e-6,
                stall_scaling: bool = False) -> AircParams:
    com = list(cfg.aero_centre_offset if com is None else com)
    inertia, inv = inertia_about_com(cfg, com)
    p = AircParams()
    p.mass, p.span, p.chord = float(cfg.mass), float(cfg.span), float(cfg.chord)
    p.reference_area = float(cfg.reference_area)
    p.com[:] = [float(c) for c in com]
    p.inertia[:] = inertia.reshape(-1).tolist()
    p.inertia_inv[:] = inv.reshape(-1).tolist()
    p.gravity[:] = [float(g) for g in gravity]
    p.epsilon = float(epsilon)
    p.stall_scaling = int(bool(stall_scaling))
    return p


def make_cost_params(goal=(150.0, 0.0), vel_param=1.0, w_goal=1000.0, w_du=100.0, du_eps=1e-2,
                     w_height=1.0, w_speed=0.01, w_vx=1000.0, w_vyz=1000.0) -> AircCostParams:
    """Weights of the primary MPC loss, /root/reference/main/control/control.py:35-70."""
    c = AircCostParams()
    c.goal[:] = [float(goal[0]), float(goal[1])]
    c.vel_param, c.w_goal, c.w_du, c.du_eps = float(vel_param), float(w_goal), float(w_du), float(du_eps)
    c.w_height, c.w_speed, c.w_vx, c.w_vyz = float(w_height), float(w_speed), float(w_vx), float(w_vyz)
    return c


def make_sweep_params(seed: int = 0, index_offset: int = 0, sigma0=(1.0, 1.0, 1.0, 0.1),
                      sigma=(0.25, 0.25, 0.25, 0.025), lo=(-5.0, -5.0, -5.0, 0.0),
                      hi=(5.0, 5.0, 5.0, 1.0), ckpt_interval: int = 0) -> AircSweepParams:
    """Sampler of the sampling-MPC sweep (include/aircraft_b200.h, AircSweepParams): per live
    control (aileron, elevator, rudder, flaps) the std of the initial offset and of the random-walk
    increment, and the actuator limits (/root/reference/src/aircraft/control/aircraft.py:24,29-41:
    +-5 deg, flaps in [0,1])."""
    s = AircSweepParams()
    s.seed, s.index_offset = int(seed) & 0xFFFFFFFFFFFFFFFF, int(index_offset)
    s.sigma0[:] = [float(v) for v in sigma0]
    s.sigma[:] = [float(v) for v in sigma]
    s.lo[:] = [float(v) for v in lo]
    s.hi[:] = [float(v) for v in hi]
    s.ckpt_interval = int(ckpt_interval)
    return s


# ---------------------------------------------------------------------------------------------
# The registry's other coefficient models (coefficient_models.py:41-135), for airc_create_model.
NPOLY = 34


def _canonical_powers():
    """sklearn PolynomialFeatures(degree=3, include_bias=False) exponents for 4 inputs, in order."""
    import itertools
    rows = []
    for deg in (1, 2, 3):
        for comb in itertools.combinations_with_replacement(range(4), deg):
            e = [0, 0, 0, 0]
            for i in comb:
                e[i] += 1
            rows.append(e)
    return np.asarray(rows, np.int64)


class _PickleStub:
    def __init__(self, *a, **k):
        pass

    def __setstate__(self, state):
        self.state = state


def load_poly_model(path, rudder_moment_arm: float = 0.5) -> np.ndarray:
    """float32[211] for AIRC_MODEL_POLY from the reference's data/networks/fitted_models_casadi.pkl
    (main/surrogates/polynomial_model.py:228-271) or from a .npz with powers/coef/intercept
    (tests/golden/make_poly_golden.py).  The pickle's CasADi Functions are generated from the very
    numbers stored next to them ('fitted_models': coef, intercept), so CasADi and sklearn objects are
    replaced by stubs while unpickling and only the plain arrays are read."""
    import pickle
    path = Path(path)
    if path.suffix == ".npz":
        d = np.load(path)
        if not {"powers", "coef", "intercept"} <= set(d.files):
            raise RuntimeError(f"{path}: not a polynomial-model table (powers, coef, intercept)")
        powers, coef, icpt = d["powers"], d["coef"], d["intercept"]
    else:
        class U(pickle.Unpickler):
            def find_class(self, module, name):
                if module.split(".")[0] in ("casadi", "sklearn"):
                    return type(name, (_PickleStub,), {"__module__": module})
                return super().find_class(module, name)
        with open(path, "rb") as f:
            fm = U(f).load()["fitted_models"]
        names = ["CX", "CY", "CZ", "Cl", "Cm", "Cn"]
        if list(fm) != names:
            raise RuntimeError(f"unexpected coefficient names in {path}: {list(fm)}")
        coef = np.stack([np.asarray(fm[k]["coef"], np.float64).reshape(-1) for k in names])
        icpt = np.array([float(fm[k]["intercept"]) for k in names])
        st = getattr(fm["CX"]["poly"], "state", {}) or {}
        if st.get("degree", 3) != 3 or st.get("include_bias", False) or st.get("n_features_in_", 4) != 4 or \
                st.get("interaction_only", False):
            raise RuntimeError(f"{path}: not the degree-3, 4-input, bias-free polynomial this build evaluates")
        powers = _canonical_powers()
    if coef.shape != (6, NPOLY) or not np.array_equal(np.asarray(powers), _canonical_powers()):
        raise RuntimeError(f"{path}: monomial set differs from degree-3 PolynomialFeatures over 4 inputs")
    out = np.empty(6 * (NPOLY + 1) + 1, np.float32)
    out[:-1] = np.concatenate([coef, icpt[:, None]], axis=1).reshape(-1)
    out[-1] = rudder_moment_arm
    return out


def load_linear_model(path) -> np.ndarray:
    """float32[36] for AIRC_MODEL_LINEAR: the six rows of data/networks/linearised.csv
    (q, alpha, beta, aileron, elevator, intercept), as LinearModel reads them (coefficient_models.py:80-84)."""
    path = Path(path)
    if path.suffix == ".npz":      # tests/golden/make_poly_golden.py
        a = np.asarray(np.load(path)["table"], np.float64)
    else:
        a = np.loadtxt(path, delimiter=",", skiprows=1, dtype=np.float64)
    if a.shape != (6, 6):
        raise RuntimeError(f"{path}: expected a 6x6 coefficient table, got {a.shape}")
    return a.astype(np.float32).reshape(-1)

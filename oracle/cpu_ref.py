"""ctypes loaders for the two C++ checkers built by oracle/Makefile -- TEST INFRASTRUCTURE.

  libairc_cpu_ref.so    float64 as-written restatement of the reference path, OpenMP
                        (oracle/cpu_ref.cpp): the timed CPU baseline, validated against
                        oracle/aircraft_oracle.py
  libairc_host_emul.so  the product's own FP32 device formulas compiled with g++
                        (oracle/host_emul.cpp): lets the CPU-only tests check the kernel algebra
"""
from __future__ import annotations

import ctypes as C
import subprocess
from pathlib import Path

import numpy as np

from aircraft_b200._cabi import AircCostParams, AircNNBlob, AircParams, AircSweepParams

HERE = Path(__file__).resolve().parent
_f64 = np.ctypeslib.ndpointer(np.float64, flags="C_CONTIGUOUS")
_f32 = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
_PP, _NP = C.POINTER(AircParams), C.POINTER(AircNNBlob)


def _load(name: str) -> C.CDLL:
    p = HERE / name
    if not p.exists():
        r = subprocess.run(["make", "-C", str(HERE)], capture_output=True, text=True)
        if r.returncode != 0 or not p.exists():
            raise RuntimeError(f"cannot build {name}:\n{r.stdout}{r.stderr}")
    return C.CDLL(str(p))


class CpuRef:
    """float64 OpenMP restatement (cpu_ref.cpp).  Arrays are AoS: x[n,13], u[n,7]."""

    def __init__(self, params: AircParams, blob: AircNNBlob, threads: int = 0):
        self.lib = _load("libairc_cpu_ref.so")
        self.params, self.blob = params, blob
        self.lib.cpuref_max_threads.restype = C.c_int
        self.max_threads = int(self.lib.cpuref_max_threads())
        self.threads = threads if threads > 0 else self.max_threads
        vp = C.c_void_p
        self.lib.cpuref_state_derivative.argtypes = [_PP, _NP, _f64, _f64, C.c_int64, _f64, C.c_int]
        self.lib.cpuref_state_update.argtypes = [_PP, _NP, _f64, _f64, C.c_double, C.c_int64, C.c_int,
                                                 C.c_uint32, _f64, C.c_int]
        self.lib.cpuref_rollout.argtypes = [_PP, _NP, _f64, _f64, C.c_double, C.c_int64, C.c_int,
                                            C.c_int, C.c_uint32, vp, vp, C.c_int]
        self.lib.cpuref_linearize.argtypes = [_PP, _NP, _f64, _f64, C.c_double, C.c_int64, C.c_int,
                                              C.c_uint32, _f64, _f64, _f64, C.c_int]
        self.lib.cpuref_linearize_dt.argtypes = [_PP, _NP, _f64, _f64, C.c_double, C.c_int64, C.c_int,
                                                 C.c_uint32, _f64, _f64, _f64, _f64, C.c_int]

        _CPp, _SPp = C.POINTER(AircCostParams), C.POINTER(AircSweepParams)
        self.lib.cpuref_rollout_cost.argtypes = [_PP, _NP, _f64, _f64, C.c_double, C.c_int64, C.c_int,
                                                 C.c_int, C.c_uint32, _CPp, _f64, C.c_int]
        self.lib.cpuref_sweep_controls.argtypes = [_SPp, _f32, C.c_int, C.c_int64, C.c_int64, _f32, C.c_int]
        self.lib.cpuref_sweep.argtypes = [_PP, _NP, _f64, _f32, C.c_double, C.c_int64, C.c_int, C.c_int,
                                          C.c_uint32, _SPp, _CPp, vp, C.POINTER(C.c_double),
                                          C.POINTER(C.c_int64), C.c_int]

    def rollout_cost(self, x0, U, dt, cost_params, substeps=1, normalise=False):
        """Primary MPC loss (main/control/control.py:35-70) of explicit control sequences."""
        x0, U = np.ascontiguousarray(x0, np.float64), np.ascontiguousarray(U, np.float64)
        B = x0.shape[0]
        flags = (1 if normalise else 0) | (2 if U.ndim == 2 else 0)
        H = U.shape[0] if U.ndim == 2 else U.shape[1]
        out = np.empty(B)
        self.lib.cpuref_rollout_cost(*self._pb(), x0, U, float(dt), B, H, substeps, flags,
                                     C.byref(cost_params), out, self.threads)
        return out

    def sweep_controls(self, sweep_params, U_nom, first, count):
        """The sampled control sequences [count,H,7] (FP32) of trajectories [first, first+count)."""
        U_nom = np.ascontiguousarray(U_nom, np.float32)
        H = U_nom.shape[0]
        U = np.empty((count, H, 7), np.float32)
        self.lib.cpuref_sweep_controls(C.byref(sweep_params), U_nom, H, first, count, U, self.threads)
        return U

    def sweep(self, x0, U_nom, dt, B, sweep_params, cost_params, substeps=1, normalise=False,
              return_costs=True):
        """Whole sampling sweep on the CPU -> (min cost, arg-min (local index), cost[B] or None)."""
        x0 = np.ascontiguousarray(x0, np.float64).reshape(13)
        U_nom = np.ascontiguousarray(U_nom, np.float32)
        H = U_nom.shape[0]
        cost = np.empty(B) if return_costs else None
        bc, bi = C.c_double(), C.c_int64()
        self.lib.cpuref_sweep(*self._pb(), x0, U_nom, float(dt), B, H, substeps, 1 if normalise else 0,
                              C.byref(sweep_params), C.byref(cost_params),
                              cost.ctypes.data if return_costs else None, C.byref(bc), C.byref(bi),
                              self.threads)
        return bc.value, bi.value, cost

    def _pb(self):
        return C.byref(self.params), C.byref(self.blob)

    def state_derivative(self, x, u):
        x, u = np.ascontiguousarray(x, np.float64), np.ascontiguousarray(u, np.float64)
        out = np.empty_like(x)
        self.lib.cpuref_state_derivative(*self._pb(), x, u, x.shape[0], out, self.threads)
        return out

    def state_update(self, x, u, dt, substeps=1, normalise=False):
        x, u = np.ascontiguousarray(x, np.float64), np.ascontiguousarray(u, np.float64)
        out = np.empty_like(x)
        self.lib.cpuref_state_update(*self._pb(), x, u, float(dt), x.shape[0], substeps,
                                     1 if normalise else 0, out, self.threads)
        return out

    def rollout(self, x0, U, dt, substeps=1, normalise=False, trajectory=True):
        x0, U = np.ascontiguousarray(x0, np.float64), np.ascontiguousarray(U, np.float64)
        B = x0.shape[0]
        flags = (1 if normalise else 0) | (2 if U.ndim == 2 else 0)
        H = U.shape[0] if U.ndim == 2 else U.shape[1]
        X = np.empty((B, H + 1, 13)) if trajectory else None
        xf = np.empty((B, 13))
        self.lib.cpuref_rollout(*self._pb(), x0, U, float(dt), B, H, substeps, flags,
                                X.ctypes.data if trajectory else None, xf.ctypes.data, self.threads)
        return X if trajectory else xf

    def linearize(self, x, u, dt, substeps=1, normalise=False, want_dt=False):
        x, u = np.ascontiguousarray(x, np.float64), np.ascontiguousarray(u, np.float64)
        n = x.shape[0]
        xn, A, Bm = np.empty((n, 13)), np.empty((n, 13, 13)), np.empty((n, 13, 7))
        if want_dt:
            Dt = np.empty((n, 13))
            self.lib.cpuref_linearize_dt(*self._pb(), x, u, float(dt), n, substeps, 1 if normalise else 0,
                                         xn, A, Bm, Dt, self.threads)
            return xn, A, Bm, Dt
        self.lib.cpuref_linearize(*self._pb(), x, u, float(dt), n, substeps, 1 if normalise else 0,
                                  xn, A, Bm, self.threads)
        return xn, A, Bm


class HostEmul:
    """The product's FP32 formulas on the CPU (host_emul.cpp).  AoS float32 arrays."""

    def __init__(self, params: AircParams, blob: AircNNBlob):
        self.lib = _load("libairc_host_emul.so")
        self.params, self.blob = params, blob
        self.lib.emul_surrogate.argtypes = [_NP, _f32, C.c_int64, _f32]
        self.lib.emul_state_derivative.argtypes = [_PP, _NP, _f32, _f32, C.c_int64, _f32]
        self.lib.emul_state_update.argtypes = [_PP, _NP, _f32, _f32, C.c_float, C.c_int64, C.c_int,
                                               C.c_uint32, _f32]
        self.lib.emul_linearize.argtypes = [_PP, _NP, _f32, _f32, C.c_float, C.c_int64, C.c_int,
                                            C.c_uint32, _f32, _f32, _f32]

        self.lib.emul_linearize_packed.argtypes = self.lib.emul_linearize.argtypes
        self.lib.emul_linearize_packed_dt.argtypes = self.lib.emul_linearize.argtypes + [_f32]
        self.lib.emul_sweep_controls.argtypes = [C.POINTER(AircSweepParams), _f32, C.c_int, C.c_int64,
                                                 C.c_int64, _f32]

    def sweep_controls(self, sweep_params, U_nom, first, count):
        """The product's own sampler code (airc_sampler.cuh) compiled for the host."""
        U_nom = np.ascontiguousarray(U_nom, np.float32)
        H = U_nom.shape[0]
        U = np.empty((count, H, 7), np.float32)
        self.lib.emul_sweep_controls(C.byref(sweep_params), U_nom, H, first, count, U)
        return U

    def linearize_packed(self, x, u, dt, substeps=1, normalise=False, want_dt=False):
        """The cooperative kernel's arithmetic (packed two-column tangents), one lane per node."""
        x, u = np.ascontiguousarray(x, np.float32), np.ascontiguousarray(u, np.float32)
        n = x.shape[0]
        xn = np.empty((n, 13), np.float32)
        A, Bm = np.zeros((n, 13, 13), np.float32), np.zeros((n, 13, 7), np.float32)
        if want_dt:
            Dt = np.zeros((n, 13), np.float32)
            self.lib.emul_linearize_packed_dt(C.byref(self.params), C.byref(self.blob), x, u, float(dt), n,
                                              substeps, 1 if normalise else 0, xn, A, Bm, Dt)
            return xn, A, Bm, Dt
        self.lib.emul_linearize_packed(C.byref(self.params), C.byref(self.blob), x, u, float(dt), n,
                                       substeps, 1 if normalise else 0, xn, A, Bm)
        return xn, A, Bm

    def linearize_tile(self, x, u, dt, normalise=False, want_dt=False):
        """The tile kernel's arithmetic (airc_jac.cuh: parked f_x / f_u blocks + tangent recursion), one substep."""
        x, u = np.ascontiguousarray(x, np.float32), np.ascontiguousarray(u, np.float32)
        n = x.shape[0]
        xn = np.empty((n, 13), np.float32)
        A, Bm = np.zeros((n, 13, 13), np.float32), np.zeros((n, 13, 7), np.float32)
        Dt = np.zeros((n, 13), np.float32) if want_dt else None
        self.lib.emul_linearize_tile.argtypes = [_PP, _NP, _f32, _f32, C.c_float, C.c_int64, C.c_uint32, _f32, _f32,
                                                 _f32, C.c_void_p]
        self.lib.emul_linearize_tile(C.byref(self.params), C.byref(self.blob), x, u, float(dt), n,
                                     1 if normalise else 0, xn, A, Bm, None if Dt is None else Dt.ctypes.data)
        return (xn, A, Bm, Dt) if want_dt else (xn, A, Bm)

    def model_state_update(self, model, table, x, u, dt, substeps=1, normalise=False):
        """airc_alt.cuh (poly / linear / default coefficient models) through the product's RK4."""
        x, u = np.ascontiguousarray(x, np.float32), np.ascontiguousarray(u, np.float32)
        out = np.empty_like(x)
        self.lib.emul_model_state_update.argtypes = [_PP, C.c_int, C.c_void_p, _f32, _f32, C.c_float, C.c_int64,
                                                     C.c_int, C.c_uint32, _f32]
        self.lib.emul_model_state_update(C.byref(self.params), model, None if table is None else table.ctypes.data,
                                         x, u, float(dt), x.shape[0], substeps, 1 if normalise else 0, out)
        return out

    def model_linearize(self, model, table, x, u, dt, substeps=1, normalise=False):
        """linearize_jet_kernel's arithmetic for the table models: (x+, A, B, dPhi/ddt)."""
        x, u = np.ascontiguousarray(x, np.float32), np.ascontiguousarray(u, np.float32)
        n = x.shape[0]
        xn, A, Bm, Dt = (np.empty((n,) + s, np.float32) for s in ((13,), (13, 13), (13, 7), (13,)))
        self.lib.emul_model_linearize.argtypes = [_PP, C.c_int, C.c_void_p, _f32, _f32, C.c_float, C.c_int64, C.c_int,
                                                  C.c_uint32, _f32, _f32, _f32, _f32]
        self.lib.emul_model_linearize(C.byref(self.params), model, None if table is None else table.ctypes.data, x, u,
                                      float(dt), n, substeps, 1 if normalise else 0, xn, A, Bm, Dt)
        return xn, A, Bm, Dt

    def model_hessian(self, model, table, x, u, lam, dt, substeps=1, normalise=False):
        x, u, lam = (np.ascontiguousarray(a, np.float32) for a in (x, u, lam))
        n = x.shape[0]
        H, g, v = np.empty((n, 21, 21), np.float32), np.empty((n, 21), np.float32), np.empty(n, np.float32)
        self.lib.emul_model_hessian.argtypes = [_PP, C.c_int, C.c_void_p, _f32, _f32, _f32, C.c_float, C.c_int64,
                                                C.c_int, C.c_uint32, _f32, _f32, _f32]
        self.lib.emul_model_hessian(C.byref(self.params), model, None if table is None else table.ctypes.data, x, u,
                                    lam, float(dt), n, substeps, 1 if normalise else 0, H, g, v)
        return H, g, v

    def model_state_derivative(self, model, table, x, u):
        x, u = np.ascontiguousarray(x, np.float32), np.ascontiguousarray(u, np.float32)
        out = np.empty_like(x)
        self.lib.emul_model_state_derivative.argtypes = [_PP, C.c_int, C.c_void_p, _f32, _f32, C.c_int64, _f32]
        self.lib.emul_model_state_derivative(C.byref(self.params), model, None if table is None else table.ctypes.data,
                                             x, u, x.shape[0], out)
        return out

    def hessian(self, x, u, lam, dt, substeps=1, normalise=False):
        """K3 arithmetic (airc_hess.cuh) on the host: H[n,21,21], grad[n,21], value[n]."""
        x, u, lam = (np.ascontiguousarray(a, np.float32) for a in (x, u, lam))
        n = x.shape[0]
        H, g, v = np.empty((n, 21, 21), np.float32), np.empty((n, 21), np.float32), np.empty(n, np.float32)
        self.lib.emul_hessian.argtypes = [_PP, _NP, _f32, _f32, _f32, C.c_float, C.c_int64, C.c_int,
                                          C.c_uint32, _f32, _f32, _f32]
        self.lib.emul_hessian(C.byref(self.params), C.byref(self.blob), x, u, lam, float(dt), n, substeps,
                              1 if normalise else 0, H, g, v)
        return H, g, v

    def hessian_shared(self, x, u, lam, dt, substeps=1, normalise=False):
        """K4 arithmetic for short horizons (jets, the network's sums shared by the 64 threads of a node)."""
        x, u, lam = (np.ascontiguousarray(a, np.float32) for a in (x, u, lam))
        n = x.shape[0]
        H, g, v = np.empty((n, 21, 21), np.float32), np.empty((n, 21), np.float32), np.empty(n, np.float32)
        self.lib.emul_hessian_shared.argtypes = [_PP, _NP, _f32, _f32, _f32, C.c_float, C.c_int64, C.c_int,
                                                 C.c_uint32, _f32, _f32, _f32]
        self.lib.emul_hessian_shared(C.byref(self.params), C.byref(self.blob), x, u, lam, float(dt), n, substeps,
                                     1 if normalise else 0, H, g, v)
        return H, g, v

    def hessian2(self, x, u, lam, dt, substeps=1, normalise=False):
        """K4 arithmetic (airc_hess2.cuh, forward over reverse) on the host: H[n,21,21], grad[n,21], value[n]."""
        x, u, lam = (np.ascontiguousarray(a, np.float32) for a in (x, u, lam))
        n = x.shape[0]
        H, g, v = np.empty((n, 21, 21), np.float32), np.empty((n, 21), np.float32), np.empty(n, np.float32)
        self.lib.emul_hessian2.argtypes = [_PP, _NP, _f32, _f32, _f32, C.c_float, C.c_int64, C.c_int,
                                           C.c_uint32, _f32, _f32, _f32]
        self.lib.emul_hessian2(C.byref(self.params), C.byref(self.blob), x, u, lam, float(dt), n, substeps,
                               1 if normalise else 0, H, g, v)
        return H, g, v

    def surrogate(self, inp):
        inp = np.ascontiguousarray(inp, np.float32)
        out = np.empty((inp.shape[0], 6), np.float32)
        self.lib.emul_surrogate(C.byref(self.blob), inp, inp.shape[0], out)
        return out

    def state_derivative(self, x, u):
        x, u = np.ascontiguousarray(x, np.float32), np.ascontiguousarray(u, np.float32)
        out = np.empty_like(x)
        self.lib.emul_state_derivative(C.byref(self.params), C.byref(self.blob), x, u, x.shape[0], out)
        return out

    def state_update(self, x, u, dt, substeps=1, normalise=False):
        x, u = np.ascontiguousarray(x, np.float32), np.ascontiguousarray(u, np.float32)
        out = np.empty_like(x)
        self.lib.emul_state_update(C.byref(self.params), C.byref(self.blob), x, u, float(dt),
                                   x.shape[0], substeps, 1 if normalise else 0, out)
        return out

    def linearize(self, x, u, dt, substeps=1, normalise=False):
        x, u = np.ascontiguousarray(x, np.float32), np.ascontiguousarray(u, np.float32)
        n = x.shape[0]
        xn = np.empty((n, 13), np.float32)
        A, Bm = np.zeros((n, 13, 13), np.float32), np.zeros((n, 13, 7), np.float32)
        self.lib.emul_linearize(C.byref(self.params), C.byref(self.blob), x, u, float(dt), n, substeps,
                                1 if normalise else 0, xn, A, Bm)
        return xn, A, Bm

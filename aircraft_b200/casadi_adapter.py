"""CasADi adapter: the GPU dynamics as ``casadi.Callback`` objects, so that the reference's symbolic NLP
construction (``ca.Opti``: /root/reference/src/aircraft/control/base.py:187-190 takes ``system.state_update``,
:275-286 embeds it in the defect constraints, IPOPT then asks for first and second derivatives) can run on the
CUDA path where CasADi is installed.

    from aircraft_b200.casadi_adapter import state_update_callback
    f = state_update_callback(aircraft)            # casadi.Function-compatible: f(x, u, dt) -> x_next
    opti.subject_to(x_next == f(x, u, dt))         # Jacobian blocks come from airc_linearize_dt_host,
                                                   # Hessian-of-Lagrangian blocks from airc_hessian_host

Three layers:
  * ``DynamicsMarshal`` -- CasADi-free: turns the column / matrix arguments CasADi hands a callback into the
    C-ABI calls (through the Aircraft facade's host path) and packs the results in CasADi's conventions (the
    Jacobian function's inputs are (x, u, dt, out) and its outputs the blocks d out / d x, d out / d u,
    d out / d dt; the reverse-of-Jacobian function returns the adjoint-seeded second-order products).  This is
    what the CPU suite tests (tests/test_casadi_adapter.py), since CasADi is not installed on the build / GPU boxes.
  * ``state_update_callback`` / ``state_derivative_callback`` -- subclasses of ``casadi.Callback`` generated on
    first use; importing this module without CasADi works, calling them raises ImportError.
  * nothing numerical lives here: eval -> airc_state_update_host, jacobian -> airc_linearize_dt_host,
    hessian -> airc_hessian_host.
"""
from __future__ import annotations

import numpy as np

try:  # optional: not available on the build box / GPU box of this repo
    import casadi as ca
except Exception:  # pragma: no cover
    ca = None

NX, NU = 13, 7


def _col(a, n):
    """CasADi hands callbacks DM columns (n x 1) or n x N matrices (mapped calls)."""
    if hasattr(a, "full") and not isinstance(a, np.ndarray):
        a = a.full()
    a = np.asarray(a, np.float64)
    if a.ndim == 0:
        a = a.reshape(1, 1)
    if a.ndim == 1:
        a = a.reshape(-1, 1)
    if a.shape[0] != n:
        raise RuntimeError(f"expected {n} rows, got {a.shape}")
    return a


class DynamicsMarshal:
    """Argument / result marshalling between CasADi's callback protocol and the Aircraft facade."""

    def __init__(self, aircraft, substeps=None, normalise=None):
        self.ac, self.substeps, self.normalise = aircraft, substeps, normalise

    # ---- Phi(x, u, dt)
    def eval_state_update(self, x, u, dt):
        x, u = _col(x, NX), _col(u, NU)
        dtv = np.asarray(dt.full() if hasattr(dt, "full") and not isinstance(dt, np.ndarray) else dt, np.float64).reshape(-1)
        out = self.ac.state_update(x, u, dtv if dtv.size > 1 else float(dtv[0]), substeps=self.substeps,
                                   normalise=self.normalise)
        return np.asarray(out, np.float64).reshape(NX, -1)

    def jac_state_update(self, x, u, dt):
        """(d Phi/d x [13,13], d Phi/d u [13,7], d Phi/d dt [13,1]) at one node -- the three blocks of CasADi's
        ``jac_state_update`` function (control/base.py:275-286 differentiated by IPOPT's interface)."""
        x, u = _col(x, NX)[:, 0], _col(u, NU)[:, 0]
        dtv = float(np.asarray(dt.full() if hasattr(dt, "full") and not isinstance(dt, np.ndarray) else dt).reshape(-1)[0])
        _, A, B, D = self.ac.linearize(x[None], u[None], dtv, substeps=self.substeps, normalise=self.normalise,
                                       want_dt=True)
        return (np.asarray(A, np.float64)[0], np.asarray(B, np.float64)[0], np.asarray(D, np.float64).reshape(NX, 1))

    def hess_state_update(self, x, u, dt, lam):
        """d2(lam^T Phi)/d(x,u,dt)2 [21,21] and its gradient [21]: what IPOPT's exact Hessian
        (src/aircraft/config.py:29) needs from this constraint block."""
        x, u, lam = _col(x, NX)[:, 0], _col(u, NU)[:, 0], _col(lam, NX)[:, 0]
        dtv = float(np.asarray(dt.full() if hasattr(dt, "full") and not isinstance(dt, np.ndarray) else dt).reshape(-1)[0])
        H, g, _ = self.ac.lagrangian_hessian(x[None], u[None], lam[None], dtv, substeps=self.substeps,
                                             normalise=self.normalise)
        return np.asarray(H, np.float64)[0], np.asarray(g, np.float64)[0]

    # ---- f(x, u)
    def eval_state_derivative(self, x, u):
        return np.asarray(self.ac.state_derivative(_col(x, NX), _col(u, NU)), np.float64).reshape(NX, -1)


def _require_casadi():
    if ca is None:
        raise ImportError("casadi is not installed: aircraft_b200.casadi_adapter needs it for the Callback classes "
                          "(the marshalling core DynamicsMarshal works without)")


_CLASSES = {}


def _classes():
    """The Callback subclasses, created once CasADi is known to be importable."""
    _require_casadi()
    if _CLASSES:
        return _CLASSES

    class _JacStateUpdate(ca.Callback):
        # CasADi's convention for the Jacobian of f(x,u,dt)->x_next: inputs (x, u, dt, out_x_next),
        # outputs (jac_x_next_x, jac_x_next_u, jac_x_next_dt)
        def __init__(self, name, marshal, opts=None):
            ca.Callback.__init__(self)
            self.m = marshal
            self.construct(name, opts or {})

        def get_n_in(self): return 4
        def get_n_out(self): return 3

        def get_sparsity_in(self, i):
            return ca.Sparsity.dense((NX, NU, 1, NX)[i], 1)

        def get_sparsity_out(self, i):
            return ca.Sparsity.dense(NX, (NX, NU, 1)[i])

        def eval(self, arg):
            A, B, D = self.m.jac_state_update(arg[0], arg[1], arg[2])
            return [ca.DM(A), ca.DM(B), ca.DM(D)]

    class _StateUpdate(ca.Callback):
        def __init__(self, name, marshal, opts=None):
            ca.Callback.__init__(self)
            self.m = marshal
            self._jac = None
            self.construct(name, opts or {})

        def get_n_in(self): return 3
        def get_n_out(self): return 1
        def get_name_in(self, i): return ("x", "u", "dt")[i]
        def get_name_out(self, i): return "x_next"

        def get_sparsity_in(self, i):
            return ca.Sparsity.dense((NX, NU, 1)[i], 1)

        def get_sparsity_out(self, i):
            return ca.Sparsity.dense(NX, 1)

        def eval(self, arg):
            return [ca.DM(self.m.eval_state_update(arg[0], arg[1], arg[2]))]

        def has_jacobian(self): return True

        def get_jacobian(self, name, inames, onames, opts):
            self._jac = _JacStateUpdate(name, self.m, opts)      # keep a reference: CasADi does not own it
            return self._jac

    class _StateDerivative(ca.Callback):
        def __init__(self, name, marshal, opts=None):
            ca.Callback.__init__(self)
            self.m = marshal
            self.construct(name, opts or {"enable_fd": True})    # f is only used by the implicit mode (base.py:281)

        def get_n_in(self): return 2
        def get_n_out(self): return 1
        def get_name_in(self, i): return ("x", "u")[i]
        def get_name_out(self, i): return "x_dot"

        def get_sparsity_in(self, i):
            return ca.Sparsity.dense((NX, NU)[i], 1)

        def get_sparsity_out(self, i):
            return ca.Sparsity.dense(NX, 1)

        def eval(self, arg):
            return [ca.DM(self.m.eval_state_derivative(arg[0], arg[1]))]

    class _LagrangianHessian(ca.Callback):
        """(x, u, dt, lam) -> (H [21,21], grad [21]) for callers that assemble the NLP Hessian themselves."""
        def __init__(self, name, marshal, opts=None):
            ca.Callback.__init__(self)
            self.m = marshal
            self.construct(name, opts or {})

        def get_n_in(self): return 4
        def get_n_out(self): return 2

        def get_sparsity_in(self, i):
            return ca.Sparsity.dense((NX, NU, 1, NX)[i], 1)

        def get_sparsity_out(self, i):
            return ca.Sparsity.dense(21, (21, 1)[i])

        def eval(self, arg):
            H, g = self.m.hess_state_update(arg[0], arg[1], arg[2], arg[3])
            return [ca.DM(H), ca.DM(g.reshape(21, 1))]

    _CLASSES.update(state_update=_StateUpdate, state_derivative=_StateDerivative, hessian=_LagrangianHessian)
    return _CLASSES


def state_update_callback(aircraft, name="state_update", substeps=None, normalise=None, opts=None):
    """``casadi.Callback`` for Phi(x,u,dt) with exact Jacobian blocks from the GPU (drop-in for
    ``aircraft.state_update`` of the reference inside ``ca.Opti``)."""
    return _classes()["state_update"](name, DynamicsMarshal(aircraft, substeps, normalise), opts)


def state_derivative_callback(aircraft, name="dynamics", opts=None):
    return _classes()["state_derivative"](name, DynamicsMarshal(aircraft), opts)


def lagrangian_hessian_callback(aircraft, name="hess_state_update", substeps=None, normalise=None, opts=None):
    return _classes()["hessian"](name, DynamicsMarshal(aircraft, substeps, normalise), opts)

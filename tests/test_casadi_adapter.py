"""casadi_adapter.py: the CasADi-free marshalling core (CasADi is not installed on the build / GPU boxes) and,
on the GPU, its numbers against the float64 oracle.  The Callback classes themselves need CasADi and are only
checked for a clean ImportError here."""
import numpy as np
import pytest

from _util import PARITY_TOL, col_err, oracle_for, t64
from aircraft_b200 import casadi_adapter as cad
from aircraft_b200.config import TRIM_COM
from aircraft_b200.dm import DM
from aircraft_b200.workloads import make_controls, make_states


class FakeAircraft:
    """Records what the marshal asks of the facade and hands back arrays of the facade's shapes."""

    def __init__(self):
        self.calls = []

    def state_update(self, x, u, dt, substeps=None, normalise=None):
        self.calls.append(("state_update", x.shape, u.shape, dt, substeps, normalise))
        return np.tile(np.arange(13.0)[:, None], (1, x.shape[1])) + dt if np.ndim(dt) == 0 else np.zeros((13, x.shape[1]))

    def state_derivative(self, x, u):
        self.calls.append(("state_derivative", x.shape, u.shape))
        return np.ones((13, x.shape[1]))

    def linearize(self, x, u, dt, substeps=None, normalise=None, want_dt=False):
        self.calls.append(("linearize", x.shape, u.shape, dt, want_dt))
        return np.zeros((1, 13)), np.full((1, 13, 13), 2.0), np.full((1, 13, 7), 3.0), np.full((1, 13), 4.0)

    def lagrangian_hessian(self, x, u, lam, dt, substeps=None, normalise=None):
        self.calls.append(("hessian", x.shape, u.shape, lam.shape, dt))
        return np.full((1, 21, 21), 5.0), np.full((1, 21), 6.0), None


def test_marshalling_without_casadi():
    ac = FakeAircraft()
    m = cad.DynamicsMarshal(ac, substeps=1, normalise=True)
    x, u = DM(np.arange(13.0)), np.arange(7.0)            # a DM column and a plain vector, as CasADi / numpy callers pass
    out = m.eval_state_update(x, u, DM(0.01))
    assert out.shape == (13, 1) and ac.calls[-1] == ("state_update", (13, 1), (7, 1), 0.01, 1, True)
    out = m.eval_state_update(np.zeros((13, 5)), np.zeros((7, 5)), 0.02)      # mapped call: 5 columns
    assert out.shape == (13, 5)
    A, B, D = m.jac_state_update(x, u, 0.01)
    assert A.shape == (13, 13) and B.shape == (13, 7) and D.shape == (13, 1) and A[0, 0] == 2.0 and D[0, 0] == 4.0
    assert ac.calls[-1] == ("linearize", (1, 13), (1, 7), 0.01, True)
    H, g = m.hess_state_update(x, u, np.float64(0.01), np.ones(13))
    assert H.shape == (21, 21) and g.shape == (21,) and ac.calls[-1][0] == "hessian"
    assert m.eval_state_derivative(x, u).shape == (13, 1)
    with pytest.raises(RuntimeError):
        m.eval_state_update(np.zeros(12), u, 0.01)


@pytest.mark.skipif(cad.ca is not None, reason="CasADi is installed here: the Callback path is live")
def test_callbacks_need_casadi():
    with pytest.raises(ImportError):
        cad.state_update_callback(FakeAircraft())
    with pytest.raises(ImportError):
        cad.lagrangian_hessian_callback(FakeAircraft())


@pytest.mark.skipif(cad.ca is None, reason="needs CasADi")
def test_callback_inside_casadi():            # pragma: no cover - runs only where CasADi exists
    import casadi as ca
    ac = FakeAircraft()
    f = cad.state_update_callback(ac)
    assert f.size1_in(0) == 13 and f.size1_in(1) == 7
    x, u, dt = ca.MX.sym("x", 13), ca.MX.sym("u", 7), ca.MX.sym("dt")
    J = ca.Function("J", [x, u, dt], [ca.jacobian(f(x, u, dt), x)])
    assert np.allclose(J(np.zeros(13), np.zeros(7), 0.01).full(), 2.0)


@pytest.mark.gpu
def test_marshal_numbers_on_the_gpu(built):
    from aircraft_b200 import Aircraft
    ac = Aircraft.from_problem(com=TRIM_COM, device="cuda:0", physical_integration_substeps=1)
    orc = oracle_for(TRIM_COM)
    m = cad.DynamicsMarshal(ac, substeps=1, normalise=True)
    x, u = make_states(64, seed=51, perturb_norm=1e-3), make_controls(64, seed=51)
    lam = np.random.default_rng(51).normal(size=(64, 13))
    xo, Ao, Bo, Do = orc.linearize(t64(x), t64(u), 0.01, 1, True, want_dt=True)
    Ho, go, _ = orc.lagrangian_hessian(t64(x), t64(u), t64(lam), 0.01, 1, True)
    xn = m.eval_state_update(x.T, u.T, 0.01)
    assert col_err(xn.T, xo.numpy()) < PARITY_TOL
    As, Bs, Ds, Hs, gs = [], [], [], [], []
    for i in range(64):
        A, B, D = m.jac_state_update(DM(x[i]), DM(u[i]), DM(0.01))
        H, g = m.hess_state_update(DM(x[i]), DM(u[i]), 0.01, DM(lam[i]))
        As.append(A); Bs.append(B); Ds.append(D[:, 0]); Hs.append(H); gs.append(g)
    assert col_err(np.stack(As), Ao.numpy()) < PARITY_TOL and col_err(np.stack(Bs), Bo.numpy()) < PARITY_TOL
    assert col_err(np.stack(Ds), Do.numpy()) < PARITY_TOL
    assert col_err(np.stack(Hs), Ho.numpy(), atol=1e-3) < PARITY_TOL and col_err(np.stack(gs), go.numpy()) < PARITY_TOL

"""The registry's other coefficient models (poly / linear / default; coefficient_models.py:32-135) in the
product's FP32 arithmetic (airc_alt.cuh): CPU-side through the host emulation, against the float64 oracle
and -- for 'poly' -- directly against the trajectory the reference stored (tests/golden/simulation.h5)."""
import numpy as np
import pytest

from _util import PARITY_TOL, col_err, oracle_for, params_for, t64
from aircraft_b200._cabi import MODEL_DEFAULT, MODEL_LINEAR, MODEL_POLY
from aircraft_b200.params import load_linear_model, load_poly_model
from aircraft_b200.trajectory_io import read_trajectories
from aircraft_b200.workloads import make_controls, make_states
from oracle.cpu_ref import HostEmul

TRIM = (0.0131991, -1.78875e-08, 0.00313384)


def _tables(golden_dir):
    return {"poly": (MODEL_POLY, load_poly_model(golden_dir / "poly_model.npz")),
            "linear": (MODEL_LINEAR, load_linear_model(golden_dir / "linear_model.npz")),
            "default": (MODEL_DEFAULT, None)}


def test_poly_table_from_the_reference_pickle_layout(golden_dir):
    t = load_poly_model(golden_dir / "poly_model.npz", rudder_moment_arm=0.5)
    assert t.shape == (211,) and t.dtype == np.float32 and t[-1] == 0.5
    d = np.load(golden_dir / "poly_model.npz")
    assert np.allclose(t[:-1].reshape(6, 35)[:, :34], d["coef"], rtol=1e-7)
    assert np.allclose(t[:-1].reshape(6, 35)[:, 34], d["intercept"], rtol=1e-7, atol=1e-12)


@pytest.mark.parametrize("model", ["poly", "linear", "default"])
def test_models_vs_oracle(built, golden_dir, model):
    p, blob = params_for(TRIM)
    emu, orc = HostEmul(p, blob), oracle_for(TRIM, coeff_model=model)
    mid, tab = _tables(golden_dir)[model]
    x, u = make_states(512, seed=71), make_controls(512, seed=71)
    x[:, 10:13] *= 5.0                                           # rates large enough to exercise the damping terms
    want = orc.state_derivative(t64(x), t64(u)).numpy()
    assert col_err(emu.model_state_derivative(mid, tab, x, u), want) < PARITY_TOL
    for ns, nm in [(1, False), (10, False), (1, True)]:
        want = orc.state_update(t64(x), t64(u), 0.01, ns, nm).numpy()
        assert col_err(emu.model_state_update(mid, tab, x, u, 0.01, ns, nm), want) < PARITY_TOL


def test_poly_reproduces_reference_stored_trajectory(built, golden_dir):
    """The product's FP32 formulas, one step at a time from the reference's own stored states
    (data/trajectories/simulation.h5: poly model, 10 substeps, dt 0.1), against the reference's next state."""
    p, blob = params_for(TRIM)
    emu = HostEmul(p, blob)
    d = read_trajectories(golden_dir / "simulation.h5")["iteration_0"]
    X, U = d["state"].T, d["control"].T
    mid, tab = _tables(golden_dir)["poly"]
    got = emu.model_state_update(mid, tab, X[:-1], U[:-1], 0.1, 10, False)
    assert col_err(got, X[1:]) < PARITY_TOL


@pytest.mark.parametrize("model", ["poly", "linear", "default"])
def test_model_derivatives_vs_oracle(built, golden_dir, model):
    """A, B, dPhi/ddt by two-column forward mode on jets and the Lagrangian Hessian by second-order jets for
    the table models (what IPOPT derives for the reference's 'poly' MPC), against jacfwd of the float64 oracle."""
    p, blob = params_for(TRIM)
    emu, orc = HostEmul(p, blob), oracle_for(TRIM, coeff_model=model)
    mid, tab = _tables(golden_dir)[model]
    n = 96
    x, u = make_states(n, seed=74, perturb_norm=1e-3), make_controls(n, seed=74)
    x[:, 10:13] *= 3.0
    lam = np.random.default_rng(74).normal(size=(n, 13)).astype(np.float32)
    for ns, nm in [(1, False), (2, True)]:
        xn, A, B, Dt = emu.model_linearize(mid, tab, x, u, 0.01, ns, nm)
        xo, Ao, Bo, Do = orc.linearize(t64(x), t64(u), 0.01, ns, nm, want_dt=True)
        assert col_err(xn, xo.numpy()) < PARITY_TOL
        assert col_err(A, Ao.numpy(), atol=1e-3) < PARITY_TOL
        assert col_err(B, Bo.numpy(), atol=1e-3) < PARITY_TOL
        assert col_err(Dt, Do.numpy(), atol=1e-3) < PARITY_TOL
    H, g, v = emu.model_hessian(mid, tab, x, u, lam, 0.01, 1, True)
    Ho, go, vo = (a.numpy() for a in orc.lagrangian_hessian(t64(x), t64(u), t64(lam), 0.01, 1, True))
    # second-order entries: 1e-4 of each entry's OWN scale over the batch.  'poly' stays below 1e-5; for
    # 'linear' / 'default' a few entries that are 7+ orders of magnitude below the matrix scale (e.g.
    # d2/d(q_z)d(aileron) ~ 3e-2 against d2/ddt2 ~ 1e6) carry FP32 cancellation noise of 1e-5 .. 6e-5
    assert col_err(H, Ho, atol=1e-3) < 1e-4 and col_err(g, go) < PARITY_TOL
    assert np.abs(v - vo).max() < PARITY_TOL * np.abs(vo).max()


REF_PICKLE = "/root/reference/data/networks/fitted_models_casadi.pkl"


@pytest.mark.skipif(not __import__("os").path.exists(REF_PICKLE), reason="the reference tree is only present in the build container")
def test_poly_table_read_from_the_reference_pickle_without_casadi(golden_dir):
    """The facade's 'poly' loader on the reference's own pickle (CasADi and sklearn classes stubbed out while
    unpickling) gives the same table as the committed fixture."""
    a = load_poly_model(REF_PICKLE, rudder_moment_arm=0.5)
    b = load_poly_model(golden_dir / "poly_model.npz", rudder_moment_arm=0.5)
    assert np.array_equal(a, b)
    lin = load_linear_model("/root/reference/data/networks/linearised.csv")
    assert np.array_equal(lin, load_linear_model(golden_dir / "linear_model.npz"))

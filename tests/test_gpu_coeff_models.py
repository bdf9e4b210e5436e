"""The registry's other coefficient models on the GPU through the C ABI (airc_create_model): parity with
the float64 oracle, and -- for 'poly' -- the reference's own stored trajectory reproduced by the kernels."""
import numpy as np
import pytest
import torch

from _util import PARITY_TOL, col_err, oracle_for, t64
from aircraft_b200.trajectory_io import read_trajectories
from aircraft_b200.workloads import make_control_sequences, make_controls, make_states

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
TRIM = (0.0131991, -1.78875e-08, 0.00313384)


def f32(a):
    return torch.tensor(np.asarray(a), dtype=torch.float32, device=DEV)


def _aircraft(model, golden_dir, **kw):
    from aircraft_b200 import Aircraft
    path = {"poly": golden_dir / "poly_model.npz", "linear": golden_dir / "linearised.csv", "default": ""}[model]
    return Aircraft.from_problem(com=TRIM, device=DEV, coeff_model_type=model, coeff_model_path=path, **kw)


@pytest.mark.parametrize("model", ["poly", "linear", "default"])
def test_models_parity(built, golden_dir, model):
    ac = _aircraft(model, golden_dir, physical_integration_substeps=1)
    orc = oracle_for(TRIM, coeff_model=model)
    n = 4096
    x, u = make_states(n, seed=72), make_controls(n, seed=72)
    x[:, 10:13] *= 5.0
    got = ac.state_derivative(f32(x.T), f32(u.T)).T.cpu()
    assert col_err(got, orc.state_derivative(t64(x), t64(u)).numpy()) < PARITY_TOL
    for ns, nm in [(1, False), (10, True)]:
        got = ac.state_update(f32(x.T), f32(u.T), 0.01, substeps=ns, normalise=nm).T.cpu()
        assert col_err(got, orc.state_update(t64(x), t64(u), 0.01, ns, nm).numpy()) < PARITY_TOL
    # rollout: teacher-forced parity at every step of the GPU's own trajectory
    B, H = 64, 30
    x0, U = make_states(B, seed=73), make_control_sequences(B, H, seed=73)
    X = ac.rollout(f32(x0), f32(U), 0.01, normalise=True, return_trajectory=True).trajectory.cpu().numpy()
    xs = X[:, :-1].reshape(-1, 13).astype(np.float64)
    want = orc.state_update(t64(xs), t64(U.reshape(-1, 7)), 0.01, 1, True).numpy()
    got = X[:, 1:].reshape(-1, 13)
    # the undamped linear fit and the cubic fit leave their range under random control walks and blow up
    # (in the reference too): compare the nodes that are still inside the flight envelope
    ok = np.isfinite(want).all(1) & np.isfinite(got).all(1) & (np.abs(xs[:, 10:13]).max(1) < 20.0)
    assert ok.sum() > 300, ok.sum()
    assert col_err(got[ok], want[ok]) < PARITY_TOL
    # diagnostics carry the model's coefficients
    d = ac.diagnostics(f32(x.T), f32(u.T))
    assert col_err(d["coefficients"].T.cpu(), orc.coefficients(t64(x), t64(u)).numpy()) < PARITY_TOL
    # derivative entry points are NN-only and say so
    with pytest.raises(RuntimeError, match="nn"):
        ac.linearize(f32(x), f32(u), 0.01)
    with pytest.raises(RuntimeError, match="nn"):
        ac.lagrangian_hessian(f32(x), f32(u), f32(x), 0.01)


def test_poly_kernels_reproduce_reference_stored_trajectory(built, golden_dir):
    """GPU kernels vs an output of the reference itself: data/trajectories/simulation.h5 (written by
    main/dynamics/dynamics.py with the poly model, default 10 substeps, dt 0.1, elevator 3 deg)."""
    ac = _aircraft("poly", golden_dir)                     # physical_integration_substeps default = 10, as the script
    assert ac.physical_integration_substeps == 10
    d = read_trajectories(golden_dir / "simulation.h5")["iteration_0"]
    X, U = d["state"].T, d["control"].T                    # [40,13], [40,7]
    # one step from each stored state (host arrays, as the script calls dyn(state, control, dt))
    nxt = np.stack([ac.state_update(X[k], U[k], 0.1) for k in range(39)])
    assert col_err(nxt, X[1:]) < PARITY_TOL
    # the whole simulation as one rollout from the stored initial state: FP32 drift over 39 x 10 RK4 steps
    traj = ac.rollout(X[:1], U[None, :39], 0.1, return_trajectory=True).trajectory[0]
    err = np.abs(np.asarray(traj) - X).max(axis=1) / np.abs(X).max()
    assert err[1] < 1e-6 and err.max() < 2e-4, err.max()
    # diagnostics on the stored states: unit quaternion -> Euler angles finite, airspeed ~ |v|
    dg = ac.diagnostics(X.T.copy(), U.T.copy())
    assert np.allclose(np.asarray(dg["airspeed"]), np.linalg.norm(X[:, 3:6], axis=1), rtol=1e-3)

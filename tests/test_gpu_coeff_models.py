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
    path = {"poly": golden_dir / "poly_model.npz", "linear": golden_dir / "linear_model.npz", "default": ""}[model]
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
    # derivatives: two-column forward mode / second-order jets (airc_hess.cuh) for the table models
    m = 512
    xs_, us_ = make_states(m, seed=75, perturb_norm=1e-3), make_controls(m, seed=75)
    xs_[:, 10:13] *= 3.0
    lam = np.random.default_rng(75).normal(size=(m, 13)).astype(np.float32)
    xn, A, B, Dt = ac.linearize(f32(xs_), f32(us_), 0.01, normalise=True, want_dt=True)
    xo, Ao, Bo, Do = orc.linearize(t64(xs_), t64(us_), 0.01, 1, True, want_dt=True)
    assert col_err(xn.cpu(), xo.numpy()) < PARITY_TOL
    # generic forward mode on jets (not K2's hand-structured tangents): a few near-cancelling entries of
    # the 'default' model reach 1.2e-5 of their scale in FP32; bar for the table models' Jacobians 3e-5
    assert col_err(A.cpu(), Ao.numpy()) < 3e-5 and col_err(B.cpu(), Bo.numpy()) < 3e-5
    assert col_err(Dt.cpu(), Do.numpy()) < 3e-5
    As = ac.linearize(f32(xs_.T), f32(us_.T), 0.01, normalise=True, layout="soa")[1]
    assert torch.equal(As.permute(2, 0, 1), A)
    H, g, _ = ac.lagrangian_hessian(f32(xs_), f32(us_), f32(lam), 0.01, normalise=True)
    Ho, go, _ = (a.numpy() for a in orc.lagrangian_hessian(t64(xs_), t64(us_), t64(lam), 0.01, 1, True))
    assert col_err(H.cpu(), Ho) < 1e-4 and col_err(g.cpu(), go) < PARITY_TOL     # see tests/test_coeff_models.py
    U3 = f32(make_control_sequences(32, 6, seed=76))
    X3 = ac.rollout(f32(make_states(32, seed=76)), U3, 0.01, normalise=True, return_trajectory=True).trajectory
    r, Ad, Bd = ac.defects(X3, U3, 0.01, normalise=True)
    fin = torch.isfinite(X3).all(-1).all(-1) & (X3[:, :, 10:13].abs().amax((1, 2)) < 20.0)   # still in the envelope
    assert fin.sum() >= 8
    assert float(r[fin].abs().max()) < 2e-5 * float(X3[fin].abs().max())
    # the sampling sweep (one-trajectory kernels for these models) against explicit rollouts of the same controls
    from aircraft_b200.params import make_cost_params, make_sweep_params
    Bs, Hs = 700, 12
    x0s = make_states(1, seed=77, level=True)[0]
    Un = np.zeros((Hs, 7), np.float32); Un[:, 1] = -0.5
    sp, cp = make_sweep_params(seed=9), make_cost_params()
    wsw = ac.sweep(f32(x0s), f32(Un), 0.01, Bs, sampler=sp, cost=cp, normalise=True, return_costs=True)
    Us = ac.sweep_controls(f32(Un), sp, 0, Bs)
    rr = ac.rollout(f32(np.repeat(x0s[None], Bs, 0)), Us, 0.01, normalise=True, cost=cp, return_trajectory=True)
    okc = torch.isfinite(rr.cost)
    assert okc.sum() > Bs // 2
    assert col_err(wsw.cost[okc].cpu()[:, None], rr.cost[okc].cpu().numpy()[:, None]) < PARITY_TOL
    j = int(wsw.argmin)
    assert j == int(torch.argmin(torch.nan_to_num(rr.cost, nan=float("inf"))))
    assert col_err(wsw.trajectory.cpu(), rr.trajectory[j].cpu().numpy()) < 5e-5
    assert torch.equal(wsw.controls, Us[j])


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

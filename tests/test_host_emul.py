"""CPU-side check of the PRODUCT's FP32 formulas (folded network, rotation-matrix form,
forward-mode tangents, RK4 with tangents) compiled with g++ from the very headers the kernels
use (oracle/host_emul.cpp), against the float64 oracle at the 1e-5 parity bar.  The MUFU
approximations are libm here; the -m gpu tests check the real device arithmetic."""
import numpy as np
import pytest

from _util import PARITY_TOL, col_err, oracle_for, params_for, t64
from aircraft_b200.workloads import make_controls, make_states
from oracle.cpu_ref import HostEmul


@pytest.fixture(scope="module")
def pair(built, com):
    p, blob = params_for(com)
    return HostEmul(p, blob), oracle_for(com)


def test_folded_surrogate_vs_reference_module(built, golden_dir):
    """W21 = W2.W1 with scalers folded in float64, logistic form of tanh: equal to the reference's
    ScaledModel on the golden inputs."""
    g = np.load(golden_dir / "nn_golden.npz")
    p, blob = params_for()
    got = HostEmul(p, blob).surrogate(g["inputs"])
    assert col_err(got, g["out_f64"], atol=1e-4) < PARITY_TOL


def test_state_derivative(pair):
    emu, orc = pair
    x, u = make_states(4096, seed=21, perturb_norm=1e-3), make_controls(4096, seed=21)
    want = orc.state_derivative(t64(x), t64(u)).numpy()
    assert col_err(emu.state_derivative(x, u), want) < PARITY_TOL


@pytest.mark.parametrize("substeps,normalise", [(1, False), (1, True), (10, True)])
def test_state_update(pair, substeps, normalise):
    emu, orc = pair
    x, u = make_states(2048, seed=22), make_controls(2048, seed=22)
    want = orc.state_update(t64(x), t64(u), 0.01, substeps, normalise).numpy()
    assert col_err(emu.state_update(x, u, 0.01, substeps, normalise), want) < PARITY_TOL


@pytest.mark.parametrize("substeps,normalise", [(1, False), (2, True)])
def test_linearize(pair, substeps, normalise):
    emu, orc = pair
    x, u = make_states(512, seed=23, perturb_norm=1e-3), make_controls(512, seed=23)
    xn, A, B = emu.linearize(x, u, 0.01, substeps, normalise)
    xo, Ao, Bo = orc.linearize(t64(x), t64(u), 0.01, substeps, normalise)
    assert col_err(xn, xo.numpy()) < PARITY_TOL
    # Jacobian entries: per-entry scale over the batch, floor 1e-3 of the matrix scale
    assert col_err(A, Ao.numpy(), atol=1e-3) < PARITY_TOL
    assert col_err(B, Bo.numpy(), atol=1e-3) < PARITY_TOL


@pytest.mark.parametrize("substeps,normalise", [(1, False), (2, True)])
def test_linearize_packed_tangents(pair, substeps, normalise):
    """The cooperative K2 arithmetic: two Jacobian columns per packed tangent (f_tangent_t<F2>,
    rk4_step_lin) with the network policy NetCoop<1>."""
    emu, orc = pair
    x, u = make_states(512, seed=24, perturb_norm=1e-3), make_controls(512, seed=24)
    xn, A, B = emu.linearize_packed(x, u, 0.01, substeps, normalise)
    xo, Ao, Bo = orc.linearize(t64(x), t64(u), 0.01, substeps, normalise)
    assert col_err(xn, xo.numpy()) < PARITY_TOL
    assert col_err(A, Ao.numpy(), atol=1e-3) < PARITY_TOL
    assert col_err(B, Bo.numpy(), atol=1e-3) < PARITY_TOL


def test_linearize_packed_with_stall_scaling(built):
    p, blob = params_for(stall_scaling=True)
    emu, orc = HostEmul(p, blob), oracle_for(stall_scaling=True)
    x, u = make_states(256, seed=25), make_controls(256, seed=25)
    x[:128, 3:6] *= np.array([0.6, 1.0, 2.0])
    xn, A, B = emu.linearize_packed(x, u, 0.01)
    xo, Ao, Bo = orc.linearize(t64(x), t64(u), 0.01)
    assert col_err(A, Ao.numpy(), atol=1e-3) < PARITY_TOL
    assert col_err(B, Bo.numpy(), atol=1e-3) < PARITY_TOL


@pytest.mark.parametrize("substeps,normalise", [(1, False), (3, True)])
def test_linearize_packed_dt_column(pair, substeps, normalise):
    """dPhi/d(dt) carried in tangent slot 14 of the cooperative kernel's arithmetic, against
    jacfwd w.r.t. dt of the float64 oracle; A and B are unchanged by the extra slot."""
    emu, orc = pair
    x, u = make_states(512, seed=26, perturb_norm=1e-3), make_controls(512, seed=26)
    xn, A, B, Dt = emu.linearize_packed(x, u, 0.02, substeps, normalise, want_dt=True)
    xo, Ao, Bo, Do = orc.linearize(t64(x), t64(u), 0.02, substeps, normalise, want_dt=True)
    assert col_err(Dt, Do.numpy(), atol=1e-3) < PARITY_TOL
    x2, A2, B2 = emu.linearize_packed(x, u, 0.02, substeps, normalise)
    assert np.array_equal(A, A2) and np.array_equal(B, B2) and np.array_equal(xn, x2)
    # d/d(dt) of an RK4 step at dt -> 0 is f(x,u): first-order consistency of the column
    _, _, _, D0 = orc.linearize(t64(x), t64(u), 1e-9, 1, False, want_dt=True)
    assert np.allclose(D0.numpy(), orc.state_derivative(t64(x), t64(u)).numpy(), rtol=1e-6, atol=1e-6)


@pytest.mark.parametrize("substeps,normalise", [(1, False), (1, True), (2, True), (3, False)])
def test_lagrangian_hessian_jets(pair, substeps, normalise):
    """K4 arithmetic against jacfwd(jacfwd(lam^T Phi)) of the float64 oracle over z = (x, u, dt), both forms:
    `hessian2` = forward over reverse (airc_hess2.cuh: the hand-written adjoint sweep of the RK4 step on dual
    numbers, eight lanes per node; what airc_hessian runs for the network model), `hessian` = second-order
    forward mode (airc_jet.cuh / airc_hess.cuh: 64 (outer, inner-pair) tasks per node; the table models),
    `hessian_shared` = the same jets with the network's value-level sums (C, dC/dy, d2C/dy2) shared by the 64
    threads of a node (NetSharedJet; what short horizons of the network model run)."""
    emu, orc = pair
    n = 40
    x, u = make_states(n, seed=27, perturb_norm=1e-3), make_controls(n, seed=27)
    lam = np.random.default_rng(27).normal(size=(n, 13)).astype(np.float32)
    Ho, go, vo = (a.numpy() for a in orc.lagrangian_hessian(t64(x), t64(u), t64(lam), 0.02, substeps, normalise))
    inert = [0, 1, 2, 16, 17, 18]                        # p never enters f; thrust is inert (aircraft.py:320)
    assert not Ho[:, inert, :].any()
    for arith in ("hessian2", "hessian_shared", "hessian"):
        H, g, v = getattr(emu, arith)(x, u, lam, 0.02, substeps, normalise)
        assert col_err(H, Ho, atol=1e-3) < PARITY_TOL, arith   # every entry against its own scale over the batch
        assert col_err(g, go) < PARITY_TOL, arith
        assert np.abs(v - vo).max() < PARITY_TOL * np.abs(vo).max(), arith
        assert np.array_equal(H, H.transpose(0, 2, 1)), arith
        assert not H[:, inert, :].any(), arith
    # the gradient is [A B dPhi/ddt]^T lam
    xo, A, B, Dt = orc.linearize(t64(x), t64(u), 0.02, substeps, normalise, want_dt=True)
    J = np.concatenate([A.numpy(), B.numpy(), Dt.numpy()[:, :, None]], axis=2)
    assert np.allclose(go, np.einsum("nr,nrz->nz", lam.astype(np.float64), J), rtol=1e-9, atol=1e-9)


def test_lagrangian_hessian_jets_stall_scaling(built):
    p, blob = params_for(stall_scaling=True)
    emu, orc = HostEmul(p, blob), oracle_for(stall_scaling=True)
    n = 64
    x, u = make_states(n, seed=28), make_controls(n, seed=28)
    x[:32, 3:6] *= np.array([0.6, 1.0, 2.0])             # large alpha: the scaling is active
    lam = np.random.default_rng(28).normal(size=(n, 13)).astype(np.float32)
    Ho, go, vo = (a.numpy() for a in orc.lagrangian_hessian(t64(x), t64(u), t64(lam), 0.01))
    for arith in (emu.hessian2, emu.hessian_shared, emu.hessian):
        H, g, v = arith(x, u, lam, 0.01)
        assert col_err(H, Ho, atol=1e-3) < PARITY_TOL and col_err(g, go) < PARITY_TOL
        assert np.abs(v - vo).max() < PARITY_TOL * np.abs(vo).max()


@pytest.mark.parametrize("normalise", [False, True])
def test_linearize_tile_structured_jacobian(pair, normalise):
    """The tile kernel's arithmetic (airc_jac.cuh): f_x / f_u blocks assembled once per stage point from the
    chain (v,q) -> v_b -> (qbar, alpha, beta) -> C -> (F, M), parked, then the RK4 tangent recursion as small
    matrix products -- against jacfwd of the float64 oracle, incl. the d/d(dt) column."""
    emu, orc = pair
    x, u = make_states(512, seed=29, perturb_norm=1e-3), make_controls(512, seed=29)
    xn, A, B, Dt = emu.linearize_tile(x, u, 0.02, normalise, want_dt=True)
    xo, Ao, Bo, Do = orc.linearize(t64(x), t64(u), 0.02, 1, normalise, want_dt=True)
    assert col_err(xn, xo.numpy()) < PARITY_TOL
    assert col_err(A, Ao.numpy(), atol=1e-3) < PARITY_TOL
    assert col_err(B, Bo.numpy(), atol=1e-3) < PARITY_TOL
    assert col_err(Dt, Do.numpy(), atol=1e-3) < PARITY_TOL
    x2, A2, B2 = emu.linearize_tile(x, u, 0.02, normalise)
    assert np.array_equal(A, A2) and np.array_equal(B, B2) and np.array_equal(xn, x2)
    eye = np.eye(13, dtype=np.float32)[:, :3]
    assert np.array_equal(A[:, :, :3], np.broadcast_to(eye, (512, 13, 3))) and not B[:, :, 3:6].any()


def test_linearize_tile_with_stall_scaling(built):
    p, blob = params_for(stall_scaling=True)
    emu, orc = HostEmul(p, blob), oracle_for(stall_scaling=True)
    x, u = make_states(256, seed=30), make_controls(256, seed=30)
    x[:128, 3:6] *= np.array([0.6, 1.0, 2.0])
    xn, A, B = emu.linearize_tile(x, u, 0.01)
    xo, Ao, Bo = orc.linearize(t64(x), t64(u), 0.01)
    assert col_err(A, Ao.numpy(), atol=1e-3) < PARITY_TOL
    assert col_err(B, Bo.numpy(), atol=1e-3) < PARITY_TOL


def test_nan_reaches_f_and_phi(built):
    """The NaN-preserving clamp of the logistic (min_nan): a NaN aileron / elevator / state must not be turned into
    a saturated hidden layer and finite coefficients."""
    p, blob = params_for()
    emu = HostEmul(p, blob)
    x, u = make_states(20, seed=92), make_controls(20, seed=92)
    for j in range(3, 13):
        x[j, j] = np.nan
    for k, j in enumerate((0, 1, 2, 6)):
        u[13 + k, j] = np.nan
    f = emu.state_derivative(x, u)
    phi = emu.state_update(x, u, 0.01, 1, False)
    for j in list(range(3, 13)) + [13, 14, 15, 16]:
        assert np.isnan(f[j]).any() and np.isnan(phi[j]).any(), j
    assert np.isfinite(f[:3]).all() and np.isfinite(f[17:]).all()

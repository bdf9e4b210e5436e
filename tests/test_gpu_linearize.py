"""Jacobian parity (K2) and full-size property tests on the GPU, through the C ABI."""
import numpy as np
import pytest
import torch

from _util import PARITY_TOL, col_err, oracle_for, t64
from aircraft_b200.config import TRIM_COM
from aircraft_b200.workloads import make_control_sequences, make_controls, make_states

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def f32(a):
    return torch.tensor(np.asarray(a), dtype=torch.float32, device=DEV)


@pytest.fixture(scope="module")
def ac(built):
    from aircraft_b200 import Aircraft
    return Aircraft.from_problem(com=TRIM_COM, device=DEV, physical_integration_substeps=1)


@pytest.fixture(scope="module")
def orc():
    return oracle_for(TRIM_COM)


@pytest.mark.parametrize("substeps,normalise,dt", [(1, False, 0.01), (1, True, 0.01), (3, True, 0.03)])
def test_linearize_parity(ac, orc, substeps, normalise, dt):
    n = 4096
    x, u = make_states(n, seed=41, perturb_norm=1e-3), make_controls(n, seed=41)
    xn, A, B = ac.linearize(f32(x), f32(u), dt, substeps=substeps, normalise=normalise)
    xo, Ao, Bo = orc.linearize(t64(x), t64(u), dt, substeps, normalise)
    assert col_err(xn.cpu(), xo.numpy()) < PARITY_TOL
    assert col_err(A.cpu(), Ao.numpy()) < PARITY_TOL
    assert col_err(B.cpu(), Bo.numpy()) < PARITY_TOL


def test_linearize_structure_and_layouts(ac):
    n = 1001
    x, u = make_states(n, seed=42, perturb_norm=1e-3), make_controls(n, seed=42)
    xn, A, B = ac.linearize(f32(x), f32(u), 0.01)
    eye = torch.eye(13, device=DEV)[:, :3]
    assert torch.equal(A[:, :, :3], eye.expand(n, 13, 3))      # dPhi/dp = [I;0] exactly
    assert torch.count_nonzero(B[:, :, 3:6]) == 0              # thrust is inert (aircraft.py:320)
    # same primal formulas; the two kernels are compiled separately, so FMA contraction may differ in the last ulp
    assert col_err(xn.cpu(), ac.state_update(f32(x.T), f32(u.T), 0.01).T.cpu()) < 1e-6
    xs, As, Bs = ac.linearize(f32(x.T), f32(u.T), 0.01, layout="soa")
    assert torch.equal(As.permute(2, 0, 1), A) and torch.equal(Bs.permute(2, 0, 1), B)
    assert torch.equal(xs.T, xn)
    # leading batch dims are kept: [windows, nodes, 13]
    xw, Aw, Bw = ac.linearize(f32(x[:1000].reshape(20, 50, 13)), f32(u[:1000].reshape(20, 50, 7)), 0.01)
    assert Aw.shape == (20, 50, 13, 13) and Bw.shape == (20, 50, 13, 7)
    assert torch.equal(Aw.reshape(1000, 13, 13), A[:1000])


def test_linearize_per_node_dt_and_stall_scaling(built):
    from aircraft_b200 import Aircraft
    ac = Aircraft.from_problem(com=TRIM_COM, device=DEV, physical_integration_substeps=1, stall_scaling=True)
    orc = oracle_for(TRIM_COM, stall_scaling=True)
    n = 2048
    x, u = make_states(n, seed=43), make_controls(n, seed=43)
    x[: n // 2, 3:6] *= np.array([0.6, 1.0, 2.0])         # push half of the nodes to high alpha
    dt = np.random.default_rng(2).uniform(0.005, 0.02, n)
    xn, A, B = ac.linearize(f32(x), f32(u), f32(dt))
    xo, Ao, Bo = orc.linearize(t64(x), t64(u), t64(dt))
    assert col_err(xn.cpu(), xo.numpy()) < PARITY_TOL
    assert col_err(A.cpu(), Ao.numpy()) < PARITY_TOL
    assert col_err(B.cpu(), Bo.numpy()) < PARITY_TOL


def test_chain_rule_property(ac):
    """Size-independent property: the Jacobian of two substeps of dt/2 equals the product of the
    Jacobians of the two successive single steps (65536 nodes)."""
    n = 65536
    x, u = make_states(n, seed=44), make_controls(n, seed=44)
    X, U = f32(x), f32(u)
    x1, A1, B1 = ac.linearize(X, U, 0.01)
    x2, A2, B2 = ac.linearize(x1, U, 0.01)
    xx, AA, BB = ac.linearize(X, U, 0.02, substeps=2)
    assert col_err(xx.cpu(), x2.cpu()) < 1e-6
    Aprod = torch.bmm(A2, A1)
    Bprod = torch.bmm(A2, B1) + B2
    assert col_err(AA.cpu(), Aprod.cpu()) < PARITY_TOL
    assert col_err(BB.cpu(), Bprod.cpu()) < PARITY_TOL


def test_full_size_c2_determinism_and_sharding_invariance(ac):
    """BASELINE config 2 (4096 x 200): bit-identical between runs; a whole-batch launch against two half-batch
    launches (what a multi-GPU sharding would do) is bit-identical while both sides are served by the same kernel
    instantiation (the latency kernel picks 8 / 4 / 2 lanes per trajectory by batch size: 2048 vs 2 x 1024 both run 8
    lanes) and agrees to FP32 rounding growth over the 200 free-running steps otherwise (4096 runs 4 lanes, its
    halves 8); unit quaternions with normalise."""
    B, H = 4096, 200
    x0, U = make_states(B, seed=45), make_control_sequences(B, H, seed=45, dtype=np.float32)
    X0, UU = f32(x0.T), f32(U.transpose(1, 2, 0))
    a = ac.rollout(X0, UU, 0.01, layout="soa", normalise=True, cost=True)
    b = ac.rollout(X0, UU, 0.01, layout="soa", normalise=True, cost=True)
    assert torch.equal(a.final, b.final) and torch.equal(a.cost, b.cost)
    h = B // 2
    lo = ac.rollout(X0[:, :h].contiguous(), UU[:, :, :h].contiguous(), 0.01, layout="soa", normalise=True, cost=True)
    hi = ac.rollout(X0[:, h:].contiguous(), UU[:, :, h:].contiguous(), 0.01, layout="soa", normalise=True, cost=True)
    halves = torch.cat([lo.final, hi.final], 1)
    ok = torch.isfinite(a.final).all(0) & torch.isfinite(halves).all(0)
    assert col_err(halves[:, ok].T.cpu(), a.final[:, ok].T.cpu().numpy()) < 2e-3     # 200 free-running FP32 steps
    q = h // 2                                          # 2048 against 2 x 1024: the same instantiation on both sides
    l2 = ac.rollout(X0[:, :q].contiguous(), UU[:, :, :q].contiguous(), 0.01, layout="soa", normalise=True, cost=True)
    h2 = ac.rollout(X0[:, q:h].contiguous(), UU[:, :, q:h].contiguous(), 0.01, layout="soa", normalise=True, cost=True)
    assert torch.equal(torch.cat([l2.final, h2.final], 1), lo.final)
    assert torch.equal(torch.cat([l2.cost, h2.cost]), lo.cost)
    fin = torch.isfinite(a.final).all(0)
    qn = (a.final[6:10, fin] ** 2).sum(0)
    assert (qn - 1).abs().max() < 1e-6
    m, i = ac.argmin(a.cost)
    ref = torch.where(torch.isnan(a.cost), torch.full_like(a.cost, float("inf")), a.cost)
    assert int(i) == int(ref.argmin())


def test_sharded_sweep_single_process(ac):
    """ShardedRollout without a process group == plain rollout + argmin, and hands back the
    winner's trajectory."""
    from aircraft_b200.sharding import ShardedRollout
    B, H = 2000, 25
    x0, U = make_states(B, seed=46), make_control_sequences(B, H, seed=46)
    sr = ShardedRollout(ac)
    res = sr.sweep(f32(x0.T), f32(U.transpose(1, 2, 0)), 0.01, B, layout="soa", gather_costs=True)
    full = ac.rollout(f32(x0), f32(U), 0.01, return_trajectory=True, cost=True)
    assert torch.equal(res.all_cost, full.cost)
    j = int(torch.where(torch.isnan(full.cost), torch.full_like(full.cost, float("inf")), full.cost).argmin())
    assert res.argmin == j and res.owner == 0
    # the winner is re-rolled by a single-trajectory launch (another instantiation of the kernel: FMA contraction may
    # differ in the last bit, which 25 free-running steps carry along)
    assert col_err(res.best_trajectory.cpu(), full.trajectory[j].cpu(), batch_axes=(0,)) < 5e-6
    assert torch.equal(res.best_controls, f32(U[j]))


@pytest.mark.parametrize("substeps,normalise,dt", [(1, False, 0.01), (3, True, 0.03)])
def test_linearize_dt_column(ac, orc, substeps, normalise, dt):
    """dPhi/d(dt) from the spare tangent slot of K2 (airc_linearize_dt): parity with jacfwd of the
    float64 oracle, unchanged A/B/x+, both layouts, host path, and a central difference in dt."""
    n = 4099
    x, u = make_states(n, seed=45, perturb_norm=1e-3), make_controls(n, seed=45)
    xn, A, B, Dt = ac.linearize(f32(x), f32(u), dt, substeps=substeps, normalise=normalise, want_dt=True)
    xo, Ao, Bo, Do = orc.linearize(t64(x), t64(u), dt, substeps, normalise, want_dt=True)
    assert Dt.shape == (n, 13)
    assert col_err(Dt.cpu(), Do.numpy()) < PARITY_TOL
    x3, A3, B3 = ac.linearize(f32(x), f32(u), dt, substeps=substeps, normalise=normalise)
    assert torch.equal(A, A3) and torch.equal(B, B3) and torch.equal(xn, x3)
    xs, As, Bs, Ds = ac.linearize(f32(x.T), f32(u.T), dt, substeps=substeps, normalise=normalise,
                                  layout="soa", want_dt=True)
    assert Ds.shape == (13, n) and torch.equal(Ds.T, Dt)
    hx, hA, hB, hD = ac.linearize(x[:257], u[:257], dt, substeps=substeps, normalise=normalise, want_dt=True)
    assert isinstance(hD, np.ndarray) and np.array_equal(hD, Dt[:257].cpu().numpy())
    # per-node dt: the column is evaluated at each node's own step size
    dts = np.linspace(0.5 * dt, 1.5 * dt, n).astype(np.float32)
    _, _, _, Dn = ac.linearize(f32(x), f32(u), f32(dts), substeps=substeps, normalise=normalise, want_dt=True)
    _, _, _, Dno = orc.linearize(t64(x), t64(u), t64(dts), substeps, normalise, want_dt=True)
    assert col_err(Dn.cpu(), Dno.numpy()) < PARITY_TOL


def test_multiple_shooting_defects(ac, orc):
    """r_k = x_{k+1} - Phi(x_k,u_k,dt) with (A,B,dPhi/ddt) blocks for many windows in one launch
    (airc_defects); a rollout's own trajectory has zero defect."""
    W, N = 7, 40
    x0 = make_states(W, seed=61)
    U = make_control_sequences(W, N, seed=61)
    X = ac.rollout(f32(x0), f32(U), 0.01, normalise=True, return_trajectory=True).trajectory   # [W,N+1,13]
    r, A, B = ac.defects(X, f32(U), 0.01, normalise=True)
    assert r.shape == (W, N, 13) and A.shape == (W, N, 13, 13) and B.shape == (W, N, 13, 7)
    assert float(r.abs().max()) < 2e-5 * float(X.abs().max())     # two kernels, last-ulp differences only
    # perturbed guess: parity with the oracle on every block
    g = torch.Generator(device="cpu").manual_seed(5)
    Xp = X.cpu() + 1e-2 * torch.randn(X.shape, generator=g)
    r, A, B, Dt = ac.defects(f32(Xp), f32(U), 0.01, normalise=True, want_dt=True)
    xs, us = Xp[:, :-1].reshape(-1, 13).numpy(), U.reshape(-1, 7)
    xo, Ao, Bo, Do = orc.linearize(t64(xs), t64(us), 0.01, 1, True, want_dt=True)
    want_r = Xp[:, 1:].reshape(-1, 13).double().numpy() - xo.numpy()
    assert col_err(r.reshape(-1, 13).cpu(), want_r, atol=1.0) < PARITY_TOL    # defect scale ~ state scale
    assert col_err(A.reshape(-1, 13, 13).cpu(), Ao.numpy()) < PARITY_TOL
    assert col_err(B.reshape(-1, 13, 7).cpu(), Bo.numpy()) < PARITY_TOL
    assert col_err(Dt.reshape(-1, 13).cpu(), Do.numpy()) < PARITY_TOL
    # the reference's guess layout [13+7, N+1] and host arrays
    guess = np.concatenate([Xp[0].numpy().T, np.concatenate([U[0], U[0][-1:]], 0).T], axis=0)
    rg, Ag, Bg = ac.defects_from_guess(guess, 0.01, normalise=True)
    assert isinstance(rg, np.ndarray) and rg.shape == (N, 13)
    assert np.array_equal(rg.astype(np.float32), r[0].cpu().numpy()) and np.array_equal(Ag.astype(np.float32), A[0].cpu().numpy())

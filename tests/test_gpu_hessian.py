"""Second-order products (K3, airc_hessian) on the GPU through the C ABI: parity with the float64
oracle's jacfwd(jacfwd(lam^T Phi)), structure, layouts, host path, consistency with K2."""
import numpy as np
import pytest
import torch

from _util import PARITY_TOL, col_err, oracle_for, t64
from aircraft_b200.config import TRIM_COM
from aircraft_b200.workloads import make_controls, make_states

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


def _inputs(n, seed):
    x, u = make_states(n, seed=seed, perturb_norm=1e-3), make_controls(n, seed=seed)
    lam = np.random.default_rng(seed).normal(size=(n, 13)).astype(np.float32)
    return x, u, lam


@pytest.mark.parametrize("substeps,normalise,dt", [(1, False, 0.01), (1, True, 0.02), (3, True, 0.03)])
def test_hessian_parity(ac, orc, substeps, normalise, dt):
    n = 515
    x, u, lam = _inputs(n, 51)
    H, g, v = ac.lagrangian_hessian(f32(x), f32(u), f32(lam), dt, substeps=substeps, normalise=normalise,
                                    want_value=True)
    Ho, go, vo = (a.numpy() for a in orc.lagrangian_hessian(t64(x), t64(u), t64(lam), dt, substeps, normalise))
    assert H.shape == (n, 21, 21) and g.shape == (n, 21) and v.shape == (n,)
    assert col_err(H.cpu(), Ho) < PARITY_TOL
    assert col_err(g.cpu(), go) < PARITY_TOL
    assert np.abs(v.cpu().numpy() - vo).max() < PARITY_TOL * np.abs(vo).max()


def test_hessian_structure_layouts_and_host_path(ac):
    n = 300
    x, u, lam = _inputs(n, 52)
    H, g, _ = ac.lagrangian_hessian(f32(x), f32(u), f32(lam), 0.01)
    assert torch.equal(H, H.transpose(1, 2))
    inert = [0, 1, 2, 16, 17, 18]
    assert torch.count_nonzero(H[:, inert, :]) == 0
    assert torch.equal(g[:, :3], f32(lam)[:, :3]) and torch.count_nonzero(g[:, 16:19]) == 0
    Hs, gs, vs = ac.lagrangian_hessian(f32(x.T), f32(u.T), f32(lam.T), 0.01, layout="soa", want_value=True)
    assert Hs.shape == (21, 21, n) and torch.equal(Hs.permute(2, 0, 1), H) and torch.equal(gs.T, g)
    Hh, gh, _ = ac.lagrangian_hessian(x, u, lam, 0.01)
    assert isinstance(Hh, np.ndarray) and np.array_equal(Hh, H.cpu().numpy()) and np.array_equal(gh, g.cpu().numpy())
    # leading batch dims are kept: [windows, nodes, ...]
    Hw, gw, _ = ac.lagrangian_hessian(f32(x.reshape(6, 50, 13)), f32(u.reshape(6, 50, 7)),
                                      f32(lam.reshape(6, 50, 13)), 0.01)
    assert Hw.shape == (6, 50, 21, 21) and torch.equal(Hw.reshape(n, 21, 21), H)
    # linear in the multipliers
    H2, g2, _ = ac.lagrangian_hessian(f32(x), f32(u), f32(2 * lam), 0.01)
    assert col_err(H2.cpu(), 2 * H.cpu().numpy()) < 1e-6
    with pytest.raises(RuntimeError):
        ac.lagrangian_hessian(f32(x), f32(u), f32(lam[:, :12]), 0.01)


def test_hessian_gradient_matches_jacobian_kernel(ac):
    """grad = [A B dPhi/ddt]^T lam: K3's first-order part against K2 (separately written tangents)."""
    n = 2048
    x, u, lam = _inputs(n, 53)
    dts = np.linspace(0.005, 0.02, n).astype(np.float32)
    _, g, v = ac.lagrangian_hessian(f32(x), f32(u), f32(lam), f32(dts), normalise=True, want_value=True)
    xn, A, B, Dt = ac.linearize(f32(x), f32(u), f32(dts), normalise=True, want_dt=True)
    J = torch.cat([A, B, Dt[:, :, None]], dim=2).double()
    want = torch.einsum("nr,nrz->nz", f32(lam).double(), J)
    assert col_err(g.cpu(), want.cpu().numpy()) < PARITY_TOL
    assert col_err(v.cpu()[:, None], (f32(lam).double() * xn.double()).sum(1).cpu().numpy()[:, None]) < PARITY_TOL

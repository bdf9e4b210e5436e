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


@pytest.mark.parametrize("n", [300, 700])             # the shared-network jets kernel / the forward-over-reverse kernel
@pytest.mark.parametrize("substeps,normalise,dt", [(1, False, 0.01), (1, True, 0.02), (3, True, 0.03)])
def test_hessian_parity(ac, orc, substeps, normalise, dt, n):
    x, u, lam = _inputs(n, 51)
    H, g, v = ac.lagrangian_hessian(f32(x), f32(u), f32(lam), dt, substeps=substeps, normalise=normalise,
                                    want_value=True)
    Ho, go, vo = (a.numpy() for a in orc.lagrangian_hessian(t64(x), t64(u), t64(lam), dt, substeps, normalise))
    assert H.shape == (n, 21, 21) and g.shape == (n, 21) and v.shape == (n,)
    assert col_err(H.cpu(), Ho) < PARITY_TOL
    assert col_err(g.cpu(), go) < PARITY_TOL
    assert np.abs(v.cpu().numpy() - vo).max() < PARITY_TOL * np.abs(vo).max()


@pytest.mark.parametrize("n", [300, 1500])            # the shared-network jets kernel / the forward-over-reverse kernel
def test_hessian_structure_layouts_and_host_path(ac, n):
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
    Hw, gw, _ = ac.lagrangian_hessian(f32(x.reshape(6, n // 6, 13)), f32(u.reshape(6, n // 6, 7)),
                                      f32(lam.reshape(6, n // 6, 13)), 0.01)
    assert Hw.shape == (6, n // 6, 21, 21) and torch.equal(Hw.reshape(n, 21, 21), H)
    # linear in the multipliers
    H2, g2, _ = ac.lagrangian_hessian(f32(x), f32(u), f32(2 * lam), 0.01)
    assert col_err(H2.cpu(), 2 * H.cpu().numpy()) < 1e-6
    with pytest.raises(RuntimeError):
        ac.lagrangian_hessian(f32(x), f32(u), f32(lam[:, :12]), 0.01)


def test_hessian_two_formulations_agree(ac, orc):
    """airc_hessian runs the network model's long batches by forward over reverse (airc_hess2.cuh, eight lanes per
    node sharing the network through shuffles) and short horizons (up to four nodes per SM) on second-order
    jets (airc_hess.cuh; one node per 64-thread block, the network's sums shared through shared memory): the same
    nodes through both, against each other and against the oracle, and the last partial warp of the long batch
    (idle lanes keep step for the shuffles).  More than 8 substeps on a long batch: the unshared jets kernel."""
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    n_short, n_long = 4 * sms, 4 * sms + 5           # 5 nodes = 40 lanes: one full warp and a quarter of the next
    x, u, lam = _inputs(n_long, 54)
    Hl, gl, vl = ac.lagrangian_hessian(f32(x), f32(u), f32(lam), 0.02, normalise=True, want_value=True)
    Hs, gs, vs = ac.lagrangian_hessian(f32(x[:n_short]), f32(u[:n_short]), f32(lam[:n_short]), 0.02, normalise=True,
                                       want_value=True)
    assert not torch.equal(Hl[:n_short], Hs)          # two different arithmetics ...
    assert col_err(Hl[:n_short].cpu(), Hs.cpu().numpy()) < 2e-6 and col_err(gl[:n_short].cpu(), gs.cpu().numpy()) < 2e-6
    Ho, go, vo = (a.numpy() for a in orc.lagrangian_hessian(t64(x), t64(u), t64(lam), 0.02, 1, True))
    assert col_err(Hl.cpu(), Ho) < PARITY_TOL and col_err(gl.cpu(), go) < PARITY_TOL
    assert col_err(Hs.cpu(), Ho[:n_short]) < PARITY_TOL
    assert np.abs(vl.cpu().numpy() - vo).max() < PARITY_TOL * np.abs(vo).max()
    assert torch.equal(Hl, Hl.transpose(1, 2))
    # more substeps than the reverse sweep keeps start states for: the jets again, same answer as the oracle
    H9, g9, _ = ac.lagrangian_hessian(f32(x[:600]), f32(u[:600]), f32(lam[:600]), 0.02, substeps=9)
    Ho9, go9, _ = (a.numpy() for a in orc.lagrangian_hessian(t64(x[:600]), t64(u[:600]), t64(lam[:600]), 0.02, 9, False))
    assert col_err(H9.cpu(), Ho9) < PARITY_TOL and col_err(g9.cpu(), go9) < PARITY_TOL
    H8, g8, _ = ac.lagrangian_hessian(f32(x[:600]), f32(u[:600]), f32(lam[:600]), 0.02, substeps=8, normalise=True)
    Ho8, go8, _ = (a.numpy() for a in orc.lagrangian_hessian(t64(x[:600]), t64(u[:600]), t64(lam[:600]), 0.02, 8, True))
    assert col_err(H8.cpu(), Ho8) < PARITY_TOL and col_err(g8.cpu(), go8) < PARITY_TOL


def test_hessian_block_size_switch(ac):
    """Long batches run 256-thread blocks (always from 6.5 waves of 32 nodes per SM on), short ones 128: the same nodes give the same
    bits either way (a node's eight lanes never see the block), including the mostly idle last block."""
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    n_long = 256 * sms + 1                            # 256-thread blocks, one live node in the last block
    x, u, lam = _inputs(n_long, 55)
    Hl, gl, vl = ac.lagrangian_hessian(f32(x), f32(u), f32(lam), 0.01, normalise=True, want_value=True)
    for lo, hi in ((0, 700), (n_long - 700, n_long)):  # the same nodes in a batch of 700: 128-thread blocks
        Hs, gs, vs = ac.lagrangian_hessian(f32(x[lo:hi]), f32(u[lo:hi]), f32(lam[lo:hi]), 0.01, normalise=True,
                                           want_value=True)
        assert torch.equal(Hl[lo:hi], Hs) and torch.equal(gl[lo:hi], gs) and torch.equal(vl[lo:hi], vs)
    assert torch.isfinite(Hl).all() and torch.equal(Hl, Hl.transpose(1, 2))


@pytest.mark.parametrize("n", [200, 1500, 40000])     # shared-network jets / forward over reverse, 128 / 256 threads
@pytest.mark.parametrize("layout", ["aos", "soa"])
def test_hessian_overwrites_every_entry(ac, n, layout):
    """airc_hessian leaves no entry of H, grad, value as it found it (the structural zeros come from a memset on the
    jets paths and from the kernel itself on the forward-over-reverse path): outputs pre-filled with NaN."""
    from aircraft_b200 import _cabi
    x, u, lam = _inputs(n, 56)
    soa = layout == "soa"
    xd, ud, ld = (f32(a.T.copy() if soa else a) for a in (x, u, lam))
    H = torch.full((n * 441,), float("nan"), device=DEV)
    g = torch.full((n * 21,), float("nan"), device=DEV)
    v = torch.full((n,), float("nan"), device=DEV)
    rc = ac._lib.airc_hessian(ac.handle, xd.data_ptr(), ud.data_ptr(), ld.data_ptr(), 0.01, None, n, 1, 1,
                              _cabi.LAYOUT_SOA if soa else _cabi.LAYOUT_AOS, H.data_ptr(), g.data_ptr(), v.data_ptr(),
                              torch.cuda.current_stream().cuda_stream)
    assert rc == 0
    assert torch.isfinite(H).all() and torch.isfinite(g).all() and torch.isfinite(v).all()
    Hn = H.view(21, 21, n).permute(2, 0, 1) if soa else H.view(n, 21, 21)
    inert = [0, 1, 2, 16, 17, 18]
    assert torch.count_nonzero(Hn[:, inert, :]) == 0 and torch.count_nonzero(Hn[:, :, inert]) == 0
    Hf, gf, vf = ac.lagrangian_hessian(f32(x), f32(u), f32(lam), 0.01, normalise=True, want_value=True)
    assert torch.equal(Hn, Hf) and torch.equal(v, vf)


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

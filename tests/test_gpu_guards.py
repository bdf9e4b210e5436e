"""Out-of-bounds check of every kernel's writes without compute-sanitizer (closed on this GPU pool: "runs under it
have left GPUs needing a reset"): each output / workspace buffer handed to the C ABI sits between two guard bands
filled with a sentinel bit pattern; after the call the bands must be untouched and the payload fully written.
Sizes are chosen to hit the tail paths (n not a multiple of the tile, of 4, of the warp) in both layouts."""
import ctypes as C

import numpy as np
import pytest
import torch

from aircraft_b200 import _cabi
from aircraft_b200.config import TRIM_COM
from aircraft_b200.params import make_cost_params, make_sweep_params
from aircraft_b200.workloads import make_control_sequences, make_controls, make_states

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
GUARD = 4096                       # floats on either side
SENT = -1.2345678e30               # never produced by the kernels


class Guarded:
    """A float32 device buffer of `n` floats between two guard bands (the payload starts 16-byte aligned)."""

    def __init__(self, n, fill=SENT):
        self.n = int(n)
        self.buf = torch.full((self.n + 2 * GUARD,), SENT, dtype=torch.float32, device=DEV)
        self.view = self.buf[GUARD:GUARD + self.n]
        if fill != SENT:
            self.view.fill_(fill)

    @property
    def ptr(self):
        return self.view.data_ptr()

    def check(self, name, written=True):
        torch.cuda.synchronize()
        lo, hi = self.buf[:GUARD], self.buf[GUARD + self.n:]
        assert bool((lo == SENT).all()) and bool((hi == SENT).all()), f"{name}: guard band overwritten"
        if written and self.n:
            assert not bool((self.view == SENT).any()), f"{name}: payload not fully written"


@pytest.fixture(scope="module")
def ac(built):
    from aircraft_b200 import Aircraft
    return Aircraft.from_problem(com=TRIM_COM, device=DEV, physical_integration_substeps=1, normalise=True)


def f32(a):
    return torch.tensor(np.ascontiguousarray(a), dtype=torch.float32, device=DEV)


@pytest.mark.parametrize("layout", [_cabi.LAYOUT_SOA, _cabi.LAYOUT_AOS])
@pytest.mark.parametrize("n,substeps", [(1, 1), (131, 1), (1028, 1), (20000, 1), (517, 3)])
def test_linearize_writes_stay_in_bounds(ac, layout, n, substeps):
    x, u = make_states(n, seed=7, perturb_norm=1e-3), make_controls(n, seed=7)
    if layout == _cabi.LAYOUT_SOA:
        x, u = x.T, u.T
    xd, ud = f32(x), f32(u)
    xn, A, Bm, Dt = Guarded(13 * n), Guarded(169 * n), Guarded(91 * n), Guarded(13 * n)
    lib, h, st = ac._lib, ac.handle, torch.cuda.current_stream().cuda_stream
    rc = lib.airc_linearize_dt(h, xd.data_ptr(), ud.data_ptr(), 0.01, None, n, substeps, 1, layout, xn.ptr, A.ptr,
                               Bm.ptr, Dt.ptr, st)
    assert rc == 0
    for nm, g in (("x_next", xn), ("A", A), ("Bm", Bm), ("dPhi_ddt", Dt)):
        g.check(nm)
    # without the d/d(dt) column (other kernel instantiation) and with an unaligned A (scalar store path)
    A2, B2 = Guarded(169 * n + 1), Guarded(91 * n)
    rc = lib.airc_linearize(h, xd.data_ptr(), ud.data_ptr(), 0.01, None, n, substeps, 1, layout, None,
                            A2.ptr + 4, B2.ptr, st)
    assert rc == 0
    A2.check("A (unaligned)", written=False)
    B2.check("Bm")
    assert torch.equal(A2.view[1:], A.view) and torch.equal(B2.view, Bm.view)
    # defects: the residual output
    xt, r = f32(x), Guarded(13 * n)
    rc = lib.airc_defects(h, xd.data_ptr(), ud.data_ptr(), xt.data_ptr(), 0.01, None, n, substeps, 1, layout, r.ptr,
                          None, None, None, st)
    assert rc == 0
    r.check("defect")


@pytest.mark.parametrize("B,H", [(1, 7), (333, 16), (4099, 40), (125000, 24)])
def test_sweep_writes_stay_in_bounds(ac, B, H):
    lib, h, st = ac._lib, ac.handle, torch.cuda.current_stream().cuda_stream
    x0 = f32(make_states(1, seed=4, level=True)[0])
    Un = f32(make_control_sequences(1, H, seed=4, sigma=0.05)[0])
    sp, cp = make_sweep_params(seed=9), make_cost_params()
    wsb = int(lib.airc_sweep_workspace_bytes(B, H, sp.ckpt_interval))
    ws = Guarded((wsb + 3) // 4)
    rec, cost = Guarded(int(lib.airc_sweep_record_floats(H))), Guarded(B)
    rc = lib.airc_sweep(h, x0.data_ptr(), Un.data_ptr(), 0.01, B, H, 1, 1, C.byref(sp), C.byref(cp), ws.ptr, wsb,
                        cost.ptr, rec.ptr, st)
    assert rc == 0
    ws.check("workspace", written=False)
    rec.check("record")
    cost.check("cost")


@pytest.mark.parametrize("layout", [_cabi.LAYOUT_SOA, _cabi.LAYOUT_AOS])
@pytest.mark.parametrize("B,H,flags", [(1, 5, 1), (77, 9, 1), (4097, 12, 1), (2049, 6, 1 | 8), (300, 6, 1 | 4)])
def test_rollout_writes_stay_in_bounds(ac, layout, B, H, flags):
    lib, h, st = ac._lib, ac.handle, torch.cuda.current_stream().cuda_stream
    x0, U = make_states(B, seed=5), make_control_sequences(B, H, seed=5)
    if layout == _cabi.LAYOUT_SOA:
        x0, U = x0.T, U.transpose(1, 2, 0)
    xd, Ud = f32(x0), f32(U)
    X, xf, cost = Guarded(13 * (H + 1) * B), Guarded(13 * B), Guarded(B)
    cp = make_cost_params()
    rc = lib.airc_rollout(h, xd.data_ptr(), Ud.data_ptr(), 0.01, B, H, 1, flags, layout, X.ptr, xf.ptr, cost.ptr,
                          C.byref(cp), st)
    assert rc == 0
    for nm, g in (("X", X), ("x_final", xf), ("cost", cost)):
        g.check(nm)


@pytest.mark.parametrize("layout", [_cabi.LAYOUT_SOA, _cabi.LAYOUT_AOS])
def test_node_kernels_and_hessian_stay_in_bounds(ac, layout):
    lib, h, st = ac._lib, ac.handle, torch.cuda.current_stream().cuda_stream
    n = 1237
    x, u = make_states(n, seed=6), make_controls(n, seed=6)
    lam = np.random.default_rng(6).normal(size=(n, 13))
    if layout == _cabi.LAYOUT_SOA:
        x, u, lam = x.T, u.T, lam.T
    xd, ud, ld = f32(x), f32(u), f32(lam)
    f, phi, dg = Guarded(13 * n), Guarded(13 * n), Guarded(_cabi.NDIAG * n)
    assert lib.airc_state_derivative(h, xd.data_ptr(), ud.data_ptr(), n, layout, f.ptr, st) == 0
    assert lib.airc_state_update(h, xd.data_ptr(), ud.data_ptr(), 0.01, None, n, 1, 1, layout, phi.ptr, st) == 0
    assert lib.airc_diagnostics(h, xd.data_ptr(), ud.data_ptr(), n, layout, dg.ptr, st) == 0
    H, g, v = Guarded(441 * n), Guarded(21 * n), Guarded(n)
    assert lib.airc_hessian(h, xd.data_ptr(), ud.data_ptr(), ld.data_ptr(), 0.01, None, n, 1, 1, layout, H.ptr, g.ptr,
                            v.ptr, st) == 0
    for nm, b in (("f", f), ("phi", phi), ("diagnostics", dg), ("H", H), ("grad", g), ("value", v)):
        b.check(nm)

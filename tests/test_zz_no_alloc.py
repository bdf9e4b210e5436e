"""airc_sweep_static -- the sampling sweep on the library's own file-scope device workspace (no cudaMalloc at call
time; include/aircraft_b200.h).  Called through the C ABI and compared with airc_sweep on a caller workspace: same
kernels, same segmentation, so the per-trajectory costs and the winner's record must agree (the tolerance is the
one test_gpu_sweep.py uses between two sweep launches; in practice the bits are equal).  Two calls in a row of
DIFFERENT shape exercise the shared array's stale contents; a third repeats the first.  Collected last on purpose
(file name), and self-contained: it passes alone in a fresh process and after the rest of the suite."""
import ctypes as C

import numpy as np
import pytest
import torch

from aircraft_b200 import _cabi
from aircraft_b200.config import TRIM_COM
from aircraft_b200.params import make_cost_params, make_sweep_params
from aircraft_b200.workloads import make_control_sequences, make_states

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
TOL = 1e-6


def f32(a):
    return torch.tensor(np.ascontiguousarray(a), dtype=torch.float32, device=DEV)


@pytest.fixture(scope="module")
def ac(built):
    from aircraft_b200 import Aircraft
    assert torch.cuda.is_available(), "the -m gpu tests need a B200"
    return Aircraft.from_problem(com=TRIM_COM, device=DEV, physical_integration_substeps=1, normalise=True)


def _both(ac, B, H, seed):
    """(cost, record) of airc_sweep on a torch-allocated workspace and of airc_sweep_static, same inputs"""
    lib, h, st = ac._lib, ac.handle, torch.cuda.current_stream().cuda_stream
    x0 = f32(make_states(1, seed=seed, level=True)[0])
    Un = f32(make_control_sequences(1, H, seed=seed, sigma=0.05)[0])
    sp, cp = make_sweep_params(seed=seed + 100), make_cost_params()
    R = int(lib.airc_sweep_record_floats(H))
    wsb = int(lib.airc_sweep_workspace_bytes(B, H, sp.ckpt_interval))
    assert wsb <= _cabi.SWEEP_STATIC_WORKSPACE_BYTES
    ws = torch.empty(wsb, dtype=torch.uint8, device=DEV)
    cost_a, rec_a = torch.full((B,), -7.0, device=DEV), torch.full((R,), -7.0, device=DEV)
    cost_b, rec_b = torch.full((B,), -9.0, device=DEV), torch.full((R,), -9.0, device=DEV)
    rc = lib.airc_sweep(h, x0.data_ptr(), Un.data_ptr(), 0.01, B, H, 1, 1, C.byref(sp), C.byref(cp), ws.data_ptr(), wsb,
                        cost_a.data_ptr(), rec_a.data_ptr(), st)
    assert rc == 0, lib.airc_last_error(h)
    rc = lib.airc_sweep_static(h, x0.data_ptr(), Un.data_ptr(), 0.01, B, H, 1, 1, C.byref(sp), C.byref(cp),
                               cost_b.data_ptr(), rec_b.data_ptr(), st)
    assert rc == 0, lib.airc_last_error(h)
    torch.cuda.synchronize()
    return cost_a.cpu().numpy(), rec_a.cpu().numpy(), cost_b.cpu().numpy(), rec_b.cpu().numpy()


def _check(ca, ra, cb, rb):
    assert np.isfinite(ca).all() and np.isfinite(cb).all()
    assert np.abs(cb - ca).max() <= TOL * np.abs(ca).max()
    assert np.array_equal(ra[1:4].view(np.uint32), rb[1:4].view(np.uint32))          # cost bits and global arg-min
    assert np.abs(rb[4:] - ra[4:]).max() <= TOL * max(1.0, np.abs(ra[4:]).max())     # trajectory + controls
    assert abs(rb[0] - ra[0]) <= TOL * abs(ra[0])


def test_static_workspace_sweep_matches_airc_sweep_on_consecutive_calls(ac):
    first = _both(ac, 4099, 40, seed=4)            # segments + checkpoints, ragged last group
    _check(*first)
    second = _both(ac, 1000, 24, seed=5)           # smaller call on the array the first one left behind
    _check(*second)
    again = _both(ac, 4099, 40, seed=4)            # and back: the results of the first call, to the bit
    _check(*again)
    assert np.array_equal(first[2], again[2]) and np.array_equal(first[3], again[3])
    big = _both(ac, 16384, 200, seed=6)            # the size the header quotes (26.9 MiB of the 32 MiB)
    _check(*big)


def test_static_workspace_sweep_rejects_sizes_beyond_the_array(ac):
    lib, h, st = ac._lib, ac.handle, torch.cuda.current_stream().cuda_stream
    B, H = 125000, 200                             # BASELINE config 4 shard: ~200 MB of workspace
    sp, cp = make_sweep_params(seed=1), make_cost_params()
    assert int(lib.airc_sweep_workspace_bytes(B, H, sp.ckpt_interval)) > _cabi.SWEEP_STATIC_WORKSPACE_BYTES
    x0 = f32(make_states(1, seed=4, level=True)[0])
    Un = f32(make_control_sequences(1, H, seed=4, sigma=0.05)[0])
    rec = torch.full((int(lib.airc_sweep_record_floats(H)),), -9.0, device=DEV)
    rc = lib.airc_sweep_static(h, x0.data_ptr(), Un.data_ptr(), 0.01, B, H, 1, 1, C.byref(sp), C.byref(cp), None,
                               rec.data_ptr(), st)
    assert rc == _cabi.EINVAL
    msg = lib.airc_last_error(h)
    msg = msg.decode() if isinstance(msg, bytes) else str(msg)
    assert "AIRC_SWEEP_STATIC_WORKSPACE_BYTES" in msg and "airc_sweep" in msg
    torch.cuda.synchronize()
    assert bool((rec == -9.0).all())               # nothing was launched

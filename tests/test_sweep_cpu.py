"""CPU-side checks of the sampling sweep: the sampler specification (numpy restatement ==
oracle/cpu_ref.cpp == the product's airc_sampler.cuh compiled for the host), the CPU reference
sweep against the float64 torch oracle, record decoding/selection, and the world-size-2 exchange
under gloo with a stand-in for the GPU sweep."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from _util import oracle_for, params_for, t64
from aircraft_b200.config import TRIM_COM
from aircraft_b200.params import make_cost_params, make_sweep_params
from aircraft_b200.workloads import make_states


def mix32(x):
    x = x.astype(np.uint32)
    x ^= x >> np.uint32(16); x *= np.uint32(0x7feb352d)
    x ^= x >> np.uint32(15); x *= np.uint32(0x846ca68b)
    x ^= x >> np.uint32(16)
    return x


def sampler_numpy(sp, Unom, first, count):
    """include/aircraft_b200.h (AircSweepParams) restated with numpy integer arithmetic.  The FP32
    fused multiply-add of the walk is emulated in float64 (exact product, one extra rounding;
    see the test for how rare a disagreement is)."""
    H = Unom.shape[0]
    g = (np.uint64(sp.index_offset) + np.arange(first, first + count, dtype=np.uint64))
    seed_lo = mix32(np.array([sp.seed & 0xFFFFFFFF], np.uint64).astype(np.uint32))[0]
    with np.errstate(over="ignore"):
        tkey = mix32((g & np.uint64(0xFFFFFFFF)).astype(np.uint32) + seed_lo) ^ np.uint32(sp.seed >> 32)
        U = np.broadcast_to(Unom[None], (count, H, 7)).astype(np.float32).copy()
        d = np.zeros((count, 4), np.float32)
        for k in range(H):
            cnt = tkey + np.uint32((k * 0x9E3779B9) & 0xFFFFFFFF)
            for c, ch in enumerate((0, 1, 2, 6)):
                w = mix32(cnt + np.uint32((c * 0x7F4A7C15) & 0xFFFFFFFF))
                n = ((w & 255) + ((w >> 8) & 255) + ((w >> 16) & 255) + (w >> 24)).astype(np.int64) - 510
                s0 = np.float32(float(sp.sigma0[c]) / 147.80053)
                s = np.float32(float(sp.sigma[c]) / 147.80053)
                if k == 0:
                    d[:, c] = s0 * n.astype(np.float32)
                else:
                    d[:, c] = (np.float64(s) * n.astype(np.float64) + d[:, c].astype(np.float64)).astype(np.float32)
                t = Unom[k, ch] + d[:, c]
                U[:, k, ch] = np.minimum(np.maximum(t, np.float32(sp.lo[c])), np.float32(sp.hi[c]))
    return U


@pytest.fixture(scope="module")
def refs(built):
    from oracle.cpu_ref import CpuRef, HostEmul
    p, blob = params_for(TRIM_COM)
    return CpuRef(p, blob), HostEmul(p, blob)


def test_sampler_three_implementations_agree(refs):
    ref, emul = refs
    H = 50
    Un = np.zeros((H, 7), np.float32)
    Un[:, 0] = np.linspace(-1, 1, H); Un[:, 4] = 0.25; Un[:, 6] = 0.5
    for kw, first, count in [(dict(seed=1), 0, 512), (dict(seed=(5 << 32) | 99, index_offset=10**6), 3, 200)]:
        sp = make_sweep_params(**kw)
        a = ref.sweep_controls(sp, Un, first, count)
        b = emul.sweep_controls(sp, Un, first, count)
        assert np.array_equal(a, b)                       # cpu_ref restatement == product header
        c = sampler_numpy(sp, Un, first, count)
        # double rounding in the numpy emulation of fmaf can flip the last bit of a walk offset
        # (probability ~2^-29 per operation); allow a handful of 1-ulp disagreements at most
        bad = a != c
        assert bad.mean() < 1e-5 and np.abs(a - c).max() <= 1e-6
        assert np.array_equal(a[:, 0], c[:, 0])           # step 0 has no fma: exact
    # distribution: unit variance of the normalised noise, independence across slots / neighbours
    sp = make_sweep_params(seed=7, sigma0=(1, 1, 1, 1), sigma=(0, 0, 0, 0), lo=(-9,) * 4, hi=(9,) * 4)
    z = ref.sweep_controls(sp, np.zeros((1, 7), np.float32), 0, 200_000)[:, 0, [0, 1, 2, 6]]
    assert np.allclose(z.std(0), 1.0, atol=0.01) and np.abs(z.mean(0)).max() < 0.01
    cc = np.corrcoef(z.T)
    assert np.abs(cc - np.eye(4)).max() < 0.01
    assert abs(np.corrcoef(z[:-1, 0], z[1:, 0])[0, 1]) < 0.01
    assert np.abs(z).max() <= 510 / 147.80053 + 1e-6


def test_cpu_reference_sweep_matches_torch_oracle(refs):
    ref, _ = refs
    orc = oracle_for(TRIM_COM)
    x0 = make_states(1, seed=4, level=True)[0]
    B, H = 48, 30
    Un = np.zeros((H, 7), np.float32); Un[:, 1] = -1.0; Un[:, 3] = np.linspace(0, 1, H)
    sp, cp = make_sweep_params(seed=3), make_cost_params(goal=(100.0, 5.0), vel_param=0.5)
    bc, bi, cost = ref.sweep(x0, Un, 0.01, B, sp, cp, normalise=True)
    U = ref.sweep_controls(sp, Un, 0, B).astype(np.float64)
    X = orc.rollout(t64(np.repeat(x0[None], B, 0)), t64(U), 0.01, normalise=True)
    want = orc.cost(X, t64(U), goal=(100.0, 5.0), vel_param=0.5).numpy()
    assert np.allclose(cost, want, rtol=1e-9, atol=1e-6)
    assert bi == int(np.argmin(want)) and abs(bc - want.min()) < 1e-6 * abs(want.min())
    c2 = ref.rollout_cost(np.repeat(x0[None], B, 0), U, 0.01, cp, normalise=True)
    assert np.allclose(c2, want, rtol=1e-9, atol=1e-6)


def _record(cost, gidx, H, fill):
    from aircraft_b200.sharding import _orderable
    R = 4 + 13 * (H + 1) + 7 * H
    r = torch.full((R,), float(fill), dtype=torch.float32)
    r[0] = cost
    key = int(_orderable(torch.tensor([cost], dtype=torch.float32))[0])
    words = torch.tensor([key - (1 << 32) if key >= (1 << 31) else key,
                          (gidx & 0xFFFFFFFF) - (1 << 32) if (gidx & 0xFFFFFFFF) >= (1 << 31) else gidx & 0xFFFFFFFF,
                          gidx >> 32], dtype=torch.int32)
    r[1:4] = words.view(torch.float32)
    return r


def test_record_decode_and_selection():
    from aircraft_b200.aircraft import decode_record
    from aircraft_b200.sharding import select_record_torch
    H = 6
    recs = torch.stack([_record(3.0, 40, H, 1), _record(-2.5, (1 << 33) + 5, H, 2), _record(-2.5, 7, H, 3),
                        _record(float("nan"), 0, H, 4)])
    best = select_record_torch(recs)
    d = decode_record(best, H)
    assert float(d.min_cost) == -2.5 and int(d.argmin) == 7
    assert d.trajectory.shape == (H + 1, 13) and d.controls.shape == (H, 7) and float(d.controls[0, 0]) == 3.0
    d2 = decode_record(recs[1], H)
    assert int(d2.argmin) == (1 << 33) + 5
    dn = decode_record(recs[1].numpy(), H)
    assert dn.argmin == (1 << 33) + 5 and dn.min_cost == -2.5


class _FakeAircraft:
    """Stands in for the GPU sweep in the gloo test: cost(g) is a fixed function of the GLOBAL
    index, so the expected winner is known whatever the sharding."""

    def __init__(self, H):
        self.H = H

    @staticmethod
    def cost_of(g):
        return ((g * 2654435761) % 1000003) / 1000.0 - (5000.0 if g == 617 else 0.0)

    def sweep(self, x0, U_nom, dt, B, *, sampler=None, **kw):
        from aircraft_b200.aircraft import decode_record
        off = int(sampler.index_offset)
        costs = [self.cost_of(off + b) for b in range(B)]
        j = int(np.argmin(costs))
        return decode_record(_record(costs[j], off + j, self.H, off + j), self.H)


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, total, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from aircraft_b200.sharding import sweep_sampled
        H = 5
        out = sweep_sampled(_FakeAircraft(H), None, None, 0.01, total, sampler=make_sweep_params(seed=1))
        q.put((rank, float(out.min_cost), int(out.argmin), float(out.controls[0, 0])))
    finally:
        dist.destroy_process_group()


def test_sweep_sampled_gloo_world2():
    world, total = 2, 1001
    port = _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, total, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    want = min(range(total), key=lambda g: (np.float32(_FakeAircraft.cost_of(g)), g))
    for rank, mc, am, fill in res:
        assert am == want == 617 and fill == 617.0
        assert mc == np.float32(_FakeAircraft.cost_of(617))

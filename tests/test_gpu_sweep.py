"""Sampling-MPC sweep (BASELINE config 4) on the GPU, through the C ABI: the in-kernel sampler
against its independent CPU restatement (bit-exact), costs against the explicit-control kernel
and the float64 CPU reference, the checkpointed winner record, shard invariance, edge cases."""
import numpy as np
import pytest
import torch

from _util import col_err, params_for
from aircraft_b200.config import TRIM_COM
from aircraft_b200.params import make_cost_params, make_sweep_params
from aircraft_b200.workloads import make_control_sequences, make_states

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
DT = 0.01


def f32(a):
    return torch.tensor(np.asarray(a), dtype=torch.float32, device=DEV)


@pytest.fixture(scope="module")
def ac(built):
    from aircraft_b200 import Aircraft
    assert torch.cuda.is_available(), "the -m gpu tests need a B200"
    return Aircraft.from_problem(com=TRIM_COM, device=DEV, physical_integration_substeps=1, normalise=True)


@pytest.fixture(scope="module")
def ref(built):
    from oracle.cpu_ref import CpuRef
    p, blob = params_for(TRIM_COM)
    return CpuRef(p, blob)


@pytest.fixture(scope="module")
def x0():
    return make_states(1, seed=4, level=True)[0]


def nominal(H, seed=9):
    """a smooth, non-trivial nominal sequence incl. non-zero (inert) thrust rows"""
    U = make_control_sequences(1, H, seed=seed, sigma=0.05)[0].astype(np.float32)
    U[:, 3] = np.linspace(0.0, 0.3, H, dtype=np.float32)
    U[:, 4] = 0.05 * np.sin(np.arange(H, dtype=np.float32))
    return U


def test_sampler_bit_exact_vs_cpu_restatement(ac, ref):
    H = 37
    Un = nominal(H)
    for kw, first, count in [(dict(seed=1), 0, 1000), (dict(seed=2**40 + 12345, index_offset=777), 5, 333),
                             (dict(seed=3, sigma0=(3, 3, 3, 1), sigma=(1, 1, 1, .5)), 2**31 - 10, 64)]:
        sp = make_sweep_params(**kw)
        got = ac.sweep_controls(f32(Un), sp, first, count).cpu().numpy()
        want = ref.sweep_controls(sp, Un, first, count)
        assert np.array_equal(got, want), kw
        soa = ac.sweep_controls(f32(Un), sp, first, count, layout="soa").cpu().numpy()
        assert np.array_equal(soa.transpose(2, 0, 1), want)
        live = want[:, :, [0, 1, 2, 6]]
        assert live.min() >= -5.0 and live.max() <= 5.0 and want[:, :, 6].min() >= 0.0 and want[:, :, 6].max() <= 1.0
        assert np.array_equal(want[:, :, 3:6], np.broadcast_to(Un[None, :, 3:6], want[:, :, 3:6].shape))


def test_sweep_costs_match_explicit_kernel_and_cpu_reference(ac, ref, x0):
    B, H = 4096 + 17, 40
    Un = nominal(H)
    sp, cp = make_sweep_params(seed=11), make_cost_params()
    r = ac.sweep(f32(x0), f32(Un), DT, B, sampler=sp, cost=cp, return_costs=True)
    cost = r.cost.cpu().numpy().astype(np.float64)
    # (1) the explicit-control kernel on the regenerated sequences (one-trajectory-per-thread kernel
    #     at this size: same mathematics, different rounding)
    U = ac.sweep_controls(f32(Un), sp, 0, B, layout="soa")
    xs = f32(np.repeat(x0[:, None], B, 1))
    e = ac.rollout(xs, U, DT, layout="soa", cost=cp, return_final=False).cost.cpu().numpy().astype(np.float64)
    assert np.abs(cost - e).max() <= 1e-5 * np.abs(e).max()
    # (2) float64 CPU reference of the whole sweep (its own sampler, its own loss)
    bc, bi, c64 = ref.sweep(x0, Un, DT, B, sp, cp, normalise=True)
    assert np.abs(cost - c64).max() <= 1e-5 * np.abs(c64).max()
    # (3) fused (min, arg-min)
    am = int(r.argmin.item())
    assert am == int(np.argmin(cost)) and float(r.min_cost.item()) == np.float32(cost.min())
    order = np.sort(c64)
    if order[1] - order[0] > 2e-5 * abs(order[0]):
        assert am == bi
    # (4) the record is the winner's trajectory and controls
    Uw = ref.sweep_controls(sp, Un, am, 1)[0]
    assert np.array_equal(r.controls.cpu().numpy(), Uw)
    Xw = ref.rollout(x0[None], Uw[None].astype(np.float64), DT, normalise=True)[0]
    assert col_err(r.trajectory.cpu().numpy(), Xw) < 5e-5      # 40 free-running FP32 steps
    one = ac.rollout(f32(x0[None]), f32(Uw[None]), DT, return_trajectory=True).trajectory[0]
    assert col_err(r.trajectory.cpu().numpy(), one.cpu().numpy()) < 2e-5


def test_full_batch_geometry_costs_match_cpu_reference(ac, ref, x0):
    """The launch geometry of BASELINE config 4's shard (125 000 trajectories: 1 954 groups of 64 on more than one
    wave, time-sliced into segments with slice hand-over, checkpoints on) at a horizon short enough that FP32 and
    float64 have not yet drifted apart: ALL 125 000 costs against the float64 CPU sweep at the parity bar, and the
    fused arg-min against it."""
    B, H = 125000, 16
    Un = nominal(H)
    sp, cp = make_sweep_params(seed=21), make_cost_params()
    r = ac.sweep(f32(x0), f32(Un), DT, B, sampler=sp, cost=cp, return_costs=True)
    cost = r.cost.cpu().numpy().astype(np.float64)
    bc, bi, c64 = ref.sweep(x0, Un, DT, B, sp, cp, normalise=True)
    assert np.isfinite(c64).all()
    err = np.abs(cost - c64).max() / np.abs(c64).max()
    print(f"125000 x {H} sweep: max |cost - cost_f64| / max |cost_f64| = {err:.2e}")
    assert err <= 1e-5
    am = int(r.argmin.item())
    assert am == int(np.argmin(cost)) and float(r.min_cost.item()) == np.float32(cost.min())
    order = np.sort(c64)
    if order[1] - order[0] > 2e-5 * abs(order[0]):
        assert am == bi
    Uw = ref.sweep_controls(sp, Un, am, 1)[0]
    assert np.array_equal(r.controls.cpu().numpy(), Uw)
    Xw = ref.rollout(x0[None], Uw[None].astype(np.float64), DT, normalise=True)[0]
    assert col_err(r.trajectory.cpu().numpy(), Xw) < 2e-5


@pytest.mark.parametrize("ckpt", [1, 7, 8, 40, 1000])
def test_checkpoint_interval_does_not_change_the_record(ac, x0, ckpt):
    B, H = 3000, 40
    Un = nominal(H)
    base = ac.sweep(f32(x0), f32(Un), DT, B, sampler=make_sweep_params(seed=5, ckpt_interval=8), return_costs=True)
    r = ac.sweep(f32(x0), f32(Un), DT, B, sampler=make_sweep_params(seed=5, ckpt_interval=ckpt), return_costs=True)
    assert torch.equal(r.cost, base.cost)
    assert torch.equal(r.raw[:4], base.raw[:4])
    assert torch.equal(r.raw, base.raw)


def test_minimal_workspace_rerolls_sequentially(ac, x0):
    B, H = 2000, 33
    Un = nominal(H)
    sp = make_sweep_params(seed=6)
    base = ac.sweep(f32(x0), f32(Un), DT, B, sampler=sp)
    import ctypes as C
    ws = torch.empty(256, dtype=torch.uint8, device=DEV)
    rec = torch.empty_like(base.raw)
    cp = make_cost_params()
    xd, Ud = f32(x0), f32(Un)          # keep the device buffers alive across the raw C call
    rc = ac._lib.airc_sweep(ac.handle, xd.data_ptr(), Ud.data_ptr(), DT, B, H, 1, 1, C.byref(sp),
                            C.byref(cp), ws.data_ptr(), ws.numel(), None, rec.data_ptr(), None)
    assert rc == 0
    torch.cuda.synchronize()
    assert torch.equal(rec, base.raw)


def test_shard_invariance_and_select_record(ac, x0):
    B, H = 5001, 25
    Un = nominal(H)
    whole = ac.sweep(f32(x0), f32(Un), DT, B, sampler=make_sweep_params(seed=21), return_costs=True)
    cut = 1777
    a = ac.sweep(f32(x0), f32(Un), DT, cut, sampler=make_sweep_params(seed=21, index_offset=0), return_costs=True)
    b = ac.sweep(f32(x0), f32(Un), DT, B - cut, sampler=make_sweep_params(seed=21, index_offset=cut), return_costs=True)
    assert torch.equal(torch.cat([a.cost, b.cost]), whole.cost)
    best = ac.select_record(torch.stack([b.raw, a.raw]))
    assert torch.equal(best, whole.raw)
    # ties resolve to the lower global index; NaN costs sort last
    r1, r2 = whole.raw.clone(), whole.raw.clone()
    r2[2:4] = torch.tensor([7, 0], dtype=torch.int32, device=DEV).view(torch.float32)
    r1[2:4] = torch.tensor([9, 0], dtype=torch.int32, device=DEV).view(torch.float32)
    assert torch.equal(ac.select_record(torch.stack([r1, r2])), r2)
    r3 = whole.raw.clone()
    r3[1] = torch.tensor([-1], dtype=torch.int32, device=DEV).view(torch.float32)[0]   # 0xFFFFFFFF = NaN key
    assert torch.equal(ac.select_record(torch.stack([r3, r1])), r1)


def test_host_entry_point_equals_device_path(ac, x0):
    B, H = 1500, 20
    Un = nominal(H)
    sp = make_sweep_params(seed=8)
    d = ac.sweep(f32(x0), f32(Un), DT, B, sampler=sp, return_costs=True)
    h = ac.sweep(x0, Un, DT, B, sampler=sp, return_costs=True)
    assert np.array_equal(h.raw, d.raw.cpu().numpy())
    assert np.array_equal(h.cost, d.cost.cpu().numpy())
    assert h.argmin == int(d.argmin.item())


def test_edge_cases(ac, x0):
    Un = nominal(10)
    r = ac.sweep(f32(x0), f32(Un), DT, 0)                       # empty sweep: header only, NaN-key
    torch.cuda.synchronize()
    assert int(r.raw[1:2].view(torch.int32).item()) == -1
    r = ac.sweep(f32(x0), f32(Un[:0]), DT, 64, return_costs=True)   # H = 0: cost of x0 itself
    assert r.trajectory.shape == (1, 13) and torch.equal(r.trajectory[0].cpu(), torch.tensor(x0, dtype=torch.float32))
    assert torch.all(r.cost == r.cost[0]) and int(r.argmin.item()) == 0
    r = ac.sweep(f32(x0), f32(Un), DT, 1, return_costs=True)    # one trajectory
    assert int(r.argmin.item()) == 0 and float(r.min_cost.item()) == float(r.cost[0].item())
    with pytest.raises(RuntimeError):
        ac.sweep(f32(x0[:12]), f32(Un), DT, 4)
    with pytest.raises(RuntimeError):
        ac.sweep(f32(x0), f32(Un[:, :6]), DT, 4)


def test_full_size_shard_properties(ac, x0):
    """BASELINE config 4's per-GPU shard at full size: size-independent properties."""
    B, H = 125_000, 200
    Un = nominal(H)
    sp = make_sweep_params(seed=2025)
    r = ac.sweep(f32(x0), f32(Un), DT, B, sampler=sp, return_costs=True)
    cost = r.cost
    assert torch.isfinite(cost).all()
    am = int(r.argmin.item())
    assert am == int(torch.argmin(cost).item()) and r.min_cost.item() == cost.min().item()
    # the record reproduces the winner: explicit rollout of its regenerated controls
    Uw = ac.sweep_controls(f32(Un), sp, am, 1)
    assert torch.equal(Uw[0], r.controls)
    one = ac.rollout(f32(x0[None]), Uw, DT, return_trajectory=True, cost=True)
    assert col_err(r.trajectory.cpu().numpy(), one.trajectory[0].cpu().numpy()) < 1e-4      # 200 free FP32 steps, two kernels
    assert abs(one.cost.item() - r.min_cost.item()) <= 1e-4 * abs(r.min_cost.item())
    # determinism
    r2 = ac.sweep(f32(x0), f32(Un), DT, B, sampler=sp, return_costs=True)
    assert torch.equal(r2.raw, r.raw) and torch.equal(r2.cost, cost)


# ---------------------------------------------------------------- packed (two-trajectory) rollout kernel
class _Packed:
    """The aircraft with explicit-control rollouts on the packed kernel (AIRC_FLAG_PACKED; by default it is chosen
    for batches of >= 2 waves only)."""

    def __init__(self, ac):
        self._ac = ac

    def rollout(self, *a, **kw):
        return self._ac.rollout(*a, packed=True, **kw)

    def __getattr__(self, name):
        return getattr(self._ac, name)


@pytest.fixture(scope="module")
def ac_packed(ac):
    return _Packed(ac)


def test_packed_rollout_parity_and_lane_independence(ac_packed, ac, ref):
    n, H = 2049, 30       # odd batch: the last thread has a dead hi half
    x0s = make_states(n, seed=77, perturb_norm=1e-3)
    U = make_control_sequences(n, H, seed=77)
    r = ac_packed.rollout(f32(x0s), f32(U), DT, return_trajectory=True, cost=True)
    Xr = ref.rollout(x0s, U, DT, normalise=True)
    got = r.trajectory.cpu().numpy()
    assert col_err(got[:, 1], Xr[:, 1]) < 1e-5                      # one step: the 1e-5 bar
    assert col_err(got, Xr, batch_axes=(0, 1)) < 5e-5               # 30 free-running FP32 steps
    cr = ref.rollout_cost(x0s, U, DT, make_cost_params(), normalise=True)
    assert np.abs(r.cost.cpu().numpy() - cr).max() <= 1e-5 * np.abs(cr).max()
    # substeps and no normalisation
    r4 = ac_packed.rollout(f32(x0s[:500]), f32(U[:500, :5]), 0.05, substeps=4, normalise=False, return_trajectory=True)
    X4 = ref.rollout(x0s[:500], U[:500, :5], 0.05, substeps=4, normalise=False)
    assert col_err(r4.trajectory.cpu().numpy(), X4, batch_axes=(0, 1)) < 2e-5
    # a trajectory's bits do not depend on its lane partner or on the batch it is part of
    for j in (0, 31, 32, 1000, n - 1):
        one = ac_packed.rollout(f32(x0s[j:j + 1]), f32(U[j:j + 1]), DT, return_trajectory=True, cost=True)
        assert torch.equal(one.trajectory[0], r.trajectory[j]) and torch.equal(one.cost[0], r.cost[j])
    # SoA == AoS, final state == last trajectory row
    rs = ac_packed.rollout(f32(x0s.T), f32(U.transpose(1, 2, 0)), DT, layout="soa", return_trajectory=True, cost=True)
    assert torch.equal(rs.trajectory.permute(2, 0, 1), r.trajectory) and torch.equal(rs.cost, r.cost)
    assert torch.equal(r.final, r.trajectory[:, -1])
    # NaN stays in its own trajectory (lane partner 32 untouched)
    xb = x0s.copy(); xb[0, 6:10] = np.nan
    rb = ac_packed.rollout(f32(xb), f32(U), DT, cost=True)
    assert torch.isnan(rb.cost[0]) and torch.equal(rb.cost[1:], r.cost[1:])


def test_sweep_equals_packed_explicit_rollout(ac_packed, x0):
    B, H = 4096 + 17, 40
    Un = nominal(H)
    sp, cp = make_sweep_params(seed=11), make_cost_params()
    r = ac_packed.sweep(f32(x0), f32(Un), DT, B, sampler=sp, cost=cp, return_costs=True)
    U = ac_packed.sweep_controls(f32(Un), sp, 0, B, layout="soa")
    e = ac_packed.rollout(f32(np.repeat(x0[:, None], B, 1)), U, DT, layout="soa", cost=cp, return_trajectory=True)
    # identical arithmetic except the summation order of the (shared) thrust-row penalty
    assert (r.cost - e.cost).abs().max().item() <= 2e-6 * e.cost.abs().max().item()
    am = int(r.argmin.item())
    assert torch.equal(e.trajectory[:, :, am], r.trajectory)


@pytest.mark.parametrize("B", [300, 125000])          # FFMA2 sweep kernel / tensor-core sweep kernel
def test_nan_in_the_nominal_sequence_propagates(ac, x0, B):
    """The sampler's clamp keeps NaN (min.NaN / max.NaN): a NaN in U_nom makes every candidate's cost NaN instead of
    silently sampling around the lower actuator limit.  On the tensor-core kernel the NaN also has to survive the
    TF32 operand split, the MMAs and the clamped logistic."""
    H = 12
    Un = nominal(H)
    Un[5, 1] = np.nan
    r = ac.sweep(f32(x0), f32(Un), DT, B, sampler=make_sweep_params(seed=4), return_costs=True)
    assert torch.isnan(r.cost).all()
    U = ac.sweep_controls(f32(Un), make_sweep_params(seed=4), 0, 8)
    assert torch.isnan(U[:, 5, 1]).all() and torch.isfinite(U[:, :5]).all()


def test_sweep_plan_graph_replay_equals_eager_sweep(ac, x0):
    """SweepPlan (aircraft_b200/sharding.py): the sweep captured in a CUDA graph and replayed -- same record and
    costs as the eager call, for new inputs copied into the plan's buffers between replays."""
    from aircraft_b200.sharding import SweepPlan
    B, H = 5000, 24
    sp, cp = make_sweep_params(seed=31), make_cost_params()
    plan = SweepPlan(ac, H, DT, B, sampler=sp, cost=cp, return_costs=True)
    for seed in (9, 10):
        Un = nominal(H, seed=seed)
        plan.x0.copy_(f32(x0))
        plan.U_nom.copy_(f32(Un))
        got = plan.replay()
        torch.cuda.synchronize()
        want = ac.sweep(f32(x0), f32(Un), DT, B, sampler=sp, cost=cp, return_costs=True)
        assert torch.equal(got.raw, want.raw) and torch.equal(got.cost, want.cost)
        assert int(got.argmin) == int(want.argmin)


def test_tensor_core_sweep_against_ffma2_sweep_and_nan_state(ac, x0):
    """The same candidates through both sweep kernels: a shard just above the tensor-core threshold and the same
    shard cut in two halves below it (FFMA2 kernel; the sampler is keyed on the global index).  Costs agree to the
    parity bar, the arg-min is the same candidate unless the two best are closer than that.  A NaN initial state
    makes every cost NaN on the tensor path too."""
    sm = torch.cuda.get_device_properties(0).multi_processor_count
    B, H = 3 * 256 * sm + 64, 20
    Un = nominal(H)
    sp, cp = make_sweep_params(seed=77), make_cost_params()
    whole = ac.sweep(f32(x0), f32(Un), DT, B, sampler=sp, cost=cp, return_costs=True)
    half = B // 2
    sp2 = make_sweep_params(seed=77, index_offset=half)
    a = ac.sweep(f32(x0), f32(Un), DT, half, sampler=sp, cost=cp, return_costs=True)
    b = ac.sweep(f32(x0), f32(Un), DT, B - half, sampler=sp2, cost=cp, return_costs=True)
    parts = torch.cat([a.cost, b.cost])
    assert torch.isfinite(whole.cost).all()
    err = float((whole.cost - parts).abs().max() / parts.abs().max())
    print(f"tensor-core sweep vs FFMA2 sweep, {B} x {H}: max cost difference {err:.2e} of the largest cost")
    assert err < 1e-5
    order = torch.sort(parts).values
    if float(order[1] - order[0]) > 2e-5 * float(order[0].abs()):
        assert int(whole.argmin) == int(torch.argmin(parts))
    xb = x0.copy()
    xb[4] = np.nan
    assert torch.isnan(ac.sweep(f32(xb), f32(Un), DT, B, sampler=sp, cost=cp, return_costs=True).cost).all()

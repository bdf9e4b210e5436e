"""The C++ float64 restatement that bench.py times as the CPU baseline (oracle/cpu_ref.cpp) must
agree with the torch float64 oracle to rounding: two independent restatements of the same
reference lines."""
import numpy as np
import pytest
import torch

from _util import oracle_for, params_for, t64
from aircraft_b200.workloads import make_control_sequences, make_controls, make_states
from oracle.cpu_ref import CpuRef


@pytest.fixture(scope="module")
def pair(built, com):
    p, blob = params_for(com)
    return CpuRef(p, blob, threads=2), oracle_for(com)


def test_state_derivative(pair):
    ref, orc = pair
    x, u = make_states(257, seed=1, perturb_norm=1e-3), make_controls(257, seed=1)
    got = ref.state_derivative(x, u)
    want = orc.state_derivative(t64(x), t64(u)).numpy()
    assert np.abs(got - want).max() < 1e-9 * max(1.0, np.abs(want).max())


@pytest.mark.parametrize("substeps,normalise", [(1, False), (1, True), (10, False), (4, True)])
def test_state_update(pair, substeps, normalise):
    ref, orc = pair
    x, u = make_states(129, seed=2), make_controls(129, seed=2)
    got = ref.state_update(x, u, 0.01, substeps, normalise)
    want = orc.state_update(t64(x), t64(u), 0.01, substeps, normalise).numpy()
    assert np.abs(got - want).max() < 1e-10 * max(1.0, np.abs(want).max())


def test_rollout_and_shared_controls(pair):
    ref, orc = pair
    x0 = make_states(16, seed=3)
    U = make_control_sequences(16, 20, seed=3)
    got = ref.rollout(x0, U, 0.01)
    want = orc.rollout(t64(x0), t64(U), 0.01).numpy()
    assert np.abs(got - want).max() < 1e-8 * np.abs(want).max()
    got = ref.rollout(x0, U[0], 0.01, normalise=True)
    want = orc.rollout(t64(x0), t64(U[0]), 0.01, normalise=True).numpy()
    assert np.abs(got - want).max() < 1e-8 * np.abs(want).max()


@pytest.mark.parametrize("substeps,normalise", [(1, False), (2, True)])
def test_linearize(pair, substeps, normalise):
    ref, orc = pair
    x, u = make_states(33, seed=4, perturb_norm=1e-3), make_controls(33, seed=4)
    xn, A, B = ref.linearize(x, u, 0.01, substeps, normalise)
    xo, Ao, Bo = orc.linearize(t64(x), t64(u), 0.01, substeps, normalise)
    assert np.abs(xn - xo.numpy()).max() < 1e-9
    assert np.abs(A - Ao.numpy()).max() < 1e-9 * np.abs(Ao.numpy()).max()
    assert np.abs(B - Bo.numpy()).max() < 1e-9 * max(1.0, np.abs(Bo.numpy()).max())


def test_stall_scaling(built):
    p, blob = params_for(stall_scaling=True)
    ref, orc = CpuRef(p, blob, threads=1), oracle_for(stall_scaling=True)
    x, u = make_states(64, seed=5), make_controls(64, seed=5)
    # push some samples to high alpha so that the logistic is active
    x[:32, 3:6] *= np.array([0.3, 1.0, 3.0])
    xn, A, B = ref.linearize(x, u, 0.01)
    xo, Ao, Bo = orc.linearize(t64(x), t64(u), 0.01)
    assert np.abs(xn - xo.numpy()).max() < 1e-9 * np.abs(xo.numpy()).max()
    assert np.abs(A - Ao.numpy()).max() < 1e-8 * np.abs(Ao.numpy()).max()


def test_linearize_dt_column(pair):
    ref, orc = pair
    x, u = make_states(64, seed=15, perturb_norm=1e-3).astype(np.float64), make_controls(64, seed=15).astype(np.float64)
    xn, A, B, Dt = ref.linearize(x, u, 0.02, 2, True, want_dt=True)
    xo, Ao, Bo, Do = orc.linearize(t64(x), t64(u), 0.02, 2, True, want_dt=True)
    assert np.allclose(Dt, Do.numpy(), rtol=1e-9, atol=1e-9)
    assert np.allclose(A, Ao.numpy(), rtol=1e-9, atol=1e-9) and np.allclose(B, Bo.numpy(), rtol=1e-9, atol=1e-9)
    # central difference in dt
    e = 1e-6
    fd = (ref.state_update(x, u, 0.02 + e, 2, True) - ref.state_update(x, u, 0.02 - e, 2, True)) / (2 * e)
    assert np.allclose(Dt, fd, rtol=1e-6, atol=1e-6)

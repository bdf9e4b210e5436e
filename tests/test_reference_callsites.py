"""The reference's own rollout loops, replayed VERBATIM against the drop-in facade.

The loop bodies below are the statements of
    /root/reference/main/control/control.py:72-93                     Controller.initialise        (MPC warm start)
    /root/reference/src/aircraft/control/moving_horizon.py:203-213    MHTT.initialise              (MHTT warm start)
    /root/reference/main/dynamics/dynamics.py:100-121                 the open-loop simulation of main()
with `self.dynamics = aircraft.state_update` as ControlProblem.__init__ sets it
(/root/reference/src/aircraft/control/base.py:187-189, which also reads dynamics.size1_in(0|1)).  They consume CasADi
DM values (.toarray(), .full(), np.isnan(state[0])); against round 1's bare ndarrays they raised AttributeError.

CPU part: the loops run against a float64-oracle-backed stand-in that returns aircraft_b200.DM, which pins the DM /
Function interface without a GPU.  GPU part (-m gpu): the same loops against aircraft_b200.Aircraft, compared with
the oracle's rollout at the parity bar.
"""
import numpy as np
import pytest
import torch

from _util import col_err, oracle_for, t64
from aircraft_b200.config import TRIM_COM
from aircraft_b200.dm import DM, Function, as_dm
from aircraft_b200.workloads import make_controls, make_states


# ------------------------------------------------------------------ the reference's loops (verbatim bodies)
class _Controller:
    """Stand-in for the attribute set ControlProblem.__init__ builds (control/base.py:187-197)."""

    def __init__(self, system, num_nodes, dt, extra_rows=0):
        self.dynamics = system.state_update
        self.state_dim = self.dynamics.size1_in(0)
        self.control_dim = self.dynamics.size1_in(1)
        self.x_dot = system.state_derivative
        self.num_nodes = num_nodes
        self.dt = dt

    def initialise_control_py(self, initial_state):          # main/control/control.py:72-93
        guess = np.zeros((self.state_dim + self.control_dim, self.num_nodes + 1))
        guess[13, :] = 1
        guess[14, :] = 0
        guess[:self.state_dim, 0] = initial_state.toarray().flatten()
        for i in range(self.num_nodes):
            guess[:self.state_dim, i + 1] = self.dynamics(
                guess[:self.state_dim, i],
                guess[self.state_dim:, i],
                self.dt

            ).toarray().flatten()
        return guess

    def initialise_mhtt(self, initial_state, current_progress):   # control/moving_horizon.py:203-213
        guess = np.zeros((self.state_dim + self.control_dim + 1, self.num_nodes + 1))
        guess[:self.state_dim, 0] = initial_state.toarray().flatten()

        # Forward propagate dynamics
        for i in range(self.num_nodes):
            guess[:self.state_dim, i + 1] = self.dynamics(
                guess[:self.state_dim, i],
                guess[self.state_dim:self.state_dim+self.control_dim, i],
                self.dt
            ).toarray().flatten()
        guess[-1, 0] = current_progress
        return guess


def simulate_dynamics_py(aircraft, state, control, dt=.1, tf=5):     # main/dynamics/dynamics.py:100-121
    dyn = aircraft.state_update
    state_list = np.zeros((aircraft.num_states, int(tf / dt)))
    times_list = np.zeros((int(tf / dt)))
    t = 0
    control_list = np.zeros((aircraft.num_controls, int(tf / dt)))
    for i in range(int(tf / dt)):
        if np.isnan(state[0]):
            print('Aircraft crashed')
            break
        else:
            if aircraft.phi(state).full().flatten() > np.deg2rad(70):
                control[0] = -1

            elif aircraft.phi(state).full().flatten() < np.deg2rad(-70):
                control[0] = 1

            state_list[:, i] = state.full().flatten()
            control_list[:, i] = control
            state = dyn(state, control, dt)
            times_list[i] = i * dt
            t += 1
    return state_list, control_list, times_list, t


# ------------------------------------------------------------------ oracle-backed stand-in (CPU)
class OracleSystem:
    """The float64 oracle behind the same Function / DM interface (test infrastructure)."""
    num_states, num_controls = 13, 7

    def __init__(self, orc, substeps=1, normalise=True):
        self.orc, self.substeps, self.normalise = orc, substeps, normalise
        self.state_update = Function("state_update", self._phi, ("x", "u", "dt"), (13, 7, 1), ("x_next",), (13,))
        self.state_derivative = Function("dynamics", self._f, ("x", "u"), (13, 7), ("x_dot",), (13,))
        self.phi = Function("phi", self._roll, ("x",), (13,), ("phi",), (1,))

    def _phi(self, x, u, dt):
        x, u = np.asarray(x, np.float64).reshape(13, -1), np.asarray(u, np.float64).reshape(7, -1)
        out = self.orc.state_update(t64(x.T), t64(u.T), float(dt), self.substeps, self.normalise).numpy().T
        return as_dm(out)

    def _f(self, x, u):
        x, u = np.asarray(x, np.float64).reshape(13, -1), np.asarray(u, np.float64).reshape(7, -1)
        return as_dm(self.orc.state_derivative(t64(x.T), t64(u.T)).numpy().T)

    def _roll(self, x):
        q = np.asarray(x, np.float64).reshape(13, -1)[6:10]
        qx, qy, qz, qw = q
        return as_dm(np.arctan2(2 * (qw * qx + qy * qz), 1 - 2 * (qx * qx + qy * qy)))


def _x0u0(seed=3):
    x = make_states(1, seed=seed, level=True)[0].astype(np.float64)
    u = make_controls(1, seed=seed)[0].astype(np.float64)
    u[3:6] = 0.0
    return x, u


def test_dm_answers_what_the_call_sites_ask():
    d = DM([1.0, 2.0, 3.0])
    assert d.shape == (3, 1) and d.full().shape == (3, 1) and type(d.full()) is np.ndarray
    assert np.array_equal(d.toarray().flatten(), [1.0, 2.0, 3.0]) and d.toarray(True).shape == (3,)
    assert d.size1() == 3 and d.size2() == 1 and d.nonzeros() == [1.0, 2.0, 3.0]
    assert float(DM(2.5)) == 2.5 and DM(2.5).is_scalar()
    assert not np.isnan(d[0]) and np.isnan(DM([np.nan, 1.0])[0])
    v = as_dm(np.arange(13.0))                      # 1-D results keep their shape; accessors still give columns
    assert v.shape == (13,) and v.full().shape == (13, 1) and np.array_equal(v.toarray().flatten(), np.arange(13.0))
    s = as_dm(np.float64(0.3).reshape(()))
    assert s.full().shape == (1, 1) and (s.full().flatten() > 0.1)
    m = DM(np.ones((13, 4)))
    assert m.T.shape == (4, 13) and isinstance(m.T, DM) and isinstance(m + 1, DM)
    with pytest.raises(TypeError):
        float(d)


def test_function_introspection_matches_casadi_usage():
    from aircraft_b200 import Aircraft
    ac = Aircraft.from_problem(com=TRIM_COM)          # no device work: only the interface
    f = ac.state_update
    assert f.size1_in(0) == 13 and f.size1_in(1) == 7 and f.size1_in(2) == 1 and f.size1_out(0) == 13
    assert f.n_in() == 3 and f.n_out() == 1 and f.name() == "state_update" and f.name_in() == ["x", "u", "dt"]
    assert ac.state_derivative.name() == "dynamics" and ac.state_derivative.name_out(0) == "x_dot"   # base.py:402-406
    assert ac.state_update is f                                         # cached
    assert ac.phi.n_in() == 1 and ac.v_frd_rel.size1_out(0) == 3 and ac.coefficients.size1_out(0) == 6
    assert ac.num_states == 13 and ac.num_controls == 7
    with pytest.raises(AttributeError):
        ac.no_such_function


def test_reference_loops_run_on_the_dm_interface():
    orc = oracle_for(TRIM_COM)
    sys_ = OracleSystem(orc)
    x0, u0 = _x0u0()
    ctl = _Controller(sys_, num_nodes=12, dt=0.01)
    g = ctl.initialise_control_py(DM(x0))
    assert g.shape == (20, 13) and np.array_equal(g[:13, 0], x0) and np.isfinite(g).all()
    U = np.zeros((12, 7)); U[:, 0] = 1.0              # guess[13,:] = 1, guess[14,:] = 0
    want = orc.rollout(t64(x0[None]), t64(U[None]), 0.01, 1, True)[0].numpy()
    assert np.allclose(g[:13].T, want, rtol=0, atol=1e-12)
    gm = ctl.initialise_mhtt(DM(x0), 0.25)
    assert gm.shape == (21, 13) and gm[-1, 0] == 0.25
    sl, cl, tl, t = simulate_dynamics_py(sys_, DM(x0), u0.copy(), dt=0.05, tf=1.0)
    assert t == 20 and np.array_equal(sl[:, 0], x0) and np.isfinite(sl).all()


# ------------------------------------------------------------------ the same loops on the GPU facade
@pytest.mark.gpu
def test_reference_loops_against_the_cuda_facade(built):
    from aircraft_b200 import Aircraft
    ac = Aircraft.from_problem(com=TRIM_COM, device="cuda:0", physical_integration_substeps=1)
    ac.normalise = True                                # ControlProblem.__init__, opts['quaternion'] == 'integration'
    orc = oracle_for(TRIM_COM)
    x0, u0 = _x0u0()
    ctl = _Controller(ac, num_nodes=50, dt=0.01)
    assert ctl.state_dim == 13 and ctl.control_dim == 7
    g = ctl.initialise_control_py(DM(x0))
    U = np.zeros((50, 7)); U[:, 0] = 1.0
    want = orc.rollout(t64(x0[None]), t64(U[None]), 0.01, 1, True)[0].numpy()
    assert col_err(g[:13].T, want) < 2e-5             # 50 free-running FP32 steps
    assert np.abs(g[:13, 1] - want[1]).max() <= 1e-5 * max(np.abs(want[1]).max(), 1.0)      # one step: the parity bar
    gm = ctl.initialise_mhtt(DM(x0), 0.1)
    want0 = orc.rollout(t64(x0[None]), t64(np.zeros((1, 50, 7))), 0.01, 1, True)[0].numpy()
    assert gm.shape == (21, 51) and col_err(gm[:13].T, want0) < 2e-5
    # the loop's 50 single-node calls give the bits of ONE rollout launch over the same controls
    ro = ac.rollout(x0[None], np.zeros((1, 50, 7)), 0.01, return_trajectory=True, normalise=True)
    assert col_err(gm[:13].T, np.asarray(ro.trajectory[0], np.float64)) < 2e-6
    # main/dynamics/dynamics.py: default 10 substeps, dt 0.1, no normalisation, DM state
    ac10 = Aircraft.from_problem(com=TRIM_COM, device="cuda:0")
    assert ac10.physical_integration_substeps == 10
    sl, cl, tl, t = simulate_dynamics_py(ac10, DM(x0), u0.copy(), dt=0.1, tf=2.0)
    assert t == 20
    orc10 = oracle_for(TRIM_COM)
    xs = x0.copy()
    for i in range(20):                                # teacher-forced against the loop's own states
        nxt = orc10.state_update(t64(sl[:, i][None]), t64(cl[:, i][None]), 0.1, 10, False)[0].numpy()
        if i + 1 < 20:
            assert np.abs(sl[:, i + 1] - nxt).max() <= 1e-5 * max(np.abs(nxt).max(), 1.0), i
    # device tensors in -> DM-compatible tensors out (no host copy until .full())
    xd = torch.tensor(x0, dtype=torch.float32, device="cuda:0")
    ud = torch.tensor(u0, dtype=torch.float32, device="cuda:0")
    out = ac.state_update(xd, ud, 0.01)
    assert out.is_cuda and out.full().shape == (13, 1) and out.toarray().flatten().shape == (13,)
    assert np.allclose(out.full().ravel(), np.asarray(ac.state_update(x0, u0, 0.01)), rtol=0, atol=0)
    assert float(ac.phi(DM(x0)).full().flatten()[0]) == pytest.approx(float(OracleSystem(orc)._roll(x0).ravel()[0]), abs=1e-6)

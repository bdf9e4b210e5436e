"""Seeded synthetic inputs for the BASELINE configs.  Generated on the CPU
in float64 (numpy), cast by the caller.  Nothing here touches the oracle.

State order  x = [p_ned(3), v_ned(3), q_frd_ned xyzw(4), omega_frd(3)]  (reference base.py:94-99)
Control order u = [aileron, elevator, rudder, thrust(3), flaps]         (reference aircraft.py:152-158)
"""
from __future__ import annotations

import numpy as np

SEED_BASE = 20250101


def _quat_from_euler(phi, theta, psi):
    """xyzw quaternion of the ZYX (yaw-pitch-roll) rotation body->NED."""
    cr, sr = np.cos(phi / 2), np.sin(phi / 2)
    cp, sp = np.cos(theta / 2), np.sin(theta / 2)
    cy, sy = np.cos(psi / 2), np.sin(psi / 2)
    w = cr * cp * cy + sr * sp * sy
    x = sr * cp * cy - cr * sp * sy
    y = cr * sp * cy + sr * cp * sy
    z = cr * cp * sy - sr * sp * cy
    return np.stack([x, y, z, w], axis=-1)


def _rotate(q, v):
    """q (x) (v,0) (x) q^-1 for unit xyzw q."""
    qv, qw = q[..., :3], q[..., 3:4]
    t = 2.0 * np.cross(qv, v)
    return v + qw * t + np.cross(qv, t)


def make_states(n: int, seed: int = 0, perturb_norm: float = 0.0, level: bool = False) -> np.ndarray:
    """n states [n,13]: V~U[30,70] m/s, alpha~U[-8,8] deg, beta~U[-5,5] deg, Euler angles
    phi~U[-30,30], theta~U[-15,15], psi~U[-180,180] deg, p = (0,0,-200)+N(0,10^2),
    omega~N(0,0.1^2).  perturb_norm>0 scales q by (1+N(0,perturb_norm^2)) so that the |q|^2 terms
    of the Jacobians are exercised.  level=True: phi=theta=beta=0 (trim seeds)."""
    g = np.random.default_rng(SEED_BASE + seed)
    d2r = np.pi / 180.0
    V = g.uniform(30.0, 70.0, n)
    alpha = g.uniform(-8.0, 8.0, n) * d2r
    beta = g.uniform(-5.0, 5.0, n) * d2r
    phi = g.uniform(-30.0, 30.0, n) * d2r
    theta = g.uniform(-15.0, 15.0, n) * d2r
    psi = g.uniform(-180.0, 180.0, n) * d2r
    if level:
        beta = np.zeros(n)
        phi = np.zeros(n)
        theta = alpha.copy()          # flight-path angle 0: pitch = alpha
    vb = np.stack([V * np.cos(alpha) * np.cos(beta), V * np.sin(beta),
                   V * np.sin(alpha) * np.cos(beta)], axis=-1)
    q = _quat_from_euler(phi, theta, psi)
    v = _rotate(q, vb)
    p = np.array([0.0, 0.0, -200.0]) + g.normal(0.0, 10.0, (n, 3))
    w = g.normal(0.0, 0.1, (n, 3))
    if level:
        w = np.zeros((n, 3))
    if perturb_norm > 0.0:
        q = q * (1.0 + g.normal(0.0, perturb_norm, (n, 1)))
    return np.concatenate([p, v, q, w], axis=-1)


def make_controls(n: int, seed: int = 0) -> np.ndarray:
    """n controls [n,7]: aileron/elevator/rudder ~U[-5,5] deg, thrust 0, flaps ~U[0,1]
    (limits: reference control/aircraft.py:24,29-41)."""
    g = np.random.default_rng(SEED_BASE + 1000 + seed)
    u = np.zeros((n, 7))
    u[:, 0:3] = g.uniform(-5.0, 5.0, (n, 3))
    u[:, 6] = g.uniform(0.0, 1.0, n)
    return u


def make_control_sequences(n: int, horizon: int, seed: int = 0, sigma: float = 0.25,
                           dtype=np.float64) -> np.ndarray:
    """[n,H,7] random walks u_{k+1} = clip(u_k + N(0,sigma^2)) from make_controls."""
    g = np.random.default_rng(SEED_BASE + 2000 + seed)
    u0 = make_controls(n, seed)
    lo = np.array([-5.0, -5.0, -5.0, 0.0, 0.0, 0.0, 0.0])
    hi = np.array([5.0, 5.0, 5.0, 0.0, 0.0, 0.0, 1.0])
    out = np.empty((n, horizon, 7), dtype)
    u = u0
    for k in range(horizon):
        out[:, k] = u
        step = g.normal(0.0, sigma, (n, 7))
        step[:, 3:6] = 0.0
        step[:, 6] *= 0.1
        u = np.clip(u + step, lo, hi)
    return out


def trim_states(aircraft, n: int, seed: int = 0, iters: int = 12):
    """"Random trim states" of BASELINE config 2: level flight at V~U[30,70], beta=phi=0; per
    sample the 2x2 problem (alpha, elevator): v_dot_z = 0, omega_dot_y = 0 is solved by Newton
    with the product's own batched state_derivative (secant Jacobian).  Returns (x[n,13],
    u[n,7]) float64 numpy.  `aircraft` is an aircraft_b200.Aircraft on a CUDA device."""
    g = np.random.default_rng(SEED_BASE + 3000 + seed)
    V = g.uniform(30.0, 70.0, n)
    psi = g.uniform(-np.pi, np.pi, n)
    alpha = np.full(n, 0.03)
    de = np.zeros(n)

    def build(alpha, de):
        vb = np.stack([V * np.cos(alpha), np.zeros(n), V * np.sin(alpha)], axis=-1)
        q = _quat_from_euler(np.zeros(n), alpha, psi)
        v = _rotate(q, vb)
        x = np.concatenate([np.tile([0.0, 0.0, -200.0], (n, 1)), v, q, np.zeros((n, 3))], axis=-1)
        u = np.zeros((n, 7))
        u[:, 1] = de
        return x, u

    def resid(alpha, de):
        x, u = build(alpha, de)
        xd = np.asarray(aircraft.state_derivative(x.T, u.T), np.float64).T
        return np.stack([xd[:, 5], xd[:, 11]], axis=-1)

    ha, he = 2e-3, 5e-2
    for _ in range(iters):
        r0 = resid(alpha, de)
        ra = (resid(alpha + ha, de) - r0) / ha
        re = (resid(alpha, de + he) - r0) / he
        det = ra[:, 0] * re[:, 1] - re[:, 0] * ra[:, 1]
        det = np.where(np.abs(det) < 1e-9, 1e-9, det)
        da = (r0[:, 0] * re[:, 1] - re[:, 0] * r0[:, 1]) / det
        dd = (ra[:, 0] * r0[:, 1] - r0[:, 0] * ra[:, 1]) / det
        alpha = np.clip(alpha - np.clip(da, -0.05, 0.05), -0.25, 0.25)
        de = np.clip(de - np.clip(dd, -2.0, 2.0), -5.0, 5.0)
    return build(alpha, de)

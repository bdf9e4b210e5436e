"""Generates tests/golden/step_golden.npz from the float64 oracle (oracle/aircraft_oracle.py,
itself pinned by test_oracle_golden.py): 64 seeded (x,u) nodes with f, Phi (two settings), A, B and
a 24-step rollout of 8 trajectories.  Committed so that the GPU parity tests have a fixed target
that does not depend on the torch version on the GPU box.

Usage:  python tests/golden/make_step_golden.py
"""
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
from aircraft_b200.config import TRIM_COM  # noqa: E402
from aircraft_b200.workloads import make_control_sequences, make_controls, make_states  # noqa: E402
from oracle.aircraft_oracle import AircraftOracle, OracleConfig  # noqa: E402


def main():
    orc = AircraftOracle(OracleConfig.from_problem(com=TRIM_COM))
    x = make_states(64, seed=101, perturb_norm=1e-3)
    u = make_controls(64, seed=101)
    tx, tu = torch.tensor(x), torch.tensor(u)
    f = orc.state_derivative(tx, tu).numpy()
    phi = orc.state_update(tx, tu, 0.01, 1, False).numpy()
    phi10n = orc.state_update(tx, tu, 0.05, 10, True).numpy()
    xn, A, B = orc.linearize(tx, tu, 0.01, 1, True)
    x0 = make_states(8, seed=102)
    U = make_control_sequences(8, 24, seed=102)
    X = orc.rollout(torch.tensor(x0), torch.tensor(U), 0.01, 1, True).numpy()
    cost = orc.cost(torch.tensor(X), torch.tensor(U)).numpy()
    np.savez_compressed(Path(__file__).resolve().parent / "step_golden.npz", x=x, u=u, f=f, phi=phi,
                        phi10n=phi10n, xn_norm=xn.numpy(), A_norm=A.numpy(), B_norm=B.numpy(),
                        x0=x0, U=U, X=X, cost=cost, com=np.array(TRIM_COM))
    print("written; |A| max", np.abs(A.numpy()).max())


if __name__ == "__main__":
    main()

"""Profiling target (for ncu): a few launches of K3 (airc_hessian) and K2 (airc_linearize_dt)."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from aircraft_b200 import Aircraft
from aircraft_b200.config import TRIM_COM
from aircraft_b200.workloads import make_controls, make_states
n = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
ac = Aircraft.from_problem(com=TRIM_COM, device=0, physical_integration_substeps=1, normalise=True)
x = torch.tensor(make_states(n, seed=1, perturb_norm=1e-3).T.copy(), dtype=torch.float32, device="cuda:0")
u = torch.tensor(make_controls(n, seed=1).T.copy(), dtype=torch.float32, device="cuda:0")
lam = torch.randn(13, n, device="cuda:0")
for _ in range(2):
    H, g, _ = ac.lagrangian_hessian(x, u, lam, 0.01, layout="soa")
    xn, A, B, Dt = ac.linearize(x, u, 0.01, layout="soa", want_dt=True)
torch.cuda.synchronize()
print("ok", float(H.abs().max()), float(A.abs().max()))

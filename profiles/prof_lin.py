"""Short profiling target for ncu: a few launches of the Jacobian kernel (airc_linearize) in one layout.
Usage: python profiles/prof_lin.py [nodes] [soa|aos] [dt]"""
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from aircraft_b200 import Aircraft  # noqa: E402
from aircraft_b200.config import TRIM_COM  # noqa: E402
from aircraft_b200.workloads import make_controls, make_states  # noqa: E402

NJ = int(sys.argv[1]) if len(sys.argv) > 1 else 655360
lay = sys.argv[2] if len(sys.argv) > 2 else "soa"
want_dt = len(sys.argv) > 3 and sys.argv[3] == "dt"
dev = torch.device("cuda", 0)
ac = Aircraft.from_problem(com=TRIM_COM, device=0, physical_integration_substeps=1, normalise=True)
x = torch.tensor(make_states(NJ, seed=300, perturb_norm=1e-3), dtype=torch.float32, device=dev)
u = torch.tensor(make_controls(NJ, seed=300), dtype=torch.float32, device=dev)
if lay == "soa":
    x, u = x.T.contiguous(), u.T.contiguous()
for _ in range(3):
    out = ac.linearize(x, u, 0.01, layout=lay, want_dt=want_dt)
torch.cuda.synchronize()
print("ok", float(out[1].abs().sum()))

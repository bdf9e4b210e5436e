"""Short profiling target for `ncu --set full`: the c4-shard rollout (fused cost) and a Jacobian
sweep, a few launches each, through the same C-ABI calls bench.py times.
Usage: python profiles/prof_target.py [rollouts] [jac_nodes]"""
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from aircraft_b200 import Aircraft  # noqa: E402
from aircraft_b200.config import TRIM_COM  # noqa: E402
from aircraft_b200.workloads import make_controls, make_states  # noqa: E402
from bench import device_control_sequences  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 125000
NJ = int(sys.argv[2]) if len(sys.argv) > 2 else 655360
dev = torch.device("cuda", 0)
ac = Aircraft.from_problem(com=TRIM_COM, device=0, physical_integration_substeps=1, normalise=True)
x0 = torch.tensor(make_states(1, seed=4, level=True).T, dtype=torch.float32, device=dev).expand(13, B).contiguous()
U = device_control_sequences(torch, B, 200, seed=100, device=dev)
x = torch.tensor(make_states(NJ, seed=300, perturb_norm=1e-3).T, dtype=torch.float32, device=dev).contiguous()
u = torch.tensor(make_controls(NJ, seed=300).T, dtype=torch.float32, device=dev).contiguous()
for _ in range(3):
    r = ac.rollout(x0, U, 0.01, layout="soa", return_final=False, cost=True)
    ac.argmin(r.cost)
    ac.linearize(x, u, 0.01, layout="soa")
torch.cuda.synchronize()
print("ok", float(r.cost.nan_to_num(1e30).min()))

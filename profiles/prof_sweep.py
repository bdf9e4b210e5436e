"""Profiling target: a few sampling sweeps of the config-4 shard (for ncu)."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from aircraft_b200 import Aircraft
from aircraft_b200.config import TRIM_COM
from aircraft_b200.params import make_sweep_params
from aircraft_b200.workloads import make_states
B = int(sys.argv[1]) if len(sys.argv) > 1 else 125000
ac = Aircraft.from_problem(com=TRIM_COM, device=0, physical_integration_substeps=1, normalise=True)
x0 = torch.tensor(make_states(1, seed=4, level=True)[0], dtype=torch.float32, device="cuda:0")
Un = torch.zeros((200, 7), dtype=torch.float32, device="cuda:0")
ws = None
for _ in range(3):
    r = ac.sweep(x0, Un, 0.01, B, sampler=make_sweep_params(seed=1), workspace=ws); ws = r.workspace
torch.cuda.synchronize()
print("ok", float(r.min_cost))

"""A/B of the small-batch rollout: one trajectory per thread (AIRC_ROLLOUT_COOP=0, the default) vs lanes per
trajectory (1 = auto by batch size as AIRC_FLAG_COOPERATIVE selects, 2/4/8 = forced)."""
import os
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
CHILD = r'''
import sys, torch
sys.path.insert(0, %r)
from aircraft_b200 import Aircraft
from aircraft_b200.config import TRIM_COM
from aircraft_b200.workloads import make_states
from bench import device_control_sequences, time_launches
dev = torch.device("cuda", 0)
ac = Aircraft.from_problem(com=TRIM_COM, device=0, physical_integration_substeps=1, normalise=True)
for B in (1, 1024, 2368, 4096, 4736, 8192, 9472, 16384):
    x0 = torch.tensor(make_states(B, seed=4).T.copy(), dtype=torch.float32, device=dev).contiguous()
    U = device_control_sequences(torch, B, 200, seed=100, device=dev)
    ms = time_launches(torch, lambda: ac.rollout(x0, U, 0.01, layout="soa", cost=True), 20, 3)
    r = ac.rollout(x0, U, 0.01, layout="soa", cost=True)
    print("  B=%%5d  %%.4f ms  %%.3e steps/s  checksum %%.8e" %% (B, ms, B*200/ms*1e3, float(r.cost.nan_to_num(0).double().sum())))
''' % str(ROOT)
for coop in ("0", "1", "2", "4", "8"):
    print("AIRC_ROLLOUT_COOP=" + coop, flush=True)
    r = subprocess.run([sys.executable, "-c", CHILD], env=dict(os.environ, AIRC_ROLLOUT_COOP=coop), capture_output=True, text=True)
    print(r.stdout.rstrip() or r.stderr[-400:], flush=True)

"""Latency cases: BASELINE config 2 (4096 x 200, explicit controls, fused cost) and config 1 (1 x 100, trajectory
returned), default dispatch vs AIRC_FLAG_COOPERATIVE, L2 flushed between launches.
Usage: python profiles/time_small.py [variant.so]   (a build variant from profiles/build_variants.py, optional)"""
import sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import torch
if len(sys.argv) > 1:
    from aircraft_b200 import _cabi
    _cabi._lib = _cabi.load_library(sys.argv[1])
from aircraft_b200 import Aircraft
from aircraft_b200.config import TRIM_COM
from aircraft_b200.workloads import make_states, trim_states
from bench import device_control_sequences, time_launches
dev = torch.device("cuda", 0)
ac = Aircraft.from_problem(com=TRIM_COM, device=0, physical_integration_substeps=1, normalise=True)
flush = torch.zeros(64 * 1024 * 1024, dtype=torch.float32, device=dev)
tag = Path(sys.argv[1]).stem if len(sys.argv) > 1 else "shipped"
for B in (4096, 1024, 16384):
    xs, _ = trim_states(ac, B, seed=0)
    x0 = torch.tensor(xs.T, dtype=torch.float32, device=dev).contiguous()
    U = device_control_sequences(torch, B, 200, seed=200, device=dev)
    a = time_launches(torch, lambda: ac.rollout(x0, U, 0.01, layout="soa", cost=True), 20, 3, flush)
    b = time_launches(torch, lambda: ac.rollout(x0, U, 0.01, layout="soa", cost=True, cooperative=True), 20, 3, flush)
    r = ac.rollout(x0, U, 0.01, layout="soa", cost=True, cooperative=True)
    print(f"[{tag}] c2-like B={B}: default {a:.4f} ms, cooperative {b:.4f} ms  ({B*200/min(a,b)/1e6:.3f}e9 steps/s)  checksum {float(r.cost.nan_to_num(0).double().sum()):.8e}", flush=True)
x1 = torch.tensor(make_states(1, seed=1, level=True), dtype=torch.float32, device=dev)
U1 = device_control_sequences(torch, 1, 100, seed=11, device=dev).permute(2, 0, 1).contiguous()
a = time_launches(torch, lambda: ac.rollout(x1, U1, 0.01, return_trajectory=True), 50, 3, None)
b = time_launches(torch, lambda: ac.rollout(x1, U1, 0.01, return_trajectory=True, cooperative=True), 50, 3, None)
print(f"[{tag}] c1: default {a*10:.3f} us/step, cooperative {b*10:.3f} us/step", flush=True)

"""A/B of the rollout kernels (default dispatch vs AIRC_FLAG_PACKED: `python profiles/ab_rollout.py [packed]`): timing on
the config-4 shard and a quick parity number against the float64 CPU reference."""
import os, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np, torch
from aircraft_b200 import Aircraft
from aircraft_b200.config import TRIM_COM, AircraftConfiguration
from aircraft_b200.params import make_sweep_params, make_params, load_nn_blob, make_cost_params
from aircraft_b200.workloads import make_states, make_control_sequences
from oracle.cpu_ref import CpuRef

dev = torch.device("cuda:0")
ac = Aircraft.from_problem(com=TRIM_COM, device=0, physical_integration_substeps=1, normalise=True)
ref = CpuRef(make_params(AircraftConfiguration.from_json(), com=TRIM_COM), load_nn_blob())
PACKED = len(sys.argv) > 1 and sys.argv[1] == "packed"
tag = "packed" if PACKED else "default"
# parity: 2048 random states, 30 teacher-forced... free-running steps vs float64
n, H = 2049, 30
x0 = make_states(n, seed=77); U = make_control_sequences(n, H, seed=77)
X = ac.rollout(torch.tensor(x0, dtype=torch.float32, device=dev), torch.tensor(U, dtype=torch.float32, device=dev),
               0.01, return_trajectory=True, cost=True, packed=PACKED)
Xr = ref.rollout(x0, U, 0.01, normalise=True)
cr = ref.rollout_cost(x0, U, 0.01, make_cost_params(), normalise=True)
got = X.trajectory.double().cpu().numpy()
err1 = (np.abs(got[:, 1] - Xr[:, 1]).max(0) / np.maximum(np.abs(Xr[:, 1]).max(0), 1e-3)).max()
errH = (np.abs(got - Xr).max((0, 1)) / np.maximum(np.abs(Xr).max((0, 1)), 1e-3)).max()
cerr = np.abs(X.cost.double().cpu().numpy() - cr).max() / np.abs(cr).max()
print(f"[{tag}] parity: 1 step {err1:.2e}, {H} free steps {errH:.2e}, cost {cerr:.2e}", flush=True)

def timeit(fn, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n

H = 200
for B in (125000, 151552, 75776, 4096):
    Un = torch.zeros((H, 7), dtype=torch.float32, device=dev)
    U = ac.sweep_controls(Un, make_sweep_params(seed=1), 0, B, layout="soa")
    xs = torch.tensor(make_states(1, seed=4, level=True)[0], dtype=torch.float32, device=dev).view(13, 1).expand(13, B).contiguous()
    ms = timeit(lambda: ac.rollout(xs, U, 0.01, layout="soa", cost=True, return_final=False, packed=PACKED))
    print(f"[{tag}] rollout B={B}: {ms:.4f} ms  {B*H/ms/1e6:.3f}e9 steps/s  {B*H*3953/ms/1e9:.2f} TFLOP/s", flush=True)

"""Device timing of the sampling sweep vs the explicit-control rollout (CUDA events)."""
import sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np, torch
from aircraft_b200 import Aircraft
from aircraft_b200.config import TRIM_COM
from aircraft_b200.params import make_sweep_params
from aircraft_b200.workloads import make_states

dev = torch.device("cuda:0")
ac = Aircraft.from_problem(com=TRIM_COM, device=0, physical_integration_substeps=1, normalise=True)
B, H = int(sys.argv[1]) if len(sys.argv) > 1 else 125000, 200
x0 = torch.tensor(make_states(1, seed=4, level=True)[0], dtype=torch.float32, device=dev)
Un = torch.zeros((H, 7), dtype=torch.float32, device=dev)

def timeit(fn, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n

for ck in (8, 4, 16, 25, 200):
    sp = make_sweep_params(seed=1, ckpt_interval=ck)
    ws = None
    r = ac.sweep(x0, Un, 0.01, B, sampler=sp); ws = r.workspace
    ms = timeit(lambda: ac.sweep(x0, Un, 0.01, B, sampler=sp, workspace=ws))
    print(f"sweep ckpt={ck:3d} B={B}: {ms:.4f} ms  {B*H/ms/1e6:.3f}e9 steps/s", flush=True)
U = ac.sweep_controls(Un, make_sweep_params(seed=1), 0, B, layout="soa")
xs = x0.view(13, 1).expand(13, B).contiguous()
ms = timeit(lambda: ac.rollout(xs, U, 0.01, layout="soa", cost=True, return_final=False))
print(f"rollout explicit U: {ms:.4f} ms  {B*H/ms/1e6:.3f}e9 steps/s")
xh, Uh = x0.cpu().numpy(), Un.cpu().numpy()
sp = make_sweep_params(seed=1)
for _ in range(3): ac.sweep(xh, Uh, 0.01, B, sampler=sp)
t = time.perf_counter()
for _ in range(10): ac.sweep(xh, Uh, 0.01, B, sampler=sp)
el = (time.perf_counter() - t) / 10
print(f"sweep_host e2e: {el*1e3:.4f} ms  {B*H/el/1e9:.3f}e9 steps/s")

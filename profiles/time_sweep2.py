"""Times airc_sweep for the given batch sizes: python profiles/time_sweep2.py 125000,1000000
(segment-count experiments are build variants: -DAIRC_AB_SWEEP_SEGMENTS=n, profiles/build_variants.py)."""
import sys, os
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from aircraft_b200 import Aircraft
from aircraft_b200.config import TRIM_COM
from aircraft_b200.params import make_sweep_params
from aircraft_b200.workloads import make_states
ac = Aircraft.from_problem(com=TRIM_COM, device=0, physical_integration_substeps=1, normalise=True)
x0 = torch.tensor(make_states(1, seed=4, level=True)[0], dtype=torch.float32, device="cuda:0")
Un = torch.zeros((200, 7), dtype=torch.float32, device="cuda:0")
def timeit(fn, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
for B in [int(b) for b in sys.argv[1].split(",")]:
    sp = make_sweep_params(seed=1)
    r = ac.sweep(x0, Un, 0.01, B, sampler=sp); ws = r.workspace
    ms = timeit(lambda: ac.sweep(x0, Un, 0.01, B, sampler=sp, workspace=ws))
    print(f"B={B}: {ms:.4f} ms {B*200/ms/1e6:.3f}e9 steps/s", flush=True)

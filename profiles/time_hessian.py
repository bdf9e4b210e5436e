"""Times K3 (airc_hessian) and K2 (airc_linearize_dt) on device-resident SoA inputs with CUDA events."""
import sys
import numpy as np
import torch
sys.path.insert(0, ".")
from aircraft_b200 import Aircraft
from aircraft_b200.config import TRIM_COM
from aircraft_b200.workloads import make_controls, make_states

ac = Aircraft.from_problem(com=TRIM_COM, device="cuda:0", physical_integration_substeps=1)
for n in (298, 65536, 1 << 20):
    x = torch.tensor(make_states(n, seed=1, perturb_norm=1e-3).T.copy(), device="cuda")
    u = torch.tensor(make_controls(n, seed=1).T.copy(), device="cuda")
    lam = torch.randn(13, n, device="cuda")
    for name, fn in (("hessian", lambda: ac.lagrangian_hessian(x, u, lam, 0.01, layout="soa", normalise=True)),
                     ("linearize_dt", lambda: ac.linearize(x, u, 0.01, layout="soa", normalise=True, want_dt=True)),
                     ("linearize", lambda: ac.linearize(x, u, 0.01, layout="soa", normalise=True))):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        reps = 10
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        print(f"{name:14s} n={n:8d}  {ms:9.4f} ms  {n / ms * 1e3:.3e} nodes/s", flush=True)

"""A/B timing of library build variants (build/variants/*.so) on the c4-shard rollout.
Each variant runs in its own process and loads its .so explicitly (profiles/build_variants.py builds them).  Usage on the GPU box:
  python profiles/ab_variants.py [names...]"""
import os
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
CHILD = r'''
import sys, torch
sys.path.insert(0, %r)
import os
from aircraft_b200 import _cabi
_cabi._lib = _cabi.load_library(os.environ["AB_LIB"])      # the variant under test, loaded explicitly
from aircraft_b200 import Aircraft
from aircraft_b200.config import TRIM_COM
from aircraft_b200.workloads import make_states
from bench import device_control_sequences, time_launches
B = int(sys.argv[1])
dev = torch.device("cuda", 0)
ac = Aircraft.from_problem(com=TRIM_COM, device=0, physical_integration_substeps=1, normalise=True)
if len(sys.argv) > 2 and sys.argv[2] == "lin":
    from aircraft_b200.workloads import make_controls
    n, nu = B, min(B, 655360)
    x = torch.tensor(make_states(nu, seed=300, perturb_norm=1e-3).T, dtype=torch.float32, device=dev)
    u = torch.tensor(make_controls(nu, seed=300).T, dtype=torch.float32, device=dev)
    x, u = x.repeat(1, n // nu).contiguous(), u.repeat(1, n // nu).contiguous()
    n = x.shape[1]
    ms = time_launches(torch, lambda: ac.linearize(x, u, 0.01, layout="soa"), 10, 3)
    xn, A, Bm = ac.linearize(x, u, 0.01, layout="soa")
    cs = float(A.double().abs().sum())
    del xn, A, Bm
    xa, ua = x.T.contiguous(), u.T.contiguous()
    ms2 = time_launches(torch, lambda: ac.linearize(xa, ua, 0.01), 10, 3)
    ms3 = time_launches(torch, lambda: ac.linearize(xa, ua, 0.01, want_dt=True), 10, 3)
    print("lin n=%%d soa %%.4f ms %%.3e evals/s %%.0f GB/s | aos %%.4f ms %%.3e evals/s | aos+dt %%.4f ms | checksum %%.6e" %% (n, ms, n/ms*1e3, n*1172/ms*1e3/1e9, ms2, n/ms2*1e3, ms3, cs))
    sys.exit(0)
if len(sys.argv) > 2 and sys.argv[2] == "sweep":
    from aircraft_b200.params import make_sweep_params
    xs = torch.tensor(make_states(1, seed=4, level=True)[0], dtype=torch.float32, device=dev)
    Un = torch.zeros((200, 7), dtype=torch.float32, device=dev)
    sp = make_sweep_params(seed=1)
    r = ac.sweep(xs, Un, 0.01, B, sampler=sp); ws = r.workspace
    best = 1e9
    for rep in range(3):
        ms = time_launches(torch, lambda: ac.sweep(xs, Un, 0.01, B, sampler=sp, workspace=ws), 20, 3)
        best = min(best, ms)
    print("sweep %%.4f ms (best of 3 x 20)  %%.3e steps/s  %%.2f TFLOP/s  min cost %%.6e" %% (best, B*200/best*1e3, B*200*3953/best*1e3/1e12, float(r.min_cost)))
    sys.exit(0)
x0 = torch.tensor(make_states(1, seed=4, level=True).T, dtype=torch.float32, device=dev).expand(13, B).contiguous()
U = device_control_sequences(torch, B, 200, seed=100, device=dev)
ms = time_launches(torch, lambda: ac.rollout(x0, U, 0.01, layout="soa", return_final=False, cost=True), 10, 3)
r = ac.rollout(x0, U, 0.01, layout="soa", return_final=False, cost=True)
print("%%.4f ms  %%.3e steps/s  %%.2f TFLOP/s  checksum %%.6e" %% (ms, B*200/ms*1e3, B*200*3953/ms*1e3/1e12, float(r.cost.nan_to_num(0).double().sum())))
''' % str(ROOT)

names = sys.argv[1:] or sorted(p.stem for p in (ROOT / "build" / "variants").glob("*.so"))
for B in [int(b) for b in os.environ.get('AB_SIZES', '125000').split(',')]:
    for n in names:
        env = dict(os.environ, AB_LIB=str(ROOT / "build" / "variants" / f"{n}.so"))
        r = subprocess.run([sys.executable, "-c", CHILD, str(B)] + ([os.environ["AB_MODE"]] if os.environ.get("AB_MODE") else []), env=env, capture_output=True, text=True)
        print(f"{n:28s} B={B}: {r.stdout.strip() or r.stderr.strip()[-300:]}", flush=True)

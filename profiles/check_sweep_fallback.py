"""The bounded-wait fallback of the time-sliced sweep (airc_sweep2.cuh): a build variant with a zero spin limit
(profiles/build_variants.py fallback=-DAIRC_S2_SPIN_LIMIT=0u) makes EVERY task whose hand-over is not already there
roll its group forward from x0 itself.  Its record and costs must be bit-identical to the shipped library's.
Usage on the GPU box: python profiles/check_sweep_fallback.py"""
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
CHILD = r'''
import sys, torch
sys.path.insert(0, %r)
from aircraft_b200 import _cabi
if sys.argv[1] != "shipped":
    _cabi._lib = _cabi.load_library(sys.argv[1])
from aircraft_b200 import Aircraft
from aircraft_b200.config import TRIM_COM
from aircraft_b200.params import make_sweep_params
from aircraft_b200.workloads import make_states
ac = Aircraft.from_problem(com=TRIM_COM, device=0, physical_integration_substeps=1, normalise=True)
x0 = torch.tensor(make_states(1, seed=4, level=True)[0], dtype=torch.float32, device="cuda:0")
Un = torch.zeros((200, 7), dtype=torch.float32, device="cuda:0")
r = ac.sweep(x0, Un, 0.01, 125000, sampler=make_sweep_params(seed=1), return_costs=True)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); ac.sweep(x0, Un, 0.01, 125000, sampler=make_sweep_params(seed=1), workspace=r.workspace); e1.record()
torch.cuda.synchronize()
torch.save((r.raw.cpu(), r.cost.cpu()), sys.argv[2])
print("%%s: %%.3f ms, min cost %%.6e" %% (sys.argv[1].split("/")[-1], e0.elapsed_time(e1), float(r.min_cost)))
''' % str(ROOT)
outs = []
for lib in ("shipped", str(ROOT / "build" / "variants" / "fallback.so")):
    f = "/tmp/sweep_%s.pt" % Path(lib).stem
    r = subprocess.run([sys.executable, "-c", CHILD, lib, f], capture_output=True, text=True)
    print(r.stdout.strip() or r.stderr[-500:], flush=True)
    outs.append(f)
import torch
a, b = torch.load(outs[0]), torch.load(outs[1])
print("record bit-identical:", torch.equal(a[0], b[0]), " costs bit-identical:", torch.equal(a[1], b[1]))
assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])

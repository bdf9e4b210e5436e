"""K4 (airc_hessian, network model) per C-ABI call on prepared device buffers over node counts, AoS and SoA,
CUDA events around every call, L2 flushed between calls.  Usage: python profiles/time_hessian_sizes.py [variant.so]
(sizes: HESS_SIZES=298,4096,...; HESS_NOFLUSH=1 keeps L2 warm between calls; HESS_MODEL=poly|linear|default: a table model)."""
import os
import sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import torch
if len(sys.argv) > 1:
    from aircraft_b200 import _cabi
    _cabi._lib = _cabi.load_library(sys.argv[1])
from aircraft_b200 import Aircraft, _cabi
from aircraft_b200.config import TRIM_COM
from aircraft_b200.workloads import make_controls, make_states
from bench import time_launches
dev = torch.device("cuda", 0)
torch.manual_seed(0)
tag = Path(sys.argv[1]).stem if len(sys.argv) > 1 else "shipped"
ac = Aircraft.from_problem(com=TRIM_COM, device=0, physical_integration_substeps=1, normalise=True,
                           coeff_model_type=os.environ.get("HESS_MODEL", "nn"))
lib, h = ac._lib, ac.handle
st = torch.cuda.current_stream().cuda_stream
flush = None if os.environ.get("HESS_NOFLUSH") else torch.zeros(64 * 1024 * 1024, dtype=torch.float32, device=dev)
sizes = [int(s) for s in os.environ.get("HESS_SIZES", "64,149,298,600,1200,2400,4800,9600,65536,262144").split(",")]
for n in sizes:
    x = torch.tensor(make_states(n, seed=5, perturb_norm=1e-3), dtype=torch.float32, device=dev)
    u = torch.tensor(make_controls(n, seed=5), dtype=torch.float32, device=dev)
    lam = torch.randn(n, 13, device=dev)
    out = []
    for layout, (xx, uu, ll) in ((_cabi.LAYOUT_AOS, (x, u, lam)),
                                 (_cabi.LAYOUT_SOA, (x.T.contiguous(), u.T.contiguous(), lam.T.contiguous()))):
        H, g, v = torch.empty(n * 441, device=dev), torch.empty(n * 21, device=dev), torch.empty(n, device=dev)
        f = lambda: lib.airc_hessian(h, xx.data_ptr(), uu.data_ptr(), ll.data_ptr(), 0.01, None, n, 1, 1, layout,
                                     H.data_ptr(), g.data_ptr(), v.data_ptr(), st)
        assert f() == 0
        reps = 30 if n <= 9600 else 10
        ms = time_launches(torch, f, reps, 3, flush)
        out.append((ms, float(H.double().abs().sum())))
    print(f"[{tag}] n={n:7d}  aos {out[0][0]*1e3:9.1f} us {n/out[0][0]*1e3:.3e} nodes/s | soa {out[1][0]*1e3:9.1f} us "
          f"{n/out[1][0]*1e3:.3e} nodes/s | checksum {out[0][1]:.6e} {out[1][1]:.6e}", flush=True)

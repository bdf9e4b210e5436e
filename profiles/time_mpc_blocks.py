"""The dynamics blocks behind IPOPT's callbacks at 298 nodes with the DEFAULT coefficient model (bench.py
extra.mpc_blocks_default_298), timed per C-ABI call without the facade's tensor preparation.
Usage: python profiles/time_mpc_blocks.py [variant.so]"""
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
tag = Path(sys.argv[1]).stem if len(sys.argv) > 1 else "shipped"
for model in ("default", "nn"):
    ac = Aircraft.from_problem(com=TRIM_COM, device=0, coeff_model_type=model, physical_integration_substeps=1, normalise=True)
    n = 298
    x = torch.tensor(make_states(n, seed=5), dtype=torch.float32, device=dev)
    u = torch.tensor(make_controls(n, seed=5), dtype=torch.float32, device=dev)
    xt = x.roll(-1, 0).contiguous()
    r, A, B = torch.empty_like(x), torch.empty(n, 13, 13, device=dev), torch.empty(n, 13, 7, device=dev)
    lib, h = ac._lib, ac.handle
    st = torch.cuda.current_stream().cuda_stream
    flush = torch.zeros(64 * 1024 * 1024, dtype=torch.float32, device=dev)
    f = lambda: lib.airc_defects(h, x.data_ptr(), u.data_ptr(), xt.data_ptr(), 0.01, None, n, 1, 1, _cabi.LAYOUT_AOS,
                                 r.data_ptr(), A.data_ptr(), B.data_ptr(), None, st)
    ms = time_launches(torch, f, 30, 3, flush)
    g = lambda: ac.defects(torch.stack([x, xt], 1)[:, :, :].reshape(n, 2, 13), u[:, None, :], 0.01)
    ms2 = time_launches(torch, g, 30, 3, flush)
    print(f"[{tag}] {model}: airc_defects 298 nodes {ms*1e3:.1f} us per call (C ABI), {ms2*1e3:.1f} us through the facade", flush=True)

"""Builds library variants for A/B timing into build/variants/<name>.so (git-ignored, travels with gpurun).
  python profiles/build_variants.py name=-DFLAG[,-DFLAG2] ...
The shipped library reads no environment knobs; a variant is a different .so that profiles/ab_variants.py loads
explicitly."""
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
OUT = ROOT / "build" / "variants"
OUT.mkdir(parents=True, exist_ok=True)


def build(spec):
    name, _, flags = spec.partition("=")
    cmd = ["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17", "-Xcompiler", "-fPIC",
           "-shared", "-I", str(ROOT / "include")] + [f for f in flags.split(",") if f] + \
          ["-o", str(OUT / f"{name}.so"), str(ROOT / "aircraft_b200" / "csrc" / "airc_api.cu")]
    r = subprocess.run(cmd, capture_output=True, text=True)
    return name, r.returncode, r.stderr[-400:]


with ThreadPoolExecutor(4) as ex:
    for name, rc, err in ex.map(build, sys.argv[1:]):
        print(name, "ok" if rc == 0 else "FAILED\n" + err, flush=True)

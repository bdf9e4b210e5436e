"""Audited per-launch counts of one kernel from an ncu capture: EXECUTED FP32 flops (from the SASS page: predicated-on
thread instructions per opcode; FFMA = 2 flop, FFMA2 = 4, FMUL/FADD = 1, FMUL2/FADD2 = 2, MUFU = 1) and DRAM bytes
(raw page), written as one entry of profiles/r2_kernel_audit.json -- the file bench.py reads for `roofline.traffic`
and the executed-flop figure, together with a hash of the CUDA sources the capture was taken on (bench.py flags the
entry as stale when the sources have changed since).
Usage: python profiles/ncu_audit.py <name> <X.ncu-rep> <units per launch> [launch index]"""
import csv
import hashlib
import io
import json
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
FLOPS = {"FFMA": 2, "FFMA2": 4, "FMUL": 1, "FADD": 1, "FMUL2": 2, "FADD2": 2, "MUFU": 1, "FMNMX": 0, "FSEL": 0}


def csrc_sha():
    h = hashlib.sha256()
    for p in sorted((ROOT / "aircraft_b200" / "csrc").glob("*")):
        h.update(p.name.encode())
        h.update(p.read_bytes())
    return h.hexdigest()[:16]


def page(rep, which):
    out = subprocess.run(["ncu", "-i", str(rep), "--page", which, "--csv"], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def main():
    name, rep, units = sys.argv[1], Path(sys.argv[2]), float(sys.argv[3])
    raw = page(rep, "raw")
    h, u, d = raw[0], raw[1], raw[2 + (int(sys.argv[4]) if len(sys.argv) > 4 else 0)]
    unit = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "us": 1e-6, "ms": 1e-3, "ns": 1e-9, "s": 1.0}

    def val(metric):
        i = h.index(metric)
        return float(d[i].replace(",", "")) * unit.get(u[i], 1.0)

    src = page(rep, "source")
    hi = next(i for i, r in enumerate(src) if r and r[0] == "Address")
    hh = src[hi]
    si, ti = hh.index("Source"), hh.index("Predicated-On Thread Instructions Executed")
    wi = hh.index("Instructions Executed")
    per_op, flops, warp_inst = {}, 0.0, 0.0
    for r in src[hi + 1:]:
        if len(r) <= ti:
            continue
        t = r[si].split()
        if not t:
            continue
        op = (t[1] if t[0].startswith("@") else t[0]).split(".")[0]
        n = float(r[ti] or 0)
        warp_inst += float(r[wi] or 0)
        per_op[op] = per_op.get(op, 0.0) + n
        flops += FLOPS.get(op, 0) * n
    entry = {
        "kernel": d[h.index("Kernel Name")],
        "capture": str(rep.name), "csrc_sha": csrc_sha(), "units_per_launch": units,
        "duration_us_under_ncu": val("gpu__time_duration.sum") * 1e6,
        "dram_bytes_per_launch": val("dram__bytes_read.sum") + val("dram__bytes_write.sum"),
        "dram_bytes_per_unit": (val("dram__bytes_read.sum") + val("dram__bytes_write.sum")) / units,
        "executed_fp32_flops_per_unit": flops / units,
        "warp_instructions_per_unit": warp_inst / units,
        "thread_instructions_per_unit": {k: round(v / units, 2) for k, v in sorted(per_op.items(), key=lambda kv: -kv[1])[:14]},
        "registers": int(val("launch__registers_per_thread")),
        "issue_active_pct": val("smsp__issue_active.avg.pct_of_peak_sustained_active"),
        "fma_pipe_pct": val("sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active"),
    }
    out = ROOT / "profiles" / "r2_kernel_audit.json"
    allv = json.loads(out.read_text()) if out.exists() else {}
    allv[name] = entry
    out.write_text(json.dumps(allv, indent=1) + "\n")
    print(json.dumps(entry, indent=1))


if __name__ == "__main__":
    main()

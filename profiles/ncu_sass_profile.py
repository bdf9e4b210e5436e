"""Per-region summary of an `ncu --page source --csv` export (SASS view): instruction counts, opcode mix and
stall samples, split at the kernel's barriers (BAR.SYNC) so that phases can be read separately.
Usage: ncu -i X.ncu-rep --page source --csv > X_src.csv ; python profiles/ncu_sass_profile.py X_src.csv"""
import csv
import sys
from collections import Counter

rows = list(csv.reader(open(sys.argv[1])))
hdr_i = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
h = rows[hdr_i]
data = rows[hdr_i + 1:]
ci = {n: h.index(n) for n in ("Source", "# Samples", "Instructions Executed", "Thread Instructions Executed")}
stall_cols = [(i, n) for i, n in enumerate(h) if n.startswith("stall_") and "Not Issued" not in n]
regions, cur = [], {"n": 0, "exec": 0, "samples": 0, "ops": Counter(), "stalls": Counter(), "static": Counter()}
for r in data:
    if len(r) < len(h):
        continue
    src = r[ci["Source"]].strip()
    op = src.split()[0] if src else "?"
    if op.startswith("@"):
        op = src.split()[1]
    ex = int(r[ci["Instructions Executed"]] or 0)
    cur["n"] += 1
    cur["exec"] += ex
    cur["samples"] += int(r[ci["# Samples"]] or 0)
    cur["ops"][op.split(".")[0] + ("." + op.split(".")[1] if op.startswith(("LDS", "STS", "LDG", "STG", "MUFU", "LDC")) and "." in op else "")] += ex
    cur["static"][op.split(".")[0]] += 1
    for i, n in stall_cols:
        cur["stalls"][n] += int(r[i] or 0)
    if op.startswith("BAR"):
        regions.append(cur)
        cur = {"n": 0, "exec": 0, "samples": 0, "ops": Counter(), "stalls": Counter(), "static": Counter()}
regions.append(cur)
tot = sum(g["exec"] for g in regions)
tots = sum(g["samples"] for g in regions)
for k, g in enumerate(regions):
    print(f"--- region {k}: {g['n']} SASS instrs, executed {g['exec']} ({100 * g['exec'] / max(tot, 1):.1f} %), "
          f"samples {g['samples']} ({100 * g['samples'] / max(tots, 1):.1f} %)")
    print("    executed mix:", ", ".join(f"{o} {100 * c / max(g['exec'], 1):.1f}%" for o, c in g["ops"].most_common(14)))
    print("    stalls:", ", ".join(f"{n[6:]} {100 * c / max(g['samples'], 1):.0f}%" for n, c in g["stalls"].most_common(6)))
allops = Counter()
for g in regions:
    allops.update(g["ops"])
print("=== whole kernel executed mix:", ", ".join(f"{o} {100 * c / max(tot, 1):.1f}%" for o, c in allops.most_common(24)))

"""Executed warp instructions and stall samples per SOURCE line of one kernel: joins `nvdisasm -gi -c` (inline
chains of every SASS address) with `ncu --page source --csv` (executed counts per address).
Usage: cuobjdump -xelf all lib.so; nvdisasm -gi -c X.cubin > dis.txt
       ncu -i X.ncu-rep --page source --csv > src.csv
       python profiles/ncu_line_profile.py dis.txt <mangled kernel name> src.csv file1.cuh[:from_line][,file2.cuh] [top]
An instruction is charged to the innermost frame of its inline chain that lies in one of the named files (at or
after from_line, so that small arithmetic helpers at the top of a file are charged to their callers)."""
import csv
import re
import sys
from collections import Counter

dis, kern, src = sys.argv[1], sys.argv[2], sys.argv[3]
files = {f.partition(":")[0]: int(f.partition(":")[2] or 0) for f in sys.argv[4].split(",")}
top = int(sys.argv[5]) if len(sys.argv) > 5 else 60
lines = open(dis).read().split("\n")
start = next(i for i, l in enumerate(lines) if l.startswith(".text." + kern + ":"))
chain, where = [], {}
pend = []
for l in lines[start + 1:]:
    if l.startswith(".text.") or l.startswith("//-----"):
        break
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        pend.append((m.group(1).rsplit("/", 1)[-1], int(m.group(2))))
        continue
    m = re.match(r"\s+/\*([0-9a-f]+)\*/", l)
    if m:
        if pend:
            chain, pend = pend, []
        where[int(m.group(1), 16)] = chain
rows = list(csv.reader(open(src)))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
h = rows[hi]
ca, ce, cs = h.index("Address"), h.index("Instructions Executed"), h.index("# Samples")
base = None
ex, sm = Counter(), Counter()
for r in rows[hi + 1:]:
    if len(r) < len(h):
        continue
    a = int(r[ca], 16)
    if base is None:
        base = a
    ch = where.get(a - base, [])
    key = next(((f, n) for f, n in ch if f in files and n >= files[f]), ("?", 0))
    ex[key] += int(r[ce] or 0)
    sm[key] += int(r[cs] or 0)
tot, tots = sum(ex.values()), sum(sm.values())
print(f"kernel {kern}: {tot} warp instructions executed, {tots} samples")
text = {}
for f in files:
    for d in ("aircraft_b200/csrc/",):
        try:
            text[f] = open(d + f).read().split("\n")
        except OSError:
            pass
for (f, n), c in ex.most_common(top):
    t = text.get(f, [""] * (n + 1))[n - 1].strip()[:90] if n else ""
    print(f"{100 * c / tot:5.1f}% instr {100 * sm[(f, n)] / max(tots, 1):5.1f}% samples  {f}:{n:<4d} {t}")

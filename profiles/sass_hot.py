"""Top SASS lines by warp-stall samples from `ncu --page source --csv`. Usage: sass_hot.py SRC.csv [N]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
h = rows[1] if 'Source' in rows[1] else rows[0]
start = rows.index(h) + 1
si, ei, st = h.index('Source'), h.index('Instructions Executed'), h.index('Warp Stall Sampling (All Samples)')
data = []
for idx, r in enumerate(rows[start:]):
    if len(r) <= st: continue
    try: data.append((int(r[st] or 0), int(r[ei] or 0), idx, r[si]))
    except ValueError: pass
tot = sum(d[0] for d in data)
print("total samples", tot, "instructions", len(data))
for s, e, i, src in sorted(data, reverse=True)[:int(sys.argv[2]) if len(sys.argv) > 2 else 25]:
    print(f"{100*s/max(tot,1):5.1f}%  exec {e:>10d}  #{i:5d}  {src[:110]}")

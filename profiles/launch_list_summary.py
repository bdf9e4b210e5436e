"""Summarises an `ncu --metrics gpu__time_duration.sum --csv --log-file X.csv python bench.py ...` launch list per kernel.
Usage: python profiles/launch_list_summary.py X.csv "<header line>" > profiles/<round>_launch_list_bench.txt"""
import collections
import csv
import re
import sys

rows = list(csv.reader(open(sys.argv[1])))
hi = [i for i, r in enumerate(rows) if 'Kernel Name' in r][0]
h, data = rows[hi], rows[hi + 1:]
kn, mv, mu = h.index('Kernel Name'), h.index('Metric Value'), h.index('Metric Unit')
unit = {'ns': 1e-6, 'us': 1e-3, 'ms': 1.0, 's': 1e3, 'usecond': 1e-3, 'nsecond': 1e-6, 'msecond': 1.0}
agg = collections.OrderedDict()
for r in data:
    if len(r) <= mv:
        continue
    name = re.sub(r'\(.*', '', r[kn])
    t = float(r[mv].replace(',', '')) * unit.get(r[mu], 1e-6)
    a = agg.setdefault(name, [0, 0.0])
    a[0] += 1
    a[1] += t
tot = sum(v[1] for v in agg.values())
print(sys.argv[2] if len(sys.argv) > 2 else "ncu launch list")
print("(per-launch times under ncu are cold-cache and serialised: shares, not absolutes)\n")
print(f"{'kernel':70s} {'launches':>8s} {'total ms':>10s} {'share':>7s} {'ms/launch':>10s}")
for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    print(f"{k[:70]:70s} {n:8d} {t:10.3f} {100 * t / tot:6.1f}% {t / n:10.4f}")


def T(pat):
    return (sum(t for k, (n, t) in agg.items() if re.search(pat, k)), sum(n for k, (n, t) in agg.items() if re.search(pat, k)))


s, ns = T('sweep2tc_kernel')
p, npp = T('sweep2_prep')
w, nw = T('sweep_winner2')
if ns and npp and nw:
    print(f"\none sweep step = sweep2_prep ({p / npp * 1e3:.1f} us) + sweep2tc_kernel ({s / ns:.3f} ms, mean over the sizes "
          f"launched) + sweep_winner2 ({w / nw * 1e3:.1f} us): kernel share {s / ns / (s / ns + p / npp + w / nw):.4f}")

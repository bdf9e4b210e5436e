"""Opcode mix of one kernel from an `ncu --page source --csv` export: executed warp-instructions
per opcode, normalised per unit of work.
Usage: python profiles/sass_mix.py SRC.csv UNITS   (UNITS = warp-level units, e.g. warps*steps)"""
import collections
import csv
import sys


def main(path, units):
    rows = list(csv.reader(open(path)))
    h = rows[1]
    si, ei, st = h.index('Source'), h.index('Instructions Executed'), h.index('Warp Stall Sampling (All Samples)')
    mix, stall = collections.Counter(), collections.Counter()
    tot = 0
    for r in rows[2:]:
        if len(r) <= ei:
            continue
        toks = r[si].split()
        if toks and toks[0].startswith('@'):
            toks = toks[1:]
        if not toks:
            continue
        op = toks[0].split('.')[0]
        if op == 'MUFU':
            op = toks[0]
        if not (r[ei] or '0').isdigit():
            continue
        n = int(r[ei] or 0)
        mix[op] += n
        stall[op] += int(r[st] or 0)
        tot += n
    ts = sum(stall.values())
    print(f"total warp-instr {tot}  per unit {tot / units:.1f}")
    for op, n in mix.most_common(28):
        print(f"  {op:12s} {n / units:9.1f} /unit  {100 * n / tot:5.1f}%   stall-samples {100 * stall[op] / max(ts, 1):5.1f}%")


if __name__ == '__main__':
    main(sys.argv[1], float(sys.argv[2]))

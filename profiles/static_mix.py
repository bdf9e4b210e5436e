"""Static opcode histogram of one kernel in the built library (no GPU needed).
Usage: python profiles/static_mix.py <substring of the mangled kernel name> [lib]"""
import collections
import re
import subprocess
import sys

lib = sys.argv[2] if len(sys.argv) > 2 else "aircraft_b200/libaircraft_b200.so"
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
on, mix = False, collections.Counter()
for line in out.splitlines():
    if "Function :" in line:
        on = sys.argv[1] in line
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if on and m:
        op = m.group(2)
        mix[op if op.startswith("MUFU") else op.split(".")[0]] += 1
tot = sum(mix.values())
print(f"{tot} instructions")
for op, n in mix.most_common(24):
    print(f"  {n:6d} {op}")

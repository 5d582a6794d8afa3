#!/usr/bin/env python3
"""Executed-instruction mix of one kernel from an ncu report (source page): python tools/ncu_opmix.py rep kernel_regex"""
import csv, collections, subprocess, sys
rep, k = sys.argv[1], sys.argv[2]
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "-k", "regex:" + k], capture_output=True, text=True).stdout
lines = raw.splitlines()
start = [i for i, l in enumerate(lines) if l.startswith('"Address"')][0]
end = next((i for i in range(start + 1, len(lines)) if lines[i].startswith('"Kernel Name"')), len(lines))
rows = list(csv.reader(lines[start:end]))
h = rows[0]
iI, iSrc = h.index("Instructions Executed"), h.index("Source")
mix = collections.Counter()
for r in rows[1:]:
    src = r[iSrc].strip().split()
    op = src[1] if src and src[0].startswith("@") else (src[0] if src else "?")
    mix[op.split(".")[0]] += int(r[iI])
tot = sum(mix.values())
print("total warp-instr", tot)
for op, c in mix.most_common(30):
    print(f"{op:12s} {c:12d} {100*c/tot:5.1f}%")

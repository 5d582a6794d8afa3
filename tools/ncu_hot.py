#!/usr/bin/env python3
"""Top stall sites of one kernel from an ncu report: python tools/ncu_hot.py rep.ncu-rep kernel_regex [N]"""
import csv, subprocess, sys
rep, k = sys.argv[1], sys.argv[2]
n = int(sys.argv[3]) if len(sys.argv) > 3 else 25
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "-k", "regex:" + k], capture_output=True, text=True).stdout
lines = raw.splitlines()
# first launch only
start = [i for i, l in enumerate(lines) if l.startswith('"Address"')][0]
end = next((i for i in range(start + 1, len(lines)) if lines[i].startswith('"Kernel Name"')), len(lines))
rows = list(csv.reader(lines[start:end]))
h = rows[0]
iS, iI, iA, iSrc = h.index("# Samples"), h.index("Instructions Executed"), h.index("Address"), h.index("Source")
stall_cols = [(c, i) for i, c in enumerate(h) if c.startswith("stall_") and "Not Issued" not in c]
data = rows[1:]
tot = sum(int(r[iS]) for r in data)
totI = sum(int(r[iI]) for r in data)
print(f"instructions(SASS lines)={len(data)} samples={tot} warp-instr executed={totI}")
order = sorted(range(len(data)), key=lambda j: -int(data[j][iS]))[:n]
for j in sorted(order):
    r = data[j]
    st = sorted(((int(r[i]), c[6:]) for c, i in stall_cols if int(r[i]) > 0), reverse=True)[:3]
    print(f"{j:5d} {int(r[iS]):6d} {100*int(r[iS])/tot:5.1f}%  ex={int(r[iI]):8d}  {r[iSrc].strip()[:70]:70s} {st}")

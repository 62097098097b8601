#!/usr/bin/env python
"""Top stall sites from `ncu -i X.ncu-rep --page source --csv` (reads the CSV on stdin or a path)."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1]) if len(sys.argv) > 1 else sys.stdin))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
h = rows[hi]
c = {k: i for i, k in enumerate(h)}
stalls = [k for k in h if k.startswith("stall_") and "Not Issued" not in k]
data = []
for r in rows[hi + 1:]:
    if len(r) < len(h) or not r[0].startswith("0x"):
        continue
    s = int(r[c["# Samples"]])
    top = max(stalls, key=lambda k: int(r[c[k]] or 0))
    data.append((s, r[c["Source"]].strip()[:80], int(r[c["Instructions Executed"]]), top, r[c["L1 Wavefronts Shared"]], r[c["L1 Wavefronts Shared Ideal"]]))
tot = sum(d[0] for d in data) or 1
n = int(sys.argv[2]) if len(sys.argv) > 2 else 30
print(f"total samples {tot}, total warp-instr {sum(d[2] for d in data)}")
for d in sorted(data, reverse=True)[:n]:
    print(f"{100 * d[0] / tot:5.1f}%  ex={d[2]:>11}  {d[3]:<22} smemwf={d[4]}/{d[5]}  {d[1]}")

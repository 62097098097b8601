#!/usr/bin/env python
"""Aggregate `ncu -i X.ncu-rep --page source --csv` by SASS opcode: executed warp-instructions, samples, stall mix."""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
h = rows[hi]; c = {k: i for i, k in enumerate(h)}
seen = set(); ex = collections.Counter(); smp = collections.Counter(); stall = collections.Counter(); wf = collections.Counter()
stalls = [k for k in h if k.startswith("stall_") and "Not Issued" not in k]
tot = totex = 0
for r in rows[hi + 1:]:
    if len(r) < len(h) or not r[0].startswith("0x") or r[0] in seen:
        continue
    seen.add(r[0])
    op = [o for o in r[c["Source"]].strip().split() if not o.startswith("@")][0].split('.')[0]
    e = int(r[c["Instructions Executed"]]); s = int(r[c["# Samples"]])
    ex[op] += e; smp[op] += s; tot += s; totex += e
    wf[op] += int(r[c["L1 Wavefronts Shared"]] or 0) if "L1 Wavefronts Shared" in c else 0
    for k in stalls:
        stall[k] += int(r[c[k]] or 0)
print("total executed warp-instr", totex, "samples", tot)
for op, e in ex.most_common(int(sys.argv[2]) if len(sys.argv) > 2 else 20):
    print(f"{op:10s} ex={e/1e6:9.1f}M {100*e/totex:5.1f}%  samples {100*smp[op]/tot:5.1f}%  smem_wavefronts={wf[op]/1e6:.1f}M")
print([(k, round(100 * v / tot, 1)) for k, v in stall.most_common(10)])

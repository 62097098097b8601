#!/usr/bin/env python
"""Per-opcode totals (instructions, shared-memory wavefronts, stall samples) from `ncu --page source --csv --print-source sass`."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
h = rows[1]; ix = {k: i for i, k in enumerate(h)}
agg = {}; tot_w = tot_i = 0
for r in rows[2:]:
    if len(r) < len(h): continue
    src = r[ix['Source']].strip()
    opn = src.split()[1] if src.startswith('@') else src.split()[0]
    w = int(r[ix['L1 Wavefronts Shared']] or 0); wi = int(r[ix['L1 Wavefronts Shared Ideal']] or 0); ie = int(r[ix['Instructions Executed']] or 0)
    a = agg.setdefault(opn, [0, 0, 0, 0]); a[0] += ie; a[1] += w; a[2] += wi; a[3] += int(r[ix['# Samples']] or 0)
    tot_w += w; tot_i += ie
print('total instr', tot_i, 'wavefronts', tot_w)
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:int(sys.argv[2]) if len(sys.argv) > 2 else 24]:
    print(f'{k:28s} inst {v[0]:>12d} wave {v[1]:>12d} ideal {v[2]:>12d} samples {v[3]}')

#!/usr/bin/env python
"""Opcode histogram (weighted by executed warp instructions) of an `ncu --page source --csv --print-source sass` dump."""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
hdr = {h: i for i, h in enumerate(rows[1])}
ops = collections.Counter(); smp = collections.Counter(); tot = 0
for r in rows[2:]:
    try: n = float(r[hdr['Instructions Executed']]); s = float(r[hdr['# Samples']])
    except Exception: continue
    src = r[hdr['Source']].split()
    if not src: continue
    op = src[1] if src[0].startswith('@') else src[0]
    op = '.'.join(op.split('.')[:3])
    ops[op] += n; smp[op] += s; tot += n
ts = sum(smp.values())
print("total", tot)
for op, n in ops.most_common(int(sys.argv[2]) if len(sys.argv) > 2 else 30):
    print(f"{op:28s} {100*n/tot:6.2f}% inst {100*smp[op]/ts:6.2f}% samples")

#!/usr/bin/env python
"""Aggregate `ncu -i rep --page source --csv --print-source sass,cuda --kernel-name K` by CUDA
source line: top lines by sampled stalls with executed instructions and lane utilisation.
usage: ncu_src.py dump.csv [N]"""
import csv, sys
rows = csv.reader(open(sys.argv[1]))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
out, fname, hdr = [], None, None
for r in rows:
    if not r: continue
    if r[0] == 'File Path': fname = r[1].split('/')[-1]; continue
    if r[0] == 'Line No': hdr = {h: i for i, h in enumerate(r)}; continue
    if hdr is None or r[0] in ('', 'Function Name'): continue
    try:
        inst = float(r[hdr['Instructions Executed']] or 0); tinst = float(r[hdr['Thread Instructions Executed']] or 0)
        samples = float(r[hdr['# Samples']] or 0)
    except (ValueError, IndexError, KeyError):
        continue
    stalls = {k[6:]: float(r[i] or 0) for k, i in hdr.items() if k.startswith('stall_') and 'Not Issued' not in k and r[i] not in ('', '-')}
    best = sorted(stalls.items(), key=lambda kv: -kv[1])[:2]
    out.append((inst, tinst, samples, fname, r[0], r[1].strip()[:90], best))
tot_i = sum(o[0] for o in out); tot_t = sum(o[1] for o in out); tot_s = sum(o[2] for o in out)
print(f"total warp inst {tot_i:.3g}, lanes/inst {tot_t/max(tot_i,1):.2f}, samples {tot_s:.0f}")
for o in sorted(out, key=lambda o: -o[2])[:top]:
    print(f"{100*o[2]/max(tot_s,1):5.1f}% smp {100*o[0]/max(tot_i,1):5.1f}% inst lanes {o[1]/max(o[0],1):5.1f} {o[3]}:{o[4]} {o[5]} {[(k,int(v)) for k,v in o[6]]}")

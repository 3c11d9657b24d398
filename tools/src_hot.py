#!/usr/bin/env python
"""Executed warp-instructions and stall samples per CUDA source line from
   ncu -i X.ncu-rep --page source --csv --print-source cuda,sass --kernel-name regex:K > x.csv
usage: src_hot.py x.csv [warps] [top] [function-substring] [sort: inst|samp|long]"""
import csv, collections, sys
rows = list(csv.reader(open(sys.argv[1])))
warps = float(sys.argv[2]) if len(sys.argv) > 2 else 0
top = int(sys.argv[3]) if len(sys.argv) > 3 else 60
want = sys.argv[4] if len(sys.argv) > 4 else ''
order = sys.argv[5] if len(sys.argv) > 5 else 'inst'
fname = None; hdr = None; func = ''
agg = collections.Counter(); src = {}; stall = collections.Counter(); longsb = collections.Counter()
for r in rows:
    if r and r[0] in ('File Name', 'File Path'): fname = r[1].split('/')[-1]; continue
    if r and r[0] in ('Function Name', 'Kernel Name'): func = r[1]; continue
    if r and r[0] == 'Line No':
        hdr = r; ie = hdr.index('Instructions Executed'); isamp = hdr.index('# Samples')
        il = hdr.index('stall_long_sb') if 'stall_long_sb' in hdr else None
        continue
    if want and want not in func: continue
    if hdr and len(r) > ie and r[0].isdigit() and r[2] == '-':
        try: n = int(r[ie])
        except ValueError: continue
        k = (fname, int(r[0])); agg[k] += n; src[k] = r[1]
        try: stall[k] += int(r[isamp])
        except ValueError: pass
        try: longsb[k] += int(r[il]) if il is not None else 0
        except ValueError: pass
tot = sum(agg.values()) or 1; st = sum(stall.values()) or 1
print('total warp-instructions', tot, 'per warp-trip', tot / warps if warps else '', 'samples', st, 'long_sb', sum(longsb.values()))
key = {'inst': agg, 'samp': stall, 'long': longsb}[order]
for (f, l), _ in key.most_common(top):
    n = agg[(f, l)]
    print(f"{100*n/tot:5.1f}% inst {100*stall[(f,l)]/st:5.1f}% samp {100*longsb[(f,l)]/st:5.1f}% long {n/warps if warps else 0:7.1f}/trip {f}:{l}  {src[(f,l)].strip()[:100]}")

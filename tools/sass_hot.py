#!/usr/bin/env python
"""Print executed-instruction hot regions from an `ncu --page source --csv --print-source sass` export."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
lo = int(sys.argv[2]) if len(sys.argv) > 2 else None
hi = int(sys.argv[3]) if len(sys.argv) > 3 else None
out, hdr, k = [], None, 0
for r in rows:
    if r and r[0] == 'Kernel Name':
        k += 1
        if k > 1:
            break
        continue
    if r and r[0] == 'Address':
        hdr = r
        continue
    if hdr and len(r) > 5:
        out.append(r)
ie, src, te = hdr.index('Instructions Executed'), hdr.index('Source'), hdr.index('Avg. Threads Executed')
tot = sum(int(r[ie]) for r in out)
if lo is None:
    print('static', len(out), 'dyn', tot)
    B = 25
    buck = [(sum(int(r[ie]) for r in out[i:i + B]), i) for i in range(0, len(out), B)]
    acc = 0
    for v, i in sorted(buck, reverse=True)[:40]:
        acc += v
        print('%5d %6.2f%% cum %6.2f%%  count/instr %d' % (i, 100 * v / tot, 100 * acc / tot, v // B))
else:
    for i in range(lo, hi):
        r = out[i]
        print(i, r[ie].rjust(10), r[te].rjust(3), r[src].strip()[:100])

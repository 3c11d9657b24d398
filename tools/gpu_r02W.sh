#!/bin/bash
set -u
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -n 1
for w in c5 c5 c4; do
timeout 600 python bench.py --workload $w --steps 3 --warmup 3 --no-cpu --no-secondary 2>/dev/null | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('$w', round(d['value']), round(d['ms_per_step'],1), [(k['class'], round(k['ms'],1)) for k in d['roofline']['kernels']])
"
done
PTB_SPECIALIZE=1 timeout 300 python tools/c3run.py 2>&1 | tail -n 1 | cut -c1-60

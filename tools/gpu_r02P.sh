#!/bin/bash
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -n 3
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu --no-secondary 2>/dev/null | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('c5', round(d['value']), round(d['ms_per_step'],1), 'e2e', round(d['e2e']['value']), d['e2e']['parts_ms'])
"

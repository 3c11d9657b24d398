#!/bin/bash
set -u
timeout 900 python -m pytest tests/test_gpu_render.py tests/test_gpu_fullsize.py tests/test_gpu_adaptive.py -m gpu -x -q 2>&1 | tail -n 1
for i in 1 2; do
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu --no-secondary 2>/dev/null | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('c5', round(d['value']), round(d['ms_per_step'],1), [(k['class'], round(k['ms'],1)) for k in d['roofline']['kernels']])
"
done

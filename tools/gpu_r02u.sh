#!/bin/bash
# CTAs-per-SM sweep of the shade kernels after the texture / reduction changes
set -u
TAG=${1:-r02u}
O=gpurun_out; mkdir -p $O
run() { # name env...
  name=$1; shift
  env "$@" timeout 400 python bench.py --workload c5 --steps 3 --warmup 3 --no-cpu --no-secondary 2>/dev/null | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('$name', round(d['ms_per_step'],1), [(k['class'], round(k['ms'],1)) for k in d['roofline']['kernels']])
"
}
run default PTB_X=0
run primary4 PTB_SHADE_CTAS=4
run primary7 PTB_SHADE_CTAS=7
run primary8 PTB_SHADE_CTAS=8
run bounce4 PTB_SHADE_CTAS_BOUNCE=4
run bounce8 PTB_SHADE_CTAS_BOUNCE=8

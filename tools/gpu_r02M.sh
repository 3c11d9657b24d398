#!/bin/bash
set -u
TAG=${1:-r02M}
O=gpurun_out; mkdir -p $O
run() { name=$1; w=$2; shift; shift
  env "$@" timeout 600 python bench.py --workload $w --steps 2 --warmup 3 --no-cpu --no-secondary 2>$O/bench_${name}_$TAG.err | tee $O/bench_${name}_$TAG.json | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('$name', round(d['value']), round(d['ms_per_step'],1), [(k['class'], k['kernel'][:10], round(k['ms'],1)) for k in d['roofline']['kernels']])
"
}
run c4_default c4 PTB_X=0
run c4_allspec c4 PTB_SPEC_MAX_WORDS=4096
run c4_allspec2 c4 PTB_SPEC_MAX_WORDS=4096

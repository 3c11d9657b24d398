#!/bin/bash
# the fallback interpreter behind the hot functions of the generated modules (function order = mangled-name order): several processes each
set -u
TAG=${1:-r02S}
O=gpurun_out; mkdir -p $O
run() { name=$1; w=$2; shift; shift
  env "$@" timeout 600 python bench.py --workload $w --steps 2 --warmup 3 --no-cpu --no-secondary 2>$O/bench_${name}_$TAG.err | tee $O/bench_${name}_$TAG.json | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('$name', round(d['value']), round(d['ms_per_step'],1), [(k['class'], round(k['ms'],1)) for k in d['roofline']['kernels']])
"
}
for i in 1 2 3; do run c5_$i c5 PTB_X=0; done
for i in 1 2 3; do run c4_$i c4 PTB_X=0; done
run c5_inline c5 PTB_SPEC_COLD=0
run c5_inline2 c5 PTB_SPEC_COLD=0
for i in 1 2; do PTB_SPECIALIZE=1 timeout 300 python tools/c3run.py 2>&1 | tail -n 1 | cut -c1-60; done
timeout 900 python -m pytest tests/test_specialize.py -m gpu -x -q 2>&1 | tail -n 1

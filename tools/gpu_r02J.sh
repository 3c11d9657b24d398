#!/bin/bash
# lean out-of-line 1x1 merges (PTB_SPEC_COLD=2) against inline, several processes each (module placement varies per process)
set -u
TAG=${1:-r02J}
O=gpurun_out; mkdir -p $O
run() { name=$1; w=$2; shift; shift
  env "$@" timeout 600 python bench.py --workload $w --steps 2 --warmup 3 --no-cpu --no-secondary 2>$O/bench_${name}_$TAG.err | tee $O/bench_${name}_$TAG.json | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('$name', round(d['value']), round(d['ms_per_step'],1), [(k['class'], round(k['ms'],1)) for k in d['roofline']['kernels']])
"
}
for i in 1 2 3; do
run c5_inline$i c5 PTB_X=0
run c5_lean$i c5 PTB_SPEC_COLD=2
done
run c4_lean c4 PTB_SPEC_COLD=2
run c4_cold c4 PTB_X=0
PTB_SPEC_COLD=2 PTB_SPECIALIZE=1 timeout 300 python tools/c3run.py 2>&1 | tail -n 1 | cut -c1-80
PTB_SPECIALIZE=1 timeout 300 python tools/c3run.py 2>&1 | tail -n 1 | cut -c1-80

#!/bin/bash
set -u
TAG=${1:-r02I}
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_specialize.py tests/test_gpu_intersect.py -m gpu -x -q > $O/pytest_spec_$TAG.log 2>&1; echo "pytest rc=$?"; tail -n 2 $O/pytest_spec_$TAG.log
run() { name=$1; w=$2; shift; shift
  env "$@" timeout 600 python bench.py --workload $w --steps 3 --warmup 3 --no-cpu --no-secondary 2>$O/bench_${name}_$TAG.err | tee $O/bench_${name}_$TAG.json | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('$name', round(d['value']), round(d['ms_per_step'],1), [(k['class'], k['kernel'][:12], round(k['ms'],1)) for k in d['roofline']['kernels']])
"
}
run c5_cold c5 PTB_X=0
run c5_old c5 PTB_SPEC_COLD_BLOCKS=0 PTB_SPEC_HINTS=0
run c5_cold2 c5 PTB_X=0
run c5_cold_nohint c5 PTB_SPEC_HINTS=0
run c4 c4 PTB_X=0
PTB_SPECIALIZE=1 timeout 300 python tools/c3run.py 2>&1 | tail -n 1 | cut -c1-120

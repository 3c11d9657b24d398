#!/bin/bash
# full GPU tests (default and with PTB_SPECIALIZE=1: every render through the specialised kernels where they apply), C4 x3, C5
set -u
TAG=${1:-r02y}
O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?"; tail -n 3 $O/pytest_gpu_$TAG.log
PTB_SPECIALIZE=1 timeout 1500 python -m pytest tests/test_gpu_render.py tests/test_gpu_fullsize.py tests/test_gpu_adaptive.py -m gpu -x -q > $O/pytest_gpu_spec1_$TAG.log 2>&1; echo "pytest(spec=1) rc=$?"; tail -n 3 $O/pytest_gpu_spec1_$TAG.log
run() { name=$1; w=$2; shift; shift
  env "$@" timeout 600 python bench.py --workload $w --steps 3 --warmup 3 --no-cpu --no-secondary 2>$O/bench_${name}_$TAG.err | tee $O/bench_${name}_$TAG.json | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('$name', round(d['value']), round(d['ms_per_step'],1), [(k['class'], k['kernel'][:12], round(k['ms'],1)) for k in d['roofline']['kernels']])
"
}
run c4_a c4 PTB_X=0
run c4_b c4 PTB_X=0
run c4_c c4 PTB_X=0
run c4_interp c4 PTB_SPEC_LARGE=0
run c5 c5 PTB_X=0

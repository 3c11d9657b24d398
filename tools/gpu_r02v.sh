#!/bin/bash
# C4 through the specialised kernels (shape functions + bounding spheres at the call sites) against the interpreter
set -u
TAG=${1:-r02v}
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_specialize.py -m gpu -x -q > $O/pytest_spec_$TAG.log 2>&1; echo "pytest rc=$?"; tail -n 3 $O/pytest_spec_$TAG.log
run() { name=$1; shift
  env "$@" timeout 600 python bench.py --workload c4 --steps 3 --warmup 3 --no-cpu --no-secondary 2>$O/bench_c4_${name}_$TAG.err | tee $O/bench_c4_${name}_$TAG.json | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('$name', round(d['value']), round(d['ms_per_step'],1), [(k['class'], k['kernel'][:22], round(k['ms'],1)) for k in d['roofline']['kernels']])
"
}
run interp PTB_X=0
run spec PTB_SPEC_MAX_WORDS=4096
run spec_nobounce PTB_SPEC_MAX_WORDS=4096 PTB_SPEC_BOUNCE=0
run spec_nobounds PTB_SPEC_MAX_WORDS=4096 PTB_SPEC_BOUNDS=0 PTB_SPEC_BOUNCE=0

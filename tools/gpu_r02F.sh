#!/bin/bash
set -u
TAG=${1:-r02F}
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_specialize.py -m gpu -x -q > $O/pytest_spec_$TAG.log 2>&1; echo "pytest rc=$?"; tail -n 2 $O/pytest_spec_$TAG.log
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
PTB_SPECIALIZE=1 PTB_SPEC_LINEINFO=1 timeout 600 ncu --set full --clock-control none --import-source on -k regex:'ptbSpecTracePrimary' -c 1 -o $O/full_c4_spec_$TAG -f python bench.py --workload c4 --steps 1 --warmup 0 --spp 16 --no-cpu --no-secondary > $O/ncu_c4_spec_$TAG.log 2>&1
tail -n 1 $O/ncu_c4_spec_$TAG.log
run c5_inline c5 PTB_X=0
run c5_cold c5 PTB_SPEC_COLD=1
run c5_inline2 c5 PTB_X=0
run c5_cold2 c5 PTB_SPEC_COLD=1

#!/bin/bash
# shape functions with the table in shared memory: C4 A/B, ncu
set -u
TAG=${1:-r02g}
O=gpurun_out; mkdir -p $O
PTB_SPEC_MAX_WORDS=512 timeout 1200 python -m pytest tests/test_specialize.py -m gpu -q -x > $O/pytest_spec_$TAG.log 2>&1; echo "rc=$?" >> $O/pytest_spec_$TAG.log
tail -n 5 $O/pytest_spec_$TAG.log
run() { name=$1; w=$2; shift; shift
  env "$@" timeout 600 python bench.py --workload $w --steps 3 --warmup 4 --no-cpu --no-secondary > $O/bench_${w}_${name}_$TAG.json 2> $O/bench_${w}_${name}_$TAG.err
  python -c "
import json;d=json.load(open('$O/bench_${w}_${name}_$TAG.json'));print('$name: $w',round(d['value']),'Mrays/s',round(d['ms_per_step'],1),'ms',{k:round(v,1) for k,v in d['stage_ms'].items()},d['trace_kernels']['specialized_passes'],d['trace_kernels']['nvrtc_compile_ms'])" || tail -3 $O/bench_${w}_${name}_$TAG.err
}
run shapes c4 PTB_SPEC_MAX_WORDS=512
run shapes_nobounce c4 PTB_SPEC_MAX_WORDS=512 PTB_SPEC_BOUNCE=0
run default c4 X=1
run default c5 X=1
PTB_SPEC_MAX_WORDS=512 PTB_SPECIALIZE=1 PTB_SPEC_LINEINFO=1 timeout 300 python bench.py --workload c4 --steps 1 --warmup 0 --spp 16 --no-cpu --no-secondary > $O/plain_ncu_$TAG.log 2>&1 && \
PTB_SPEC_MAX_WORDS=512 PTB_SPECIALIZE=1 PTB_SPEC_LINEINFO=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:'ptbSpec' -c 2 -o $O/full_c4_$TAG -f python bench.py --workload c4 --steps 1 --warmup 0 --spp 16 --no-cpu --no-secondary > $O/ncu_c4_$TAG.log 2>&1
tail -n 2 $O/ncu_c4_$TAG.log

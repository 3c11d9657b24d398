#!/bin/bash
set -u
TAG=${1:-r02j}
O=gpurun_out; mkdir -p $O
run() { name=$1; w=$2; shift; shift
  env "$@" timeout 600 python bench.py --workload $w --steps 3 --warmup 3 --no-cpu --no-secondary > $O/bench_${w}_${name}_$TAG.json 2> $O/bench_${w}_${name}_$TAG.err
  python -c "
import json;d=json.load(open('$O/bench_${w}_${name}_$TAG.json'));print('$name: $w',round(d['value']),'Mrays/s',round(d['ms_per_step'],1),'ms',{k:round(v,1) for k,v in d['stage_ms'].items()})" || tail -3 $O/bench_${w}_${name}_$TAG.err
}
run base6 c5 X=1
run p7 c5 PTB_SHADE_CTAS=7
run p8 c5 PTB_SHADE_CTAS=8
run t10 c5 PTB_SPEC_MIN_CTAS=10
run t12 c5 PTB_SPEC_MIN_CTAS=12
run t6 c5 PTB_SPEC_MIN_CTAS=6
run tail3 c5 PTB_TAIL_LEVEL=3
run tail5 c5 PTB_TAIL_LEVEL=5

#!/bin/bash
# shade-kernel occupancy variants, paths in flight
set -u
TAG=${1:-r02i}
O=gpurun_out; mkdir -p $O
run() { name=$1; w=$2; shift; shift
  env "$@" timeout 600 python bench.py --workload $w --steps 3 --warmup 3 --no-cpu --no-secondary > $O/bench_${w}_${name}_$TAG.json 2> $O/bench_${w}_${name}_$TAG.err
  python -c "
import json;d=json.load(open('$O/bench_${w}_${name}_$TAG.json'));print('$name: $w',round(d['value']),'Mrays/s',round(d['ms_per_step'],1),'ms',{k:round(v,1) for k,v in d['stage_ms'].items()})" || tail -3 $O/bench_${w}_${name}_$TAG.err
}
run base c5 X=1
run p5 c5 PTB_SHADE_CTAS=5
run p6 c5 PTB_SHADE_CTAS=6
run b8 c5 PTB_SHADE_CTAS_BOUNCE=8
run b4 c5 PTB_SHADE_CTAS_BOUNCE=4
run fl256 c5 PTB_PATHS_IN_FLIGHT=268435456
run fl64 c5 PTB_PATHS_IN_FLIGHT=67108864
run p6 c4 PTB_SHADE_CTAS=6
run fl256 c4 PTB_PATHS_IN_FLIGHT=268435456

#!/bin/bash
# specialised C4 camera kernel against instruction-cache misses: trip sync, fewer CTAs
set -u
TAG=${1:-r02x}
O=gpurun_out; mkdir -p $O
run() { name=$1; shift
  env "$@" timeout 600 python bench.py --workload c4 --steps 2 --warmup 3 --no-cpu --no-secondary 2>$O/bench_c4_${name}_$TAG.err | tee $O/bench_c4_${name}_$TAG.json | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('$name', round(d['value']), round(d['ms_per_step'],1), [(k['class'], k['kernel'][:12], round(k['ms'],1)) for k in d['roofline']['kernels']])
"
}
run spec PTB_SPEC_MAX_WORDS=4096 PTB_SPEC_BOUNCE=0
run spec_sync PTB_SPEC_MAX_WORDS=4096 PTB_SPEC_BOUNCE=0 PTB_SPEC_TRIP_SYNC=1
run spec_c4 PTB_SPEC_MAX_WORDS=4096 PTB_SPEC_BOUNCE=0 PTB_SPEC_MIN_CTAS=4
run spec_c6 PTB_SPEC_MAX_WORDS=4096 PTB_SPEC_BOUNCE=0 PTB_SPEC_MIN_CTAS=6
run spec_sync_c6 PTB_SPEC_MAX_WORDS=4096 PTB_SPEC_BOUNCE=0 PTB_SPEC_MIN_CTAS=6 PTB_SPEC_TRIP_SYNC=1

#!/bin/bash
set -u
TAG=${1:-r02w}
O=gpurun_out; mkdir -p $O
PTB_SPEC_MAX_WORDS=4096 PTB_SPEC_BOUNCE=0 PTB_SPECIALIZE=1 PTB_SPEC_LINEINFO=1 timeout 600 ncu --set full --clock-control none --import-source on -k regex:'ptbSpecTracePrimary' -c 1 -o $O/full_c4_spec_$TAG -f python bench.py --workload c4 --steps 1 --warmup 0 --spp 16 --no-cpu --no-secondary > $O/ncu_c4_spec_$TAG.log 2>&1
tail -n 1 $O/ncu_c4_spec_$TAG.log

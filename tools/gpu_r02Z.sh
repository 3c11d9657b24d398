#!/bin/bash
# ncu full captures of the last code of the round (for profiles/ncu_kernels.json)
set -u
TAG=${1:-r02Z}
O=gpurun_out; mkdir -p $O
for w in c5 c4; do
PTB_SPECIALIZE=1 PTB_SPEC_LINEINFO=1 timeout 300 python bench.py --workload $w --steps 1 --warmup 0 --spp 16 --no-cpu --no-secondary > $O/plain_ncu_${w}_$TAG.log 2>&1 && \
PTB_SPECIALIZE=1 PTB_SPEC_LINEINFO=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:'race|shadeKernel' -c 4 -o $O/full_${w}_$TAG -f python bench.py --workload $w --steps 1 --warmup 0 --spp 16 --no-cpu --no-secondary > $O/ncu_full_${w}_$TAG.log 2>&1
tail -n 1 $O/ncu_full_${w}_$TAG.log
done
PTB_SPECIALIZE=1 PTB_SPEC_LINEINFO=1 timeout 300 python tools/c3run.py > $O/plain_c3_$TAG.log 2>&1 && \
PTB_SPECIALIZE=1 PTB_SPEC_LINEINFO=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:'ntersect' -s 3 -c 1 -o $O/full_c3_$TAG -f python tools/c3run.py > $O/ncu_c3_$TAG.log 2>&1
tail -n 1 $O/ncu_c3_$TAG.log

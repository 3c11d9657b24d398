#!/bin/bash
# ncu full capture: first primary + bounce launches of workload $1, tag $2
set -u
W=${1:-c4}; TAG=${2:-cur}
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'trace|shadeKernel' -c 4 -o gpurun_out/full_${W}_${TAG} -f python bench.py --workload $W --steps 1 --warmup 0 --spp 8 --no-cpu > gpurun_out/ncu_full_${W}_${TAG}.log 2>&1
tail -n 2 gpurun_out/ncu_full_${W}_${TAG}.log | cut -c1-200

#!/bin/bash
# GPU box script: ncu full capture of the first primary + bounce launches of a workload ($1, default c4).
set -u
W=${1:-c4}; TAG=${2:-cur}
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'primaryKernel|bounceKernel' -c 3 -o gpurun_out/full_${W}_${TAG} -f python bench.py --workload $W --steps 1 --warmup 0 --spp 8 --no-cpu > gpurun_out/ncu_full_${W}_${TAG}.log 2>&1
tail -n 3 gpurun_out/ncu_full_${W}_${TAG}.log | cut -c1-300
for spp in 2 4 8 16; do
 timeout 300 python bench.py --workload $W --no-cpu --steps 2 --warmup 2 --samples-per-pass $spp 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.read());print('$W spp/pass $spp',d['value'],'Mrays/s',d['ms_per_step'],'ms',d['stage_ms'])"
done

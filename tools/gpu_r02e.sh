#!/bin/bash
# pixel-major order + warp-segmented accumulate + terminal materials: parity, bench, ncu of all C5 kernels
set -u
TAG=${1:-r02e}
O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q -x > $O/pytest_gpu_$TAG.log 2>&1; echo "all rc=$?" >> $O/pytest_gpu_$TAG.log
tail -n 12 $O/pytest_gpu_$TAG.log
for w in c5 c4; do
timeout 600 python bench.py --workload $w --steps 3 --warmup 3 --no-cpu --no-secondary > $O/bench_${w}_$TAG.json 2> $O/bench_${w}_$TAG.err
python -c "
import json;d=json.load(open('$O/bench_${w}_$TAG.json'));print('$w',round(d['value']),'Mrays/s',round(d['ms_per_step'],1),'ms e2e',round(d['e2e']['ms_per_step'],1),{k:round(v,1) for k,v in d['stage_ms'].items()})" || tail -3 $O/bench_${w}_$TAG.err
done
PTB_SPECIALIZE=1 PTB_SPEC_LINEINFO=1 timeout 300 python bench.py --steps 1 --warmup 0 --spp 16 --no-cpu --no-secondary > $O/plain_ncu_$TAG.log 2>&1 && \
PTB_SPECIALIZE=1 PTB_SPEC_LINEINFO=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:'ptbSpec|shadeKernel' -c 4 -o $O/full_c5_$TAG -f python bench.py --steps 1 --warmup 0 --spp 16 --no-cpu --no-secondary > $O/ncu_c5_$TAG.log 2>&1
tail -n 2 $O/ncu_c5_$TAG.log

#!/bin/bash
# GPU box script: smoke, parity tests, benches (ours + reference arm), ncu launch list + full captures.
# Usage: bash tools/gpu_full.sh TAG     -> outputs in gpurun_out/*_TAG.*
set -u
TAG=${1:-cur}
O=gpurun_out; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/smi_$TAG.txt 2>&1
timeout 600 python __graft_entry__.py smoke > $O/smoke_$TAG.log 2>&1; echo "smoke rc=$?" >> $O/smoke_$TAG.log
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu_$TAG.log
tail -n 3 $O/smoke_$TAG.log $O/pytest_gpu_$TAG.log
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_ref_$TAG.json 2> $O/bench_ref_$TAG.err
timeout 900 python bench.py > $O/bench_c5_$TAG.json 2> $O/bench_c5_$TAG.err; echo "rc=$?" >> $O/bench_c5_$TAG.err
timeout 900 python bench.py --workload c4 > $O/bench_c4_$TAG.json 2> $O/bench_c4_$TAG.err; echo "rc=$?" >> $O/bench_c4_$TAG.err
for w in c5 c4; do python -c "
import json;d=json.load(open('$O/bench_${w}_$TAG.json'));print('$w',round(d['value']),'Mrays/s e2e',round(d['e2e']['value']),'frac',round(d['roofline']['frac'],4),d['roofline']['kernel'],'cpu',d['cpu_baseline'] and round(d['cpu_baseline']['value'],1))"; done
# ncu: launch list of a short run of the same command (16 spp so the list is bounded), then full captures
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/launches_c5_$TAG.csv python bench.py --steps 1 --warmup 1 --spp 16 --no-cpu > $O/ncu_list_$TAG.log 2>&1
for w in c5 c4; do
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'race|shadeKernel' -c 4 -o $O/full_${w}_$TAG -f python bench.py --workload $w --steps 1 --warmup 0 --spp 8 --no-cpu > $O/ncu_full_${w}_$TAG.log 2>&1
done
ls -la $O | tail -n 20

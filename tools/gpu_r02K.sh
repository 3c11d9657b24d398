#!/bin/bash
# Full default bench (all legs) + reference arm, timed; ncu launch list; full captures for profiles/ncu_kernels.json
set -u
TAG=${1:-r02K}
O=gpurun_out; mkdir -p $O
t0=$(date +%s)
timeout 600 python bench.py --impl reference > $O/bench_ref_$TAG.json 2> $O/bench_ref_$TAG.err; echo "ref rc=$? $(( $(date +%s) - t0 )) s"
t0=$(date +%s)
timeout 900 python bench.py > $O/bench_default_$TAG.json 2> $O/bench_default_$TAG.err; echo "default rc=$? $(( $(date +%s) - t0 )) s"
tail -n 3 $O/bench_default_$TAG.err
timeout 300 python bench.py --steps 1 --warmup 1 --spp 16 --no-cpu --no-secondary > $O/plain_list_$TAG.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_c5_$TAG.csv python bench.py --steps 1 --warmup 1 --spp 16 --no-cpu --no-secondary > $O/ncu_list_$TAG.log 2>&1
for w in c5 c4; do
PTB_SPECIALIZE=1 PTB_SPEC_LINEINFO=1 timeout 300 python bench.py --workload $w --steps 1 --warmup 0 --spp 16 --no-cpu --no-secondary > $O/plain_ncu_${w}_$TAG.log 2>&1 && \
PTB_SPECIALIZE=1 PTB_SPEC_LINEINFO=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:'race|shadeKernel' -c 4 -o $O/full_${w}_$TAG -f python bench.py --workload $w --steps 1 --warmup 0 --spp 16 --no-cpu --no-secondary > $O/ncu_full_${w}_$TAG.log 2>&1
tail -n 1 $O/ncu_full_${w}_$TAG.log
done
PTB_SPECIALIZE=1 PTB_SPEC_LINEINFO=1 timeout 300 python tools/c3run.py > $O/plain_c3_$TAG.log 2>&1 && \
PTB_SPECIALIZE=1 PTB_SPEC_LINEINFO=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:'ntersect' -s 3 -c 1 -o $O/full_c3_$TAG -f python tools/c3run.py > $O/ncu_c3_$TAG.log 2>&1
tail -n 1 $O/plain_c3_$TAG.log; tail -n 1 $O/ncu_c3_$TAG.log
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?"; tail -n 2 $O/pytest_gpu_$TAG.log

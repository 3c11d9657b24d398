#!/bin/bash
# A/B of the two-ended (direction-sorted) ray queues: PTB_BIN_AXIS = -1 (one class) / 0 / 1 / 2 on C5 and C4, then the GPU tests
set -u
TAG=${1:-r02l}
O=gpurun_out; mkdir -p $O
for w in c5 c4; do
for ax in -1 1 0 2; do
PTB_BIN_AXIS=$ax timeout 400 python bench.py --workload $w --steps 3 --warmup 3 --no-cpu --no-secondary > $O/bench_${w}_ax${ax}_$TAG.json 2> $O/bench_${w}_ax${ax}_$TAG.err
echo "$w axis $ax rc=$?"; python - <<P
import json
for l in open("$O/bench_${w}_ax${ax}_$TAG.json"):
    if l.startswith("{"):
        d=json.loads(l); print(d["value"], d["ms_per_step"], [(k["class"], round(k["ms"],1)) for k in d["roofline"]["kernels"]])
P
done
done
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?"; tail -n 5 $O/pytest_gpu_$TAG.log

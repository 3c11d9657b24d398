#!/bin/bash
set -u
TAG=${1:-r02t}
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_specialize.py tests/test_gpu_render.py tests/test_gpu_intersect.py -m gpu -x -q > $O/pytest_bounds_$TAG.log 2>&1; echo "pytest rc=$?"; tail -n 3 $O/pytest_bounds_$TAG.log
for v in 1 0; do
PTB_FAST_BOUNDS=$v timeout 400 python bench.py --workload c4 --steps 3 --warmup 2 --no-cpu --no-secondary > $O/bench_c4_b${v}_$TAG.json 2> $O/bench_c4_b${v}_$TAG.err
python - <<P
import json
for l in open("$O/bench_c4_b${v}_$TAG.json"):
    if l.startswith("{"):
        d=json.loads(l); print("bounds=$v c4", d["value"], d["ms_per_step"], [(k["class"], round(k["ms"],1)) for k in d["roofline"]["kernels"]])
P
done
PTB_FAST_BOUNDS=1 timeout 600 ncu --set full --clock-control none --import-source on -k regex:'tracePrimary' -c 1 -o $O/full_c4_b1_$TAG -f python bench.py --workload c4 --steps 1 --warmup 0 --spp 16 --no-cpu --no-secondary > $O/ncu_c4_b1_$TAG.log 2>&1
tail -n 1 $O/ncu_c4_b1_$TAG.log

#!/bin/bash
# bounding spheres of sphere runs: GPU tests, then C4 with (default) and without (PTB_FAST_BOUNDS=0) them
set -u
TAG=${1:-r02r}
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_specialize.py tests/test_gpu_render.py tests/test_gpu_csg.py -m gpu -x -q > $O/pytest_bounds_$TAG.log 2>&1; echo "pytest rc=$?"; tail -n 3 $O/pytest_bounds_$TAG.log
for v in 1 0; do
PTB_FAST_BOUNDS=$v timeout 400 python bench.py --workload c4 --steps 3 --warmup 2 --no-cpu --no-secondary > $O/bench_c4_b${v}_$TAG.json 2> $O/bench_c4_b${v}_$TAG.err
python - <<P
import json
for l in open("$O/bench_c4_b${v}_$TAG.json"):
    if l.startswith("{"):
        d=json.loads(l); print("bounds=$v c4", d["value"], d["ms_per_step"], [(k["class"], round(k["ms"],1)) for k in d["roofline"]["kernels"]])
P
done
timeout 400 python bench.py --workload c5 --steps 3 --warmup 3 --no-cpu --no-secondary 2>/dev/null | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('c5', d['value'], d['ms_per_step'], [(k['class'], round(k['ms'],1)) for k in d['roofline']['kernels']])
"

#!/bin/bash
# specialised-evaluator tests, then C5 bench twice (kernel split) and C3 leg
set -u
TAG=${1:-r02p}
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_specialize.py tests/test_gpu_render.py -m gpu -x -q > $O/pytest_spec_$TAG.log 2>&1; echo "pytest rc=$?"; tail -n 3 $O/pytest_spec_$TAG.log
for w in c5 c5; do
timeout 400 python bench.py --workload $w --steps 3 --warmup 3 --no-cpu --no-secondary > $O/bench_${w}_$TAG.json 2> $O/bench_${w}_$TAG.err
python - <<P
import json
for l in open("$O/bench_${w}_$TAG.json"):
    if l.startswith("{"):
        d=json.loads(l); print("$w", d["value"], d["ms_per_step"], [(k["class"], round(k["ms"],1)) for k in d["roofline"]["kernels"]])
P
done
PTB_SPECIALIZE=1 timeout 300 python tools/c3run.py 2>&1 | tail -n 2

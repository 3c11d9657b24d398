#!/bin/bash
set -u
TAG=${1:-r02U}
O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?"; tail -n 2 $O/pytest_gpu_$TAG.log
for i in 1 2; do
timeout 900 python bench.py > $O/bench_default_${i}_$TAG.json 2> $O/bench_default_${i}_$TAG.err
python - <<P
import json
for l in open("$O/bench_default_${i}_$TAG.json"):
    if l.startswith("{"):
        d=json.loads(l); print("value", round(d["value"]), round(d["ms_per_step"],1), "e2e", round(d["e2e"]["value"]), round(d["e2e"]["ms_per_step"],1), d["e2e"]["parts_ms"], "c4", round(d["secondary"]["c4"]["value"]), "c3", round(d["secondary"]["c3"]["value"]))
P
done

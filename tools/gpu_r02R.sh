#!/bin/bash
# 8 x B200, final code: the bench as the driver launches it at N = 8 (ours), 6 timed steps
set -u
TAG=${1:-r02R}
O=gpurun_out; mkdir -p $O
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus 8 --steps 6 --warmup 3 > $O/bench_c5_gpus8_$TAG.json 2> $O/bench_c5_gpus8_$TAG.err; echo "N=8 rc=$?"
python - <<P
import json
for l in open("$O/bench_c5_gpus8_$TAG.json"):
    if l.startswith("{"):
        d=json.loads(l); print(8, "value", round(d["value"]), "ms", round(d["ms_per_step"],1), "e2e", round(d["e2e"]["value"]), round(d["e2e"]["ms_per_step"],1), d["e2e"].get("parts_ms"))
P

#!/bin/bash
# e2e with page-locked inputs: 1 GPU and (when the box has them) 2 GPUs
set -u
TAG=${1:-r02C}
O=gpurun_out; mkdir -p $O
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu --no-secondary > $O/bench_c5_e2e1_$TAG.json 2> $O/bench_c5_e2e1_$TAG.err
N=$(nvidia-smi -L | wc -l)
if [ $N -ge 2 ]; then
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 3 --warmup 3 > $O/bench_c5_e2e2_$TAG.json 2> $O/bench_c5_e2e2_$TAG.err
fi
python - <<P
import json,glob
for f in sorted(glob.glob("$O/bench_c5_e2e?_$TAG.json")):
    for l in open(f):
        if l.startswith("{"):
            d=json.loads(l); print(f[-14:], "value", round(d["value"]), "ms", round(d["ms_per_step"],1), "e2e", round(d["e2e"]["value"]), round(d["e2e"]["ms_per_step"],1), d["e2e"].get("parts_ms"))
P

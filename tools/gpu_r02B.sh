#!/bin/bash
# 8 x B200: multi-GPU tests (ctypes + C++-only caller), then the bench as the driver launches it at N = 8, 4, 2 (ours; reference arm at 8)
set -u
TAG=${1:-r02B}
O=gpurun_out; mkdir -p $O
nvidia-smi -L | head -8 > $O/smi_$TAG.txt
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > $O/pytest_multi_$TAG.log 2>&1; echo "pytest rc=$?"; tail -n 3 $O/pytest_multi_$TAG.log
for N in 8 4 2; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N bench.py --gpus $N --steps 3 --warmup 3 > $O/bench_c5_gpus${N}_$TAG.json 2> $O/bench_c5_gpus${N}_$TAG.err; echo "N=$N rc=$?"
python - <<P
import json
for l in open("$O/bench_c5_gpus${N}_$TAG.json"):
    if l.startswith("{"):
        d=json.loads(l); print($N, "value", round(d["value"]), "ms", round(d["ms_per_step"],1), "e2e", round(d["e2e"]["value"]), round(d["e2e"]["ms_per_step"],1), d.get("scaling"), d["e2e"].get("parts_ms"))
P
done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29519 bench.py --gpus 8 --steps 3 --warmup 3 --impl reference > $O/bench_ref_gpus8_$TAG.json 2> $O/bench_ref_gpus8_$TAG.err; echo "ref rc=$?"; tail -c 400 $O/bench_ref_gpus8_$TAG.json

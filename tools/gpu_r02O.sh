#!/bin/bash
# per-step e2e parts on every rank, 8 GPUs
PTB200_E2E_DEBUG=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus 8 --steps 6 --warmup 3 2>&1 >gpurun_out/bench_c5_gpus8_r02O.json | grep "e2e step" | sort | awk '{print $4, $6, $8}' | tr '\n' ';'
echo
tail -c 700 gpurun_out/bench_c5_gpus8_r02O.json | head -c 400

#!/bin/bash
# per-step e2e parts on every rank, 2 GPUs
PTB200_E2E_DEBUG=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 4 --warmup 3 2>&1 >/dev/null | grep "e2e step"

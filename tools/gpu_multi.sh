#!/bin/bash
# N-GPU bench through torch.distributed.run (as the driver launches it)
N=${1:-2}
mkdir -p gpurun_out
for arm in ${ARMS:-ours reference}; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 2 --warmup 3 --impl $arm > gpurun_out/bench_multi_${N}_$arm.json 2> gpurun_out/bench_multi_${N}_$arm.err; echo "rc=$?"
tail -n 1 gpurun_out/bench_multi_${N}_$arm.json | cut -c1-900
tail -n 3 gpurun_out/bench_multi_${N}_$arm.err | cut -c1-300
done

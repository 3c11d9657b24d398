#!/bin/bash
# multi-GPU inside the library on 2 GPUs: C++-only test binary, ctypes tests, torchrun bench (ours + reference arm)
set -u
TAG=${1:-r02h}
O=gpurun_out; mkdir -p $O
nvidia-smi -L > $O/smi_$TAG.txt 2>&1
timeout 600 ./path-trace_b200/build/multi_gpu_main 2 192 108 64 > $O/multi_main_$TAG.log 2>&1; echo "rc=$?" >> $O/multi_main_$TAG.log; tail -n 4 $O/multi_main_$TAG.log
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -q -x > $O/pytest_multi_$TAG.log 2>&1; echo "rc=$?" >> $O/pytest_multi_$TAG.log; tail -n 12 $O/pytest_multi_$TAG.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 > $O/bench_c5_gpus2_$TAG.json 2> $O/bench_c5_gpus2_$TAG.err; echo "rc=$?" >> $O/bench_c5_gpus2_$TAG.err
tail -n 6 $O/bench_c5_gpus2_$TAG.err
python -c "
import json;d=json.load(open('$O/bench_c5_gpus2_$TAG.json'));print('2 GPUs: c5',round(d['value']),'Mrays/s',round(d['ms_per_step'],1),'ms e2e',round(d['e2e']['value']),round(d['e2e']['ms_per_step'],1),'ms',d['e2e']['parts_ms'],d['parallelism'])"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 1 --warmup 0 > $O/bench_ref_gpus2_$TAG.json 2> $O/bench_ref_gpus2_$TAG.err; echo "ref rc=$?"; cut -c1-300 $O/bench_ref_gpus2_$TAG.json
timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu --no-secondary > $O/bench_c5_gpus1_$TAG.json 2> $O/bench_c5_gpus1_$TAG.err
python -c "
import json;d=json.load(open('$O/bench_c5_gpus1_$TAG.json'));print('1 GPU: c5',round(d['value']),'Mrays/s',round(d['ms_per_step'],1),'ms e2e',round(d['e2e']['value']),round(d['e2e']['ms_per_step'],1),'ms',d['e2e']['parts_ms'])"

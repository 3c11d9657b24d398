#!/bin/bash
# GPU box script: parity tests + short benches.  Outputs into gpurun_out/.
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -n 15 gpurun_out/pytest_gpu.log
for w in c4 c5; do
  timeout 600 python bench.py --workload $w --no-cpu --steps 2 --warmup 3 > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; echo "rc=$?" >> gpurun_out/bench_$w.err
  python -c "
import json;d=json.load(open('gpurun_out/bench_$w.json'));print('$w',d['value'],'Mrays/s',d['ms_per_step'],'ms frac',d['roofline']['frac'],d['stage_ms'])"
done

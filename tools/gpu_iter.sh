#!/bin/bash
# GPU box: parity tests, short benches of c4/c5, e2e probe, samples-per-pass sweep.  Outputs in gpurun_out/.
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -n 4 gpurun_out/pytest_gpu.log
for w in c4 c5; do
  for spp in ${SPPS:-0}; do
  timeout 600 python bench.py --workload $w --no-cpu --steps 2 --warmup 3 --samples-per-pass $spp > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err; echo "rc=$?" >> gpurun_out/bench_$w.err
  python -c "
import json;d=json.load(open('gpurun_out/bench_$w.json'));print('$w spp/pass $spp',round(d['value']),'Mrays/s',round(d['ms_per_step'],1),'ms e2e',round(d['e2e']['value']),d['e2e']['parts_ms'],'frac',round(d['roofline']['frac'],4),d['stage_ms'])"
  done
done
timeout 300 python tools/gpu_e2e_probe.py 64 2>&1 | tail -4

#!/bin/bash
# GPU box: specialised-kernel parity tests, then benches with (PTB_SPECIALIZE=1) and without (=0) them.
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_specialize.py -m gpu -x -q > gpurun_out/pytest_spec.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_spec.log
tail -n 25 gpurun_out/pytest_spec.log
for mode in ${MODES:-1 0}; do
for w in ${WL:-c5 c4}; do
  PTB_SPECIALIZE=$mode timeout 600 python bench.py --workload $w --no-cpu --steps 2 --warmup 3 > gpurun_out/bench_${w}_spec$mode.json 2> gpurun_out/bench_${w}_spec$mode.err; echo "rc=$?" >> gpurun_out/bench_${w}_spec$mode.err
  python -c "
import json;d=json.load(open('gpurun_out/bench_${w}_spec$mode.json'));print('$w PTB_SPECIALIZE=$mode',round(d['value']),'Mrays/s',round(d['ms_per_step'],1),'ms e2e',round(d['e2e']['value']),'frac',round(d['roofline']['frac'],4),d['stage_ms'])" || tail -3 gpurun_out/bench_${w}_spec$mode.err
done
done

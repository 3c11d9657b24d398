#!/bin/bash
# Round 2, third GPU call: second-generation specialised kernels — parity, A/B benches, ncu.
set -u
TAG=${1:-r02c}
O=gpurun_out; mkdir -p $O
timeout 1200 python -m pytest tests/test_specialize.py tests/test_gpu_intersect.py -m gpu -q -x > $O/pytest_spec_$TAG.log 2>&1; echo "pytest rc=$?" >> $O/pytest_spec_$TAG.log
tail -n 15 $O/pytest_spec_$TAG.log
run() { # name, env...
  name=$1; shift
  env "$@" timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu --no-secondary > $O/bench_c5_${name}_$TAG.json 2> $O/bench_c5_${name}_$TAG.err
  python -c "
import json;d=json.load(open('$O/bench_c5_${name}_$TAG.json'));print('$name: c5',round(d['value']),'Mrays/s',round(d['ms_per_step'],1),'ms',{k:round(v,1) for k,v in d['stage_ms'].items()},d['trace_kernels']['specialized_passes'])" || tail -3 $O/bench_c5_${name}_$TAG.err
}
run gen2 PTB_SPEC_GEN=2
run gen1 PTB_SPEC_GEN=1
run gen2_bounce PTB_SPEC_GEN=2 PTB_SPEC_BOUNCE=1
run gen2_ctas6 PTB_SPEC_GEN=2 PTB_SPEC_MIN_CTAS=6
run gen2_bounce_ctas6 PTB_SPEC_GEN=2 PTB_SPEC_BOUNCE=1 PTB_SPEC_MIN_CTAS=6
for g in 2 1; do
  echo "c3 gen$g: $(PTB_SPEC_GEN=$g PTB_SPECIALIZE=1 timeout 300 python tools/c3run.py 2>&1 | tail -n 1)"
done
echo "c3 interpreter: $(PTB_SPECIALIZE=0 timeout 300 python tools/c3run.py 2>&1 | tail -n 1)"
PTB_SPECIALIZE=1 PTB_SPEC_LINEINFO=1 timeout 300 python bench.py --steps 1 --warmup 0 --spp 8 --no-cpu --no-secondary > $O/plain_ncu_$TAG.log 2>&1 && \
PTB_SPECIALIZE=1 PTB_SPEC_LINEINFO=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:'ptbSpec|shadeKernel' -c 3 -o $O/full_c5_$TAG -f python bench.py --steps 1 --warmup 0 --spp 8 --no-cpu --no-secondary > $O/ncu_c5_$TAG.log 2>&1
tail -n 3 $O/ncu_c5_$TAG.log
PTB_SPECIALIZE=1 PTB_SPEC_LINEINFO=1 PTB200_C3_LOG2=24 timeout 900 ncu --set full --clock-control none --import-source on -k regex:'ntersect' -s 2 -c 1 -o $O/full_c3_$TAG -f python tools/c3run.py > $O/ncu_c3_$TAG.log 2>&1
tail -n 3 $O/ncu_c3_$TAG.log
ls -la $O | tail -n 8

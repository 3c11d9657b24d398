#!/bin/bash
# GPU box: rebuild the kernels with each set of -D flags and run a short bench of $WL (default "c4 c5")
set -u
WL=${WL:-"c4 c5"}
mkdir -p gpurun_out
while [ $# -gt 0 ]; do
  V="$1"; shift
  rm -f path-trace_b200/build/ptb_kernels.o path-trace_b200/build/ptb_capi.o
  make -C path-trace_b200 EXTRA="$V" -j8 > gpurun_out/build.log 2>&1 || { echo "build failed: $V"; tail -5 gpurun_out/build.log; continue; }
  for w in $WL; do
    timeout 300 python bench.py --workload $w --no-cpu --steps 2 --warmup 2 2>gpurun_out/err.log | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read());print('[$V] $w %.0f Mrays/s %.0f ms primary %.0f bounce %.0f'%(d['value'],d['ms_per_step'],d['stage_ms']['primary'],d['stage_ms']['bounce']))
except Exception as e: print('[$V] $w failed', e)"
  done
done

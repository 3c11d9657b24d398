#!/bin/bash
# Round 2, second GPU call: new parity tests (full-size goldens, KATs, render_pixels), ncu of the C3 ray-batch kernel.
set -u
TAG=${1:-r02b}
O=gpurun_out; mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -q -s -k "golden or kat or render_pixels or counter or philox or reflect" > $O/pytest_new_$TAG.log 2>&1; echo "pytest rc=$?" >> $O/pytest_new_$TAG.log
grep -E "relRMSE|passed|failed|Error|assert" $O/pytest_new_$TAG.log | head -40
cat > /tmp/c3run.py <<'PY'
import sys, os
sys.path.insert(0, 'path-trace_b200'); sys.path.insert(0, 'tests'); sys.path.insert(0, '.')
import torch, bench
os.environ.setdefault("PTB200_C3_LOG2", "24")
print(bench.run_c3(torch, 3, 72.0, 6557.8, False))
PY
PTB_SPECIALIZE=1 timeout 300 python /tmp/c3run.py > $O/c3_plain_$TAG.log 2>&1 && \
PTB_SPECIALIZE=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:'ntersect' -s 2 -c 2 -o $O/full_c3_$TAG -f python /tmp/c3run.py > $O/ncu_c3_$TAG.log 2>&1
tail -n 3 $O/c3_plain_$TAG.log; tail -n 5 $O/ncu_c3_$TAG.log
ls -la $O | tail -n 6

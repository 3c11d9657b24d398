#!/bin/bash
# final state: GPU suite, default bench line, reference arm
set -u
TAG=${1:-r02T}
O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?"; tail -n 2 $O/pytest_gpu_$TAG.log
t0=$(date +%s); timeout 900 python bench.py > $O/bench_default_$TAG.json 2> $O/bench_default_$TAG.err; echo "default rc=$? $(( $(date +%s) - t0 )) s"
t0=$(date +%s); timeout 600 python bench.py --impl reference > $O/bench_ref_$TAG.json 2> $O/bench_ref_$TAG.err; echo "ref rc=$? $(( $(date +%s) - t0 )) s"
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -n 2

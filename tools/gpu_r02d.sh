#!/bin/bash
# no-alloc entry point: alone in a fresh process, then after all the other GPU tests; then a short bench.
set -u
TAG=${1:-r02d}
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > $O/pytest_noalloc_alone_$TAG.log 2>&1; echo "alone rc=$?" >> $O/pytest_noalloc_alone_$TAG.log
tail -n 12 $O/pytest_noalloc_alone_$TAG.log
timeout 1500 python -m pytest tests -m gpu -q -x > $O/pytest_gpu_$TAG.log 2>&1; echo "all rc=$?" >> $O/pytest_gpu_$TAG.log
tail -n 12 $O/pytest_gpu_$TAG.log
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu --no-secondary > $O/bench_c5_$TAG.json 2> $O/bench_c5_$TAG.err
python -c "
import json;d=json.load(open('$O/bench_c5_$TAG.json'));print('c5',round(d['value']),'Mrays/s',round(d['ms_per_step'],1),'ms',{k:round(v,1) for k,v in d['stage_ms'].items()})" || tail -3 $O/bench_c5_$TAG.err

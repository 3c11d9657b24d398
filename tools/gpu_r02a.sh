#!/bin/bash
# Round 2, first GPU call: smoke, parity tests, default bench (all legs), tail-level sweep.
set -u
TAG=${1:-r02a}
O=gpurun_out; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/smi_$TAG.txt 2>&1
timeout 600 python __graft_entry__.py smoke > $O/smoke_$TAG.log 2>&1; echo "smoke rc=$?" >> $O/smoke_$TAG.log
timeout 1200 python -m pytest tests -m gpu -q > $O/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu_$TAG.log
tail -n 3 $O/smoke_$TAG.log; tail -n 30 $O/pytest_gpu_$TAG.log
timeout 900 python bench.py --steps 3 --warmup 3 > $O/bench_c5_$TAG.json 2> $O/bench_c5_$TAG.err; echo "rc=$?" >> $O/bench_c5_$TAG.err
tail -n 5 $O/bench_c5_$TAG.err
python - <<PY
import json
try:
    d=json.load(open('$O/bench_c5_$TAG.json'))
    print('c5',round(d['value']),'Mrays/s e2e',round(d['e2e']['value']),'ms',round(d['ms_per_step'],1),d['stage_ms'])
    for k in d['roofline']['kernels']: print('  ',k['class'],round(k['ms'],1),'ms share',round(k['share'],3),k['bound'],'frac',round(k['frac'],4))
    s=d['secondary']
    print('c4',round(s['c4']['value']),s['c4']['stage_ms'])
    for k in s['c4']['roofline']['kernels']: print('  ',k['class'],round(k['ms'],1),'ms share',round(k['share'],3),k['bound'],'frac',round(k['frac'],4))
    c3=s['c3']; print('c3',round(c3['value']),'Mrays/s',c3['ms_per_step'],'ms frac',c3['roofline']['frac'],c3['vs_exact_policy'],c3['flagged_grazing'],c3.get('cpu_baseline'))
except Exception as e: print('bench parse failed',e)
PY
for t in 0 2 4; do
PTB_TAIL_LEVEL=$t timeout 600 python bench.py --steps 3 --warmup 2 --no-cpu --no-secondary > $O/bench_c5_tail${t}_$TAG.json 2> $O/bench_c5_tail${t}_$TAG.err
python -c "
import json;d=json.load(open('$O/bench_c5_tail${t}_$TAG.json'));print('tail level $t: c5',round(d['value']),'Mrays/s',round(d['ms_per_step'],1),'ms',d['stage_ms'],d['gpu_launches'])"
done
PTB_TAIL_LEVEL=0 timeout 600 python bench.py --workload c4 --steps 3 --warmup 2 --no-cpu --no-secondary > $O/bench_c4_tail0_$TAG.json 2> $O/bench_c4_tail0_$TAG.err
python -c "
import json;d=json.load(open('$O/bench_c4_tail0_$TAG.json'));print('tail level 0: c4',round(d['value']),'Mrays/s',round(d['ms_per_step'],1),'ms',d['stage_ms'])"
ls -la $O | tail -n 12

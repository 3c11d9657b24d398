#!/bin/bash
# texture + render tests, then C5 / C4 bench (kernel split), then ncu full capture of one C5 pass
set -u
TAG=${1:-r02n}
O=gpurun_out; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_render.py tests/test_gpu_kat.py -m gpu -x -q > $O/pytest_tex_$TAG.log 2>&1; echo "pytest rc=$?"; tail -n 3 $O/pytest_tex_$TAG.log
for w in c5 c5 c4; do
timeout 400 python bench.py --workload $w --steps 3 --warmup 3 --no-cpu --no-secondary > $O/bench_${w}_$TAG.json 2> $O/bench_${w}_$TAG.err
python - <<P
import json
for l in open("$O/bench_${w}_$TAG.json"):
    if l.startswith("{"):
        d=json.loads(l); print("$w", d["value"], d["ms_per_step"], [(k["class"], round(k["ms"],1)) for k in d["roofline"]["kernels"]])
P
done
PTB_SPECIALIZE=1 PTB_SPEC_LINEINFO=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:'race|shadeKernel' -c 4 -o $O/full_c5_$TAG -f python bench.py --workload c5 --steps 1 --warmup 0 --spp 16 --no-cpu --no-secondary > $O/ncu_full_c5_$TAG.log 2>&1
tail -n 1 $O/ncu_full_c5_$TAG.log

#!/bin/bash
# C5 bench (kernel split) + ncu full capture of the four wavefront kernels of one C5 pass
set -u
TAG=${1:-r02m}
O=gpurun_out; mkdir -p $O
for rep in 1 2; do
timeout 400 python bench.py --workload c5 --steps 3 --warmup 3 --no-cpu --no-secondary > $O/bench_c5_${rep}_$TAG.json 2> $O/bench_c5_${rep}_$TAG.err
python - <<P
import json
for l in open("$O/bench_c5_${rep}_$TAG.json"):
    if l.startswith("{"):
        d=json.loads(l); print(d["value"], d["ms_per_step"], [(k["class"], round(k["ms"],1)) for k in d["roofline"]["kernels"]])
P
done
if [ "${2:-ncu}" = "ncu" ]; then
PTB_SPECIALIZE=1 PTB_SPEC_LINEINFO=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:'race|shadeKernel' -c 4 -o $O/full_c5_$TAG -f python bench.py --workload c5 --steps 1 --warmup 0 --spp 16 --no-cpu --no-secondary > $O/ncu_full_c5_$TAG.log 2>&1
tail -n 1 $O/ncu_full_c5_$TAG.log
fi

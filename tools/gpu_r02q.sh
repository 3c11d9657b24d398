#!/bin/bash
# C4 with two rays per lane in the interpreting trace kernels (PTB_TRACE_RAYS=2) at 8 / 5 / 4 CTAs per SM; PTB_SPECIALIZE=0 keeps C5 interpreted too
set -u
TAG=${1:-r02q}
O=gpurun_out; mkdir -p $O
cp path-trace_b200/libptb200.so /tmp/libptb200_default.so
for v in default rays2_c8 rays2_c5 rays2_c4; do
if [ $v = default ]; then cp /tmp/libptb200_default.so path-trace_b200/libptb200.so; else cp path-trace_b200/variants/libptb200_$v.so path-trace_b200/libptb200.so; fi
for w in c4 c5; do
PTB_SPECIALIZE=0 timeout 400 python bench.py --workload $w --steps 2 --warmup 2 --no-cpu --no-secondary > $O/bench_${w}_${v}_$TAG.json 2> $O/bench_${w}_${v}_$TAG.err
python - <<P
import json
for l in open("$O/bench_${w}_${v}_$TAG.json"):
    if l.startswith("{"):
        d=json.loads(l); print("$v $w", d["value"], d["ms_per_step"], [(k["class"], round(k["ms"],1)) for k in d["roofline"]["kernels"]])
P
done
done
cp /tmp/libptb200_default.so path-trace_b200/libptb200.so

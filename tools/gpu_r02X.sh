#!/bin/bash
# final state of the round: default bench line (twice), reference arm
set -u
TAG=${1:-r02X}
O=gpurun_out; mkdir -p $O
for i in 1 2; do
timeout 900 python bench.py > $O/bench_default_${i}_$TAG.json 2> $O/bench_default_${i}_$TAG.err
python - <<P
import json
for l in open("$O/bench_default_${i}_$TAG.json"):
    if l.startswith("{"):
        d=json.loads(l); print("value", round(d["value"]), round(d["ms_per_step"],1), "e2e", round(d["e2e"]["value"]), round(d["e2e"]["ms_per_step"],1), d["e2e"]["parts_ms"], "c4", round(d["secondary"]["c4"]["value"]), "c3", round(d["secondary"]["c3"]["value"]), [(k["class"], round(k["ms"],1)) for k in d["roofline"]["kernels"]])
P
done
timeout 600 python bench.py --impl reference > $O/bench_ref_$TAG.json 2> $O/bench_ref_$TAG.err; tail -c 200 $O/bench_ref_$TAG.json

#!/bin/bash
# GPU box script: tests, bench, ncu launch list + one full capture.  Outputs into gpurun_out/.
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 600 python bench.py > gpurun_out/bench_c5.json 2> gpurun_out/bench_c5.err; echo "rc=$?" >> gpurun_out/bench_c5.err
timeout 600 python bench.py --workload c4 --no-cpu > gpurun_out/bench_c4.json 2> gpurun_out/bench_c4.err; echo "rc=$?" >> gpurun_out/bench_c4.err
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
# ncu: launch list (short run: 16 spp so the list is bounded), then one full capture of primary + first bounce
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_c5.csv python bench.py --steps 1 --warmup 1 --spp 16 --no-cpu > gpurun_out/ncu_list.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'primaryKernel|bounceKernel' -c 3 -o gpurun_out/full_c5 -f python bench.py --steps 1 --warmup 0 --spp 8 --no-cpu > gpurun_out/ncu_full.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'primaryKernel|bounceKernel' -c 3 -o gpurun_out/full_c4 -f python bench.py --workload c4 --steps 1 --warmup 0 --spp 8 --no-cpu > gpurun_out/ncu_full_c4.log 2>&1
ls -la gpurun_out

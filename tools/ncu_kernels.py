#!/usr/bin/env python
"""Distils ncu --set full captures into profiles/ncu_kernels.json, the per-kernel counters bench.py attaches to its
roofline entries (DRAM bytes per launch = `traffic`, FMA-pipe and issue-active percentages, instructions per ray).

    python tools/ncu_kernels.py c5=gpurun_out/full_c5.ncu-rep c4=gpurun_out/full_c4.ncu-rep c3=gpurun_out/full_c3.ncu-rep

The captures are of `bench.py --workload W --steps 1 --warmup 0 --spp 16` (one pass of 16 samples per pixel at 4K:
the launch size of the benched passes) and of tools/c3run.py; the first launch of each kernel class is taken."""
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CLASSES = [("trace_primary", ("ptbSpecTracePrimary", "tracePrimaryKernel")), ("shade_primary", ("shadeKernel<(bool)1", "shadeKernel<1")),
           ("trace_bounce", ("ptbSpecTraceQueue", "traceQueueKernel")), ("shade_bounce", ("shadeKernel<(bool)0", "shadeKernel<0")),
           ("intersect", ("ptbSpecIntersect", "intersectFastKernel"))]
WANT = {
    "gpu__time_duration.sum": "ms_under_ncu", "launch__registers_per_thread": "registers", "smsp__inst_executed.sum": "warp_instructions",
    "smsp__issue_active.avg.pct_of_peak_sustained_active": "issue_active_pct", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active": "fma_pipe_pct",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active": "alu_pipe_pct", "smsp__thread_inst_executed_per_inst_executed.ratio": "threads_per_instruction",
    "sm__warps_active.avg.pct_of_peak_sustained_active": "warps_active_pct", "dram__bytes_read.sum": "dram_read", "dram__bytes_write.sum": "dram_write",
    "lts__t_sector_hit_rate.pct": "l2_hit_pct",
}
UNIT = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "ms": 1.0, "us": 1e-3, "ns": 1e-6, "s": 1e3}


def main():
    out = {}
    path = os.path.join(ROOT, "profiles", "ncu_kernels.json")
    if os.path.exists(path):
        out = json.load(open(path))
    for arg in sys.argv[1:]:
        name, rep = arg.split("=")
        raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        rows = list(csv.reader(io.StringIO(raw)))
        hdr, units = rows[0], rows[1]
        entry = {}
        for r in rows[2:]:
            kernel = r[hdr.index("Kernel Name")]
            for cls, names in CLASSES:
                if cls in entry or not any(n in kernel for n in names):
                    continue
                e = {"kernel": kernel.split("(")[0], "source": os.path.basename(rep)}
                for metric, key in WANT.items():
                    if metric in hdr and r[hdr.index(metric)] not in ("", "n/a"):
                        v = float(r[hdr.index(metric)].replace(",", ""))
                        e[key] = v * UNIT.get(units[hdr.index(metric)], 1.0) if key.startswith("dram_") or key == "ms_under_ncu" else v
                if "dram_read" in e and "dram_write" in e:
                    e["dram_bytes_per_launch"] = e.pop("dram_read") + e.pop("dram_write")
                entry[cls] = e
                break
        out[name] = entry
    json.dump(out, open(path, "w"), indent=1, sort_keys=True)
    print(json.dumps(out, indent=1, sort_keys=True)[:1500])


if __name__ == "__main__":
    main()

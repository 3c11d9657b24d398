#!/usr/bin/env python
"""Summarise an .ncu-rep (ncu --set full capture) into a small text file for profiles/.

    python tools/ncu_summary.py gpurun_out/full_c4.ncu-rep > profiles/r01_full_c4.txt

Per captured launch: duration, issue/pipe utilisation, warp execution efficiency, occupancy,
registers, DRAM bytes, and the SASS opcode mix (share of executed warp instructions)."""
import collections
import csv
import io
import subprocess
import sys

RAW = [
    "gpu__time_duration.sum",
    "launch__grid_size", "launch__block_size", "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
    "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum",
    "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "smsp__thread_inst_executed_per_inst_executed.ratio",
    "smsp__thread_inst_executed_per_inst_executed.pct",
    "dram__bytes_read.sum", "dram__bytes_write.sum",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "smsp__inst_executed_op_shared_ld.sum", "smsp__inst_executed_op_shared_st.sum",
    "smsp__inst_executed_op_local_ld.sum", "smsp__inst_executed_op_local_st.sum",
]


def ncu(args):
    return subprocess.run(["ncu"] + args, capture_output=True, text=True).stdout


def main():
    rep = sys.argv[1]
    raw = list(csv.reader(io.StringIO(ncu(["-i", rep, "--page", "raw", "--csv"]))))
    hdr, units = raw[0], raw[1]
    print("# %s" % rep)
    for n, r in enumerate(raw[2:]):
        name = r[hdr.index("Kernel Name")].split("(")[0]
        print("\n## launch %d: %s" % (n, name))
        for w in RAW:
            if w in hdr:
                print("  %-70s %s %s" % (w, r[hdr.index(w)], units[hdr.index(w)]))
    src = list(csv.reader(io.StringIO(ncu(["-i", rep, "--page", "source", "--csv", "--print-source", "sass"]))))
    blocks, cur, h = [], None, None
    for r in src:
        if r and r[0] == "Kernel Name":
            cur = {"name": r[1].split("(")[0], "rows": []}
            blocks.append(cur)
        elif r and r[0] == "Address":
            h = r
        elif cur is not None and h and len(r) > 5:
            cur["rows"].append(r)
    if not h:
        return
    ie, so = h.index("Instructions Executed"), h.index("Source")
    for n, b in enumerate(blocks):
        ops = collections.Counter()
        tot = 0
        for r in b["rows"]:
            s = r[so].strip().split()
            op = s[1] if s[0].startswith("@") else s[0]
            ops[op.split(".")[0]] += int(r[ie])
            tot += int(r[ie])
        print("\n## launch %d: %s — SASS opcode mix (%d static, %d executed warp instructions)" % (n, b["name"], len(b["rows"]), tot))
        line = []
        for k, v in ops.most_common(24):
            line.append("%s %.1f%%" % (k, 100.0 * v / max(tot, 1)))
        print("  " + ", ".join(line))
        fp = sum(ops[k] for k in ("FFMA", "FMUL", "FADD", "FSETP", "FSEL", "MUFU", "FMNMX"))
        print("  FP32 arithmetic share (FFMA+FMUL+FADD+FSETP+FSEL+FMNMX+MUFU): %.1f%%" % (100.0 * fp / max(tot, 1)))


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""tools/ncu_summary.py <report.ncu-rep> -- compact text summary of an ncu --set full report:
key counters, stall reasons per issue, dynamic opcode histogram (from the source page)."""
import collections
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[0]
for row in rows[2:3]:
    d = dict(zip(hdr, row))
    print("kernel:", d.get("Kernel Name", "")[:90], " grid", d.get("Grid Size"), "block", d.get("Block Size"))
    keys = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
            "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
            "lts__throughput.avg.pct_of_peak_sustained_elapsed",
            "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
            "sm__throughput.avg.pct_of_peak_sustained_elapsed",
            "smsp__issue_active.avg.pct_of_peak_sustained_active",
            "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
            "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
            "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
            "sm__inst_executed_pipe_cbu.avg.pct_of_peak_sustained_active",
            "sm__warps_active.avg.pct_of_peak_sustained_active",
            "smsp__warps_eligible.avg.per_cycle_active", "launch__registers_per_thread",
            "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
            "smsp__inst_executed.sum", "sm__cycles_elapsed.max",
            "l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_ld.sum",
            "l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_st.sum",
            "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
            "l1tex__t_sector_pipe_lsu_mem_global_op_ld_hit_rate.pct", "lts__t_sector_hit_rate.pct"]
    for k in keys:
        if k in d:
            u = rows[1][hdr.index(k)]
            print(f"  {k:68s} {d[k]} {u}")
    print("  stalls per issue:", ", ".join(
        f"{k[34:-23]}={float(v):.2f}" for k, v in d.items()
        if k.startswith("smsp__average_warps_issue_stalled") and k.endswith("per_issue_active.ratio")
        and float(v) >= 0.05))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
hdr = rows[hi]
ci, si = hdr.index("Instructions Executed"), hdr.index("Source")
ops, tot = collections.Counter(), 0
for r in rows[hi + 1:]:
    if r and r[0] == "Address":
        break
    if len(r) <= ci or not r[ci]:
        continue
    toks = r[si].split()
    op = toks[1] if toks[0].startswith("@") else toks[0]
    op = op.split(".")[0]
    ops[op] += int(r[ci]); tot += int(r[ci])
print(f"  dynamic warp instructions (first launch): {tot}")
print("  " + "  ".join(f"{op} {100 * n / tot:.1f}%" for op, n in ops.most_common(14)))

#!/usr/bin/env python
"""tools/ncu_counters.py <workload>=<report.ncu-rep> ... -- per-launch counters of each workload's
dominant kernel out of `ncu --set full` reports, written to profiles/kernel_counters_r02.json
(read by bench.py for `roofline.traffic` and `roofline.issue_frac`), stamped with the commit the
captures were taken at."""
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
out_path = os.path.join(ROOT, "profiles", "kernel_counters_r02.json")
data = json.load(open(out_path)) if os.path.exists(out_path) else {}
commit = subprocess.run(["git", "-C", ROOT, "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()


def kernel_name(full):
    # "void <unnamed>::os_fused_persist_kernel<Plan<16384, ...>>(<unnamed>::OsParams)" -> short form
    t = full.replace("void ", "").replace("<unnamed>::", "")
    t = t[:t.rfind("(")] if "(" in t else t
    return t.replace("Plan<", "<").replace(" ", "")


for spec in sys.argv[1:]:
    name, rep = spec.split("=", 1)
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    # the launch with the longest duration is the dominant kernel of the capture
    best = max(rows[2:], key=lambda r: float(r[hdr.index("gpu__time_duration.sum")].replace(",", "")))
    d = dict(zip(hdr, best))
    u = dict(zip(hdr, units))

    def val(k):
        x = float(d[k].replace(",", ""))
        scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "inst": 1.0, "": 1.0}
        return x * scale.get(u[k], 1.0)

    data[name] = {
        "kernel": kernel_name(d["Kernel Name"]),
        "dram_bytes_per_launch": val("dram__bytes_read.sum") + val("dram__bytes_write.sum"),
        "warp_instructions_per_launch": val("smsp__inst_executed.sum"),
        "issue_active_pct": float(d["smsp__issue_active.avg.pct_of_peak_sustained_active"]),
        "duration_us_under_ncu": float(d["gpu__time_duration.sum"].replace(",", "")) *
        {"us": 1.0, "ms": 1e3, "ns": 1e-3, "s": 1e6}.get(u["gpu__time_duration.sum"], 1.0),
        "source": os.path.relpath(rep, ROOT) + " (ncu --set full --clock-control none, one launch)",
        "commit": commit,
    }
json.dump(data, open(out_path, "w"), indent=1)
print(json.dumps(data, indent=1))

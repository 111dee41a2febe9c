#!/usr/bin/env python
"""tools/ncu_phases.py <report.ncu-rep> -- where a kernel's warp-time goes, phase by phase.

Cuts the SASS of the first kernel in the report at its synchronisation instructions
(BAR, WARPSYNC, SYNCS, LDTM/STTM waits) and prints, per segment: warp instructions executed,
stall samples, and the three largest stall reasons.  Needs an ncu --set full report."""
import collections
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
min_share = float(sys.argv[2]) if len(sys.argv) > 2 else 0.5
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
hdr = rows[hi]
col = {n: i for i, n in enumerate(hdr)}
stall_cols = [n for n in hdr if n.startswith("stall_") and "Not Issued" not in n]
segs, cur = [], None


def new_seg(label):
    global cur
    cur = {"label": label, "inst": 0, "samples": 0, "stalls": collections.Counter(), "ops": collections.Counter()}
    segs.append(cur)


new_seg("start")
for r in rows[hi + 1:]:
    if r and r[0] == "Address":
        break
    if len(r) < len(hdr):
        continue
    text = r[col["Source"]]
    toks = text.split()
    op = toks[1] if toks[0].startswith("@") else toks[0]
    base = op.split(".")[0]
    n = int(r[col["Instructions Executed"]] or 0)
    s = int(r[col["# Samples"]] or 0)
    cur["inst"] += n
    cur["samples"] += s
    cur["ops"][base] += n
    for c in stall_cols:
        v = int(r[col[c]] or 0)
        if v:
            cur["stalls"][c[6:]] += v
    if base in ("BAR", "WARPSYNC", "SYNCS", "UTCBAR", "EXIT") or op.startswith("BAR"):
        new_seg(f"after {op} @{r[0][-5:]}")
tot_s = sum(s["samples"] for s in segs) or 1
tot_i = sum(s["inst"] for s in segs) or 1
print(f"total warp instructions {tot_i}, samples {tot_s}")
print(f"{'segment':34s} {'inst%':>6s} {'time%':>6s}  top stalls / top opcodes")
for s in segs:
    share = 100.0 * s["samples"] / tot_s
    if share < min_share:
        continue
    top = ", ".join(f"{k}={100.0 * v / max(s['samples'], 1):.0f}%" for k, v in s["stalls"].most_common(4))
    ops = " ".join(f"{k}:{v * 100 // max(s['inst'], 1)}" for k, v in s["ops"].most_common(4))
    print(f"{s['label'][:34]:34s} {100.0 * s['inst'] / tot_i:6.1f} {share:6.1f}  {top} | {ops}")

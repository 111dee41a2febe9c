#!/bin/bash
# tools/ab.sh <workload> <label=ENV...> ... -- A/B bench lines inside ONE gpurun call (boxes differ
# by a few per cent, so variants are only comparable when they ran on the same GPU), two rounds.
# usage: tools/ab.sh c4 "base=" "noedge=VELES_B200_NO_EDGE_FAST=1"
w=$1; shift
for round in 1 2; do
  for spec in "$@"; do
    label=${spec%%=*}; envs=${spec#*=}
    env $envs timeout 150 python bench.py --workload $w --no-cpu-baseline --no-e2e --steps 30 \
      > gpurun_out/ab_${w}_${label}.json 2> gpurun_out/ab_${w}_${label}.err
    python - <<PY
import json
try:
    d = json.load(open("gpurun_out/ab_${w}_${label}.json"))
    r = d.get("roofline") or {}
    print("$w $label round $round: ms_per_step %.4f value %.1f kernel_ms %.4f"
          % (d["ms_per_step"], d["value"], r.get("kernel_ms_per_launch", 0)))
except Exception as e:
    print("$w $label: failed", e)
PY
  done
done

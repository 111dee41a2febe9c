#!/bin/bash
# tools/profile.sh <tag> <kernel-regex> <bench args...> -- plain run, ncu launch list, ncu full
# capture of the top kernel (B200_PROFILING.md recipe).  Run under gpurun; outputs in gpurun_out/.
set -u
TAG=$1; K=$2; shift 2
CMD="python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline --min-warm-s 0 $*"
$CMD > gpurun_out/plain_$TAG.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv \
    --log-file gpurun_out/launches_$TAG.csv $CMD > gpurun_out/ncu_list_$TAG.log 2>&1
$CMD > gpurun_out/plain2_$TAG.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$K -s 3 -c 2 \
    -f -o gpurun_out/prof_$TAG $CMD > gpurun_out/ncu_full_$TAG.log 2>&1
tail -1 gpurun_out/ncu_full_$TAG.log

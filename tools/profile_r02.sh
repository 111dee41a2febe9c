#!/bin/bash
# tools/profile_r02.sh -- the round-2 ncu evidence, one gpurun call on one GPU.
# Every ncu run follows a plain run of the same command line (B200_PROFILING.md).
set -u
B="python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline --no-configs --min-warm-s 0 --no-profile"
out=gpurun_out
# launch list of the headline command (cold-cache, serialised: compare shares)
timeout 120 $B > $out/plain_c4.log 2>&1 && \
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/launches_c4_r02.csv $B > $out/ncu_list_c4.log 2>&1
for w in c4 c3 c5; do
  case $w in c4) K=persist;; c3) K=os_fused_kernel;; c5) K=os_fused_real;; esac
  timeout 120 $B --workload $w > $out/plain_$w.log 2>&1 && \
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:$K -s 4 -c 1 -o $out/prof_${w}_r02 $B --workload $w > $out/ncu_full_$w.log 2>&1
  tail -1 $out/ncu_full_$w.log
done
# forced direct kernel on the C3 shape
timeout 120 $B --workload c3 --algorithm 0 > $out/plain_c3d.log 2>&1 && \
timeout 300 ncu --set full --clock-control none --import-source on -k regex:direct_kernel -s 4 -c 1 -o $out/prof_c3direct_r02 $B --workload c3 --algorithm 0 > $out/ncu_full_c3d.log 2>&1
tail -1 $out/ncu_full_c3d.log

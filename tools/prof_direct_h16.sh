#!/bin/bash
# tools/prof_direct_h16.sh -- the direct kernel in its HBM-bound regime (4096 x 16384, h = 16): bench line,
# then one ncu --set full capture of the same command (one GPU).
B="python bench.py --workload c3h16 --algorithm 0 --steps 3 --warmup 3 --no-e2e --no-cpu-baseline --no-configs --min-warm-s 0 --no-profile"
out=gpurun_out
timeout 120 python bench.py --workload c3h16 --algorithm 0 --no-e2e --no-cpu-baseline --steps 30 > $out/bench_c3h16.json 2> $out/bench_c3h16.err
tail -c 600 $out/bench_c3h16.json
timeout 120 $B > $out/plain_c3h16.log 2>&1 && \
timeout 300 ncu --set full --clock-control none --import-source on -k regex:direct_kernel -s 4 -c 1 -f -o $out/prof_c3h16_r02 $B > $out/ncu_full_c3h16.log 2>&1
tail -1 $out/ncu_full_c3h16.log

#!/bin/bash
# tools/refresh_bench.sh -- every bench line of profiles/bench_*_r01.json in one go (run under gpurun).
set -u
python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err
for w in c1 c2 c3 c4 c5 peaks minmax wavelet; do
  python bench.py --workload $w > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err || echo "bench $w failed"
done
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err
for w in default c1 c2 c3 c4 c5 peaks minmax wavelet reference; do echo "$w: $(cut -c1-150 gpurun_out/bench_$w.json)"; done

#!/bin/bash
# tools/scale_run.sh N -- bench lines of the partitioned configs on N GPUs of one box (run under
# gpurun --gpus N); results in gpurun_out/scaleN_<workload>.json
N=$1; port=29600
for w in c3 c4 c5; do
  port=$((port + 1))
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $port \
    bench.py --gpus $N --workload $w --steps 20 --warmup 3 --no-cpu-baseline 2> gpurun_out/scale${N}_$w.err | tail -1 > gpurun_out/scale${N}_$w.json
  echo "$w: $(cut -c1-200 gpurun_out/scale${N}_$w.json)"
done

#!/bin/bash
# multi-GPU bench lines: args = N (ranks), workload, extra flags
N=${1:-2}; WL=${2:-c2}; shift 2
mkdir -p gpurun_out
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
  bench.py --gpus $N --workload $WL --steps 5 --warmup 3 "$@" > gpurun_out/bench_${WL}_n$N.json 2> gpurun_out/bench_${WL}_n$N.err
echo "exit $?"; tail -5 gpurun_out/bench_${WL}_n$N.err; cut -c1-1200 gpurun_out/bench_${WL}_n$N.json

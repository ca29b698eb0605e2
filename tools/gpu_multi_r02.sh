#!/bin/bash
# Multi-GPU pass (gpurun --gpus N): 2-rank NCCL parity test, host-feed ceiling at 1/2/4/N ranks, bench at N ranks.
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
timeout 900 python -m pytest tests/test_multi_gpu_nccl.py -m gpu -x -q > gpurun_out/pytest_nccl.log 2>&1; echo "nccl pytest exit $?" >> gpurun_out/pytest_nccl.log
tail -5 gpurun_out/pytest_nccl.log
: > gpurun_out/h2d_ceiling.jsonl
for n in 1 2 4 8; do
  if [ $n -le $NG ]; then
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500+n)) tools/h2d_ceiling.py 2>/dev/null | grep '^{' >> gpurun_out/h2d_ceiling.jsonl
  fi
done
cat gpurun_out/h2d_ceiling.jsonl
if [ "$1" != "nobench" ]; then
  timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus $NG --secondary ${2:-c2,c3} > gpurun_out/bench_n$NG.json 2> gpurun_out/bench_n$NG.err
  echo "bench exit $?"; tail -3 gpurun_out/bench_n$NG.err; cut -c1-600 gpurun_out/bench_n$NG.json
fi

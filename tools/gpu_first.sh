#!/bin/bash
# first GPU pass of a session: parity tests, then a timing sweep over the config shapes
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit --format=csv > gpurun_out/smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
for cfg in "2000000 10 64" "1000000 16 128" "1000000 9 128" "65536 32 256" "2000000 7 64" "4000000 11 10"; do
  set -- $cfg
  timeout 300 python tools/quick_bench.py --N $1 --D $2 --K $3 >> gpurun_out/quick_bench.jsonl 2>> gpurun_out/quick_bench.err
done
cat gpurun_out/quick_bench.jsonl | cut -c1-400

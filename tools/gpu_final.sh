#!/bin/bash
# the driver's round-end sequence: reference arm, own arm (default flags), smoke
mkdir -p gpurun_out
timeout 900 python bench.py --impl reference --gpus 1 --steps 3 --warmup 1 > gpurun_out/final_ref.json 2> gpurun_out/final_ref.err; echo "ref exit $?"; cut -c1-400 gpurun_out/final_ref.json
timeout 900 python bench.py > gpurun_out/final_bench.json 2> gpurun_out/final_bench.err; echo "bench exit $?"; tail -n 3 gpurun_out/final_bench.err
python __graft_entry__.py smoke 2>&1 | tail -3

#!/bin/bash
# GPU pass for the cooperative (team) kernels: parity tests, then timing sweeps of D = 17..32 shapes
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_team_kernels.py -m gpu -x -q > gpurun_out/pytest_team.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_team.log
tail -25 gpurun_out/pytest_team.log
for cfg in "32 256 262144" "32 64 262144" "24 128 262144" "21 64 262144" "28 128 262144"; do
  set -- $cfg
  timeout 300 python tools/quick_bench.py --D $1 --K $2 --N $3 --variants '[{"rc":3},{"rc":3,"pt":8},{"rc":4}]' 2>&1 | tail -4
done > gpurun_out/quick_team.jsonl
cat gpurun_out/quick_team.jsonl | cut -c1-420

#!/bin/bash
# GPU pass: parity tests, smoke, bench (c2), ncu launch list + full capture of the pair kernels
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke.log; tail -3 gpurun_out/smoke.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; echo "bench exit $?"; tail -2 gpurun_out/bench_c2.err
cat gpurun_out/bench_c2.json | cut -c1-3000
# profiling runs use a smaller shard (ncu replays kernels)
PROF="python bench.py --workload small --steps 2 --warmup 3 --no-e2e --no-cpu"
$PROF > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches.csv $PROF > gpurun_out/ncu_launches.log 2>&1
$PROF > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:pair_kernel -s 6 -c 2 -o gpurun_out/prof_pair -f $PROF > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/ncu_full.log

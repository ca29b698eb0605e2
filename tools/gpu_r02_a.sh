#!/bin/bash
# Round-2 GPU pass A: parity tests, smoke, default bench (c5 headline + c2/c3/c4), reference arm,
# ncu launch list of the default command's shape, ncu --set full of the c5 pair kernels and the c4 team kernels.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -8 gpurun_out/pytest_gpu.log
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke.log; tail -3 gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench exit $?"; tail -2 gpurun_out/bench_default.err
cut -c1-1500 gpurun_out/bench_default.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "ref exit $?"
cut -c1-600 gpurun_out/bench_reference.json
# launch list (c5 at 1/8 of the shard: ncu serialises and replays)
PROF="python bench.py --workload c5 --points 1562500 --secondary none --steps 2 --warmup 3 --no-e2e --no-cpu"
$PROF > gpurun_out/plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/launches_c5.csv $PROF > gpurun_out/ncu_launches.log 2>&1
# full captures: c5 (one STATS + one LOGLIK launch), c4 team kernels
PROF="python bench.py --workload c5 --points 1562500 --secondary none --steps 2 --warmup 3 --no-e2e --no-cpu --no-sgd"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:pair_kernel -s 4 -c 2 -o gpurun_out/prof_pair_c5 -f $PROF > gpurun_out/ncu_full_c5.log 2>&1
tail -2 gpurun_out/ncu_full_c5.log
PROF="python bench.py --workload c4 --points 89856 --secondary none --steps 2 --warmup 3 --no-e2e --no-cpu --no-sgd"
$PROF > gpurun_out/plain_c4.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:team_ -s 6 -c 3 -o gpurun_out/prof_team_c4 -f $PROF > gpurun_out/ncu_full_c4.log 2>&1
tail -2 gpurun_out/ncu_full_c4.log
ls -la gpurun_out/

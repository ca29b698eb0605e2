#!/bin/bash
# Final round-2 GPU pass: parity suite, smoke, default bench, reference arm, ncu launch list + full captures (c5 at 1/8 shard).
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke exit $?" >> gpurun_out/smoke.log; tail -4 gpurun_out/smoke.log
timeout 1200 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench exit $?"; tail -3 gpurun_out/bench_default.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "ref exit $?"
PROF="python bench.py --workload c5 --points 1562500 --secondary none --steps 2 --warmup 3 --no-e2e --no-cpu --no-sgd"
$PROF > gpurun_out/plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'dgmm|pair_kernel|prepare_kernel|finalize|mstep|pack_noise|full_kernel' -c 200 --csv --log-file gpurun_out/launches_c5.csv $PROF > gpurun_out/ncu_launches.log 2>&1
$PROF > gpurun_out/plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:pair_kernel -s 4 -c 2 -o gpurun_out/prof_pair_c5 -f $PROF > gpurun_out/ncu_full_c5.log 2>&1
tail -2 gpurun_out/ncu_full_c5.log
PROF="python bench.py --workload c4 --points 89856 --secondary none --steps 2 --warmup 3 --no-e2e --no-cpu --no-sgd"
$PROF > gpurun_out/plain_c4.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'team_loglik|team_stats' -s 4 -c 2 -o gpurun_out/prof_team_c4 -f $PROF > gpurun_out/ncu_full_c4.log 2>&1
tail -2 gpurun_out/ncu_full_c4.log

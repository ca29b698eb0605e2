#!/bin/bash
# ncu evidence for the bench line: launch list (gpu__time_duration) of the bench command and one
# --set full capture of the last STATS launch and the first LOGLIK launch, both on the FULL c2 shard;
# then the single-GPU bench lines of the other named shapes (c3, c5 shards).
mkdir -p gpurun_out
PROF="python bench.py --workload c2 --steps 2 --warmup 3 --no-e2e --no-cpu"
$PROF > gpurun_out/plain_c2.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/launches_c2.csv $PROF > gpurun_out/ncu_launches_c2.log 2>&1
$PROF > gpurun_out/plain2_c2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:pair_kernel -s 4 -c 2 -o gpurun_out/prof_pair_c2 -f $PROF > gpurun_out/ncu_full_c2.log 2>&1
tail -3 gpurun_out/ncu_full_c2.log
for wl in c5 c3; do
  timeout 600 python bench.py --workload $wl --steps 5 --warmup 3 --no-e2e --no-cpu > gpurun_out/bench_$wl.json 2> gpurun_out/bench_$wl.err; echo "bench $wl exit $?"
  cut -c1-600 gpurun_out/bench_$wl.json; tail -2 gpurun_out/bench_$wl.err
done

#!/bin/bash
# ncu --set full capture of one STATS and one LOGLIK launch on the full c5 and c3 shards (DRAM traffic
# per launch for roofline.traffic of those bench lines); each only after the plain command exited 0.
mkdir -p gpurun_out
for wl in c5 c3; do
  PROF="python bench.py --workload $wl --steps 2 --warmup 3 --no-e2e --no-cpu --no-sgd"
  $PROF > gpurun_out/plain_$wl.log 2>&1 &&
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:pair_kernel -s 4 -c 2 -o gpurun_out/prof_pair_$wl -f $PROF > gpurun_out/ncu_full_$wl.log 2>&1
  tail -2 gpurun_out/ncu_full_$wl.log
  ncu -i gpurun_out/prof_pair_$wl.ncu-rep --page raw --csv > gpurun_out/prof_pair_$wl.csv 2>/dev/null
  ls -la gpurun_out/prof_pair_$wl.*
done

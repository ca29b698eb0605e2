#!/bin/bash
# Round-2 GPU pass C: full parity suite, ncu launch lists of the SGD step (c2 and c5 shapes), compute-sanitizer.
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -12 gpurun_out/pytest_gpu.log
for wl in c2 c5; do
  PROF="python bench.py --workload $wl --points 1048576 --secondary none --steps 2 --warmup 3 --no-e2e --no-cpu"
  $PROF > gpurun_out/plain_sgd_$wl.log 2>&1 &&
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'dgmm|pair_kernel|prepare_kernel|finalize|sgd_|gather_points|mstep' -c 400 --csv --log-file gpurun_out/launches_sgd_$wl.csv $PROF > gpurun_out/ncu_launches_sgd_$wl.log 2>&1
  tail -2 gpurun_out/ncu_launches_sgd_$wl.log
done
bash tools/gpu_sanitizer.sh

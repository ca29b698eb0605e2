#!/bin/bash
# ncu capture of one stats launch and one loglik launch (args: N D K)
N=${1:-2000000}; D=${2:-10}; K=${3:-64}; TAG=${4:-d${D}k${K}}
mkdir -p gpurun_out
CMD="python tools/quick_bench.py --N $N --D $D --K $K"
$CMD > gpurun_out/plain_$TAG.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:pair_kernel -s 3 -c 3 -o gpurun_out/prof_$TAG -f $CMD > gpurun_out/ncu_$TAG.log 2>&1
tail -2 gpurun_out/ncu_$TAG.log

#!/bin/bash
# A/B of the shipped library against an experimental build (DGMM_LIB) on a list of "D K N" shapes.
mkdir -p gpurun_out
for cfg in "$@"; do
  set -- $cfg
  for lib in "" "density_deconvolution_b200/libdgmm_b200_exp.so"; do
    echo "== D=$1 K=$2 N=$3 lib=${lib:-shipped}"
    DGMM_LIB=$lib timeout 300 python tools/quick_bench.py --D $1 --K $2 --N $3 2>&1 | tail -1 | cut -c1-330
  done
done 2>&1 | tee gpurun_out/ab_lib.txt

#!/bin/bash
# bench-only refresh on the final build: default line + reference arm + the Gaia workload
mkdir -p gpurun_out
timeout 1200 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench exit $?"; tail -3 gpurun_out/bench_default.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; echo "ref exit $?"
timeout 600 python bench.py --workload gaia --secondary none --steps 5 > gpurun_out/bench_gaia.json 2> gpurun_out/bench_gaia.err; echo "gaia exit $?"

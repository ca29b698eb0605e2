#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_kmeans.py -m gpu -x -q > gpurun_out/pytest_kmeans.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_kmeans.log
tail -15 gpurun_out/pytest_kmeans.log
timeout 600 python tools/kmeans_bench.py > gpurun_out/kmeans_bench.jsonl 2> gpurun_out/kmeans_bench.err; cat gpurun_out/kmeans_bench.jsonl; tail -3 gpurun_out/kmeans_bench.err

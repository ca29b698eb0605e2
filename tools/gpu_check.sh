#!/bin/bash
# parity tests + short bench lines (no CPU arm, no e2e) for c2, c5, c3
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -8 gpurun_out/pytest_gpu.log
for wl in c2 c5 c3; do
  timeout 600 python bench.py --workload $wl --steps 5 --warmup 3 --no-e2e --no-cpu > gpurun_out/chk_$wl.json 2> gpurun_out/chk_$wl.err; echo "bench $wl exit $?"
  python - <<PY
import json
for l in open("gpurun_out/chk_$wl.json"):
    if l.startswith("{"):
        d=json.loads(l); r=d["roofline"]
        print("$wl", "EM ms %.2f frac %.3f | loglik ms %.2f frac %.3f" % (r["launch_ms"], r["frac"], r["loglik_kernel"]["launch_ms"], r["loglik_kernel"]["frac"]), "| sgd", d.get("sgd"))
PY
  tail -n 2 gpurun_out/chk_$wl.err
done

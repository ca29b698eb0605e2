#!/bin/bash
# quick GPU pass: parity tests of the ops, then the timing sweep over the config shapes
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_ops.py -m gpu -x -q > gpurun_out/pytest_ops.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_ops.log
tail -4 gpurun_out/pytest_ops.log
rm -f gpurun_out/quick_bench.jsonl
for cfg in "2000000 10 64" "1000000 16 128" "1000000 9 128" "2000000 7 64" "4000000 11 10"; do
  set -- $cfg
  timeout 300 python tools/quick_bench.py --N $1 --D $2 --K $3 >> gpurun_out/quick_bench.jsonl 2>> gpurun_out/quick_bench.err
done
python - <<'PY'
import json
for l in open('gpurun_out/quick_bench.jsonl'):
    d = json.loads(l)
    if 'N' in d:
        print("D=%d K=%d  loglik %.2f ms frac %.3f | stats %.2f ms frac %.3f  (pairs/s %.3g / %.3g)" % (d['D'], d['K'], d['t_loglik_ms'], d['frac_ll_nominal'], d['t_stats_ms'], d['frac_em_nominal'], d['loglik_pairs_per_s'], d['stats_pairs_per_s']))
PY
tail -3 gpurun_out/quick_bench.err

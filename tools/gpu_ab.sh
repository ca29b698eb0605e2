#!/bin/bash
# A/B of the main (rc 0) and alternative (rc 1) placements on the config shapes
CFGS=("2000000 10 64" "1000000 9 128" "1000000 11 128" "1000000 12 128" "2000000 7 64" "2000000 10 32")
mkdir -p gpurun_out; rm -f gpurun_out/ab.jsonl
timeout 900 python -m pytest tests/test_gpu_ops.py -m gpu -x -q > gpurun_out/pytest_ops.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_ops.log
tail -3 gpurun_out/pytest_ops.log
for cfg in "${CFGS[@]}"; do
  set -- $cfg
  timeout 300 python tools/quick_bench.py --N $1 --D $2 --K $3 --variants '[{"rc":0},{"rc":1},{"rc":3}]' >> gpurun_out/ab.jsonl 2>> gpurun_out/ab.err
done
python - <<'PY'
import json
for l in open('gpurun_out/ab.jsonl'):
    d = json.loads(l)
    if 'N' in d:
        ps = d['plan_stats']
        print("D=%d K=%d rc=%s accp=%d TPB=%d grid=%d | loglik %.2f ms frac %.3f | stats %.2f ms frac %.3f" % (d['D'], d['K'], d['variant'].get('rc'), ps['accp'], ps['TPB'], ps['grid'], d['t_loglik_ms'], d['frac_ll_nominal'], d['t_stats_ms'], d['frac_em_nominal']))
    elif 'error' in d: print(d)
PY
tail -3 gpurun_out/ab.err

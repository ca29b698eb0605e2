#!/bin/bash
# A/B of the log-likelihood kernel placements: rc 0 = (128 threads, 3 blocks/SM), rc 1 = (128, 4)
mkdir -p gpurun_out; rm -f gpurun_out/ab_ll.jsonl
for cfg in "4000000 10 64" "2000000 9 128" "2000000 11 128" "2000000 12 128" "4000000 7 64" "4000000 10 32"; do
  set -- $cfg
  timeout 300 python tools/quick_bench.py --N $1 --D $2 --K $3 --variants '[{"rc":0},{"rc":1}]' >> gpurun_out/ab_ll.jsonl 2>> gpurun_out/ab_ll.err
done
python - <<'PY'
import json
for l in open('gpurun_out/ab_ll.jsonl'):
    d = json.loads(l)
    if 'N' in d:
        pl = d['plan_loglik']
        print("D=%d K=%d rc=%s TPB=%d grid=%d | loglik %.2f ms frac %.3f | stats %.2f ms frac %.3f" % (d['D'], d['K'], d['variant'].get('rc'), pl['TPB'], pl['grid'], d['t_loglik_ms'], d['frac_ll_nominal'], d['t_stats_ms'], d['frac_em_nominal']))
    elif 'error' in d: print(d)
PY
tail -n 3 gpurun_out/ab_ll.err

#!/bin/bash
# tile-size sweep (points per tile / ring depth) for the pair kernels
mkdir -p gpurun_out; rm -f gpurun_out/pt.jsonl
V='[{}, {"pt":32}, {"pt":32,"nst":2}, {"pt":64,"nst":2}, {"pt":8}]'
for cfg in "4000000 10 64" "2000000 9 128" "1000000 16 128" "2000000 12 128" "4000000 7 64"; do
  set -- $cfg
  timeout 300 python tools/quick_bench.py --N $1 --D $2 --K $3 --variants "$V" >> gpurun_out/pt.jsonl 2>> gpurun_out/pt.err
done
python - <<'PY'
import json
for l in open('gpurun_out/pt.jsonl'):
    d = json.loads(l)
    if 'N' in d:
        pl = d['plan_loglik']; ps = d['plan_stats']
        print("D=%d K=%d %-22s LL: PT=%d NST=%d grid=%d %.2f ms frac %.3f | ST: PT=%d grid=%d %.2f ms frac %.3f" % (d['D'], d['K'], json.dumps(d['variant']), pl['PT'], pl['NST'], pl['grid'], d['t_loglik_ms'], d['frac_ll_nominal'], ps['PT'], ps['grid'], d['t_stats_ms'], d['frac_em_nominal']))
    elif 'error' in d: print(d)
PY
tail -n 3 gpurun_out/pt.err

#!/bin/bash
# parity suite + minibatch-size timing with the automatic tile size against the fixed default one
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
for cfg in "10 64 32" "9 128 16" "16 128 8" "7 512 8" "10 128 16" "8 64 32"; do
  set -- $cfg
  python tools/quick_bench.py --D $1 --K $2 --N 65536 --variants "[{},{\"pt\":$3}]" | grep variant
done

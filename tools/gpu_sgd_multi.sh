#!/bin/bash
# SGD step at N ranks (peer-memory reduction vs NCCL all-reduce) on small resident shards: c2 and c5 shapes.
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
timeout 600 python -m pytest tests/test_multi_gpu_nccl.py -m gpu -x -q 2>&1 | tail -2
for wl in c2 c5; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port 29701 bench.py --gpus $NG --workload $wl --points 2097152 --secondary none --steps 3 --no-e2e --no-cpu 2> gpurun_out/sgd_multi_$wl.err | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print(json.dumps(dict(workload='$wl', n_gpus=d['n_gpus'], ms_per_em_step=d['ms_per_step'], sgd=d['sgd'])))
" | tee -a gpurun_out/sgd_multi.jsonl
done

"""Host-memory feed ceiling of the end-to-end (per-call streaming) arm: every rank copies a pinned 1 GiB buffer
to its GPU ``reps`` times; rank 0 prints per-rank and aggregate GB/s at this world size.  Run under torchrun at
1 / 2 / 4 / 8 ranks on one box (``tools/gpu_multi_r02.sh``)."""
import json
import os
import time

import torch
import torch.distributed as dist


def main():
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    n = 1 << 28
    h = torch.empty(n, dtype=torch.float32, pin_memory=True)
    h.fill_(1.0)
    d = torch.empty(n, dtype=torch.float32, device=dev)
    d.copy_(h, non_blocking=True)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    reps = 8
    t0 = time.perf_counter()
    for _ in range(reps):
        d.copy_(h, non_blocking=True)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    gbs = torch.tensor([n * 4 * reps / dt / 1e9], dtype=torch.float64, device=dev)
    if world > 1:
        allv = [torch.zeros_like(gbs) for _ in range(world)]
        dist.all_gather(allv, gbs)
        vals = [v.item() for v in allv]
    else:
        vals = [gbs.item()]
    if rank == 0:
        print(json.dumps(dict(ranks=world, per_rank_gbs=[round(v, 2) for v in vals], aggregate_gbs=round(sum(vals), 2),
                              bytes_per_copy=n * 4, reps=reps)))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

"""Cross-GPU exchange of the hot path: ONE all-reduce(SUM) of the packed statistics buffer per
EM iteration / SGD step.  One process per GPU; points are sharded contiguously
by rank; parameters are replicated and every rank applies the identical update, so no
broadcast is needed.  ``torch.distributed`` (NCCL on GPUs, gloo in the CPU tests) is the
plumbing."""
import torch
import torch.distributed as dist


def is_distributed():
    return dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1


def world_size():
    return dist.get_world_size() if is_distributed() else 1


def rank():
    return dist.get_rank() if is_distributed() else 0


def shard_bounds(n, r=None, ws=None):
    """Contiguous ``[lo, hi)`` slice of ``n`` points owned by rank ``r``."""
    r = rank() if r is None else r
    ws = world_size() if ws is None else ws
    base, rem = divmod(n, ws)
    lo = r * base + min(r, rem)
    return lo, lo + base + (1 if r < rem else 0)


def allreduce_stats(stats):
    """In-place SUM of the packed statistics across ranks (no-op for one process).  Every slot of the buffer is
    additive; the ``(ll_hi, ll_lo)`` float pair that carries the DOUBLE ``sum_i ll_i`` is reduced as a double of its
    own (a float32 sum of the ``hi`` parts would round the total to ~1e-7 relative, which at ``|ll| ~ 1e9`` is hundreds
    of units - enough to fake or hide the ``|ll - prev| < tol`` convergence test) and split again afterwards."""
    if is_distributed():
        tail = (stats[-4].double() + stats[-3].double()).reshape(1)
        dist.all_reduce(stats, op=dist.ReduceOp.SUM)
        dist.all_reduce(tail, op=dist.ReduceOp.SUM)
        hi = tail.to(torch.float32)
        stats[-4:-3] = hi
        stats[-3:-2] = torch.where(torch.isfinite(hi), (tail - hi.double()).to(torch.float32), torch.zeros_like(hi))
    return stats


def allreduce_scalar_sum(value, device):
    if not is_distributed():
        return value
    t = torch.tensor([value], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return t.item()

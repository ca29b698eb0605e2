"""Initialisation helpers (reference: deconv/gmm/util.py) on the fused k-means kernel
(``dgmm_kmeans_assign_f32`` / ``dgmm_kmeans_update_f32``): one pass over the
points yields labels, per-centroid counts, coordinate sums and the total distance, so neither the
reference's ``[B,K,D]`` difference / mask tensors nor its Python loop over centroids exist here.
Under ``torch.distributed`` every rank passes its own shard and the packed ``K(1+D)+2`` floats are
all-reduced, so all ranks end with identical centroids (ranks must make the same number of
batches).  There is no torch fallback: the tensors must live on the GPU."""
import torch

from . import ops as _real_ops

from . import dist as _dist


def _allreduce(t, op="sum"):
    if _dist.is_distributed():
        import torch.distributed as td
        td.all_reduce(t, op={"sum": td.ReduceOp.SUM, "min": td.ReduceOp.MIN, "max": td.ReduceOp.MAX}[op])
    return t


def _from_rank0(t):
    if _dist.is_distributed():
        import torch.distributed as td
        td.broadcast(t, src=0)
    return t


def k_means(X, k, max_iters=50, tol=1e-9, device=None, generator=None, ops=None, return_labels=False):
    """Lloyd's k-means from uniformly drawn centroids in the bounding box
    (deconv/gmm/util.py:4-33).  Returns ``(one-hot resp [N,k] float, centroids [k,D])``.
    Deliberate fix: an empty cluster keeps its previous centroid
    instead of becoming NaN (util.py:24), which made ``DeconvGMM.fit`` fail at iteration 0."""
    ops = ops or _real_ops
    n, d = X.shape
    lo = _allreduce(X.min(dim=0)[0], "min")
    hi = _allreduce(X.max(dim=0)[0], "max")
    centroids = (_from_rank0(torch.rand(k, d, device=X.device, generator=generator)) * (hi - lo) + lo).contiguous()
    prev = None
    for _ in range(max_iters):
        # labels and distances are taken against the centroids the pass started from (util.py:20-21)
        labels, stats = ops.kmeans_assign(X, centroids, want_labels=True)
        _allreduce(stats)
        ops.kmeans_update("lloyd", stats, centroids)
        total = ops.kmeans_total_distance(stats)
        if prev is not None and abs(total - prev) < tol:
            break
        prev = total
    if return_labels:
        return labels, centroids
    resp = torch.zeros(n, k, device=X.device, dtype=X.dtype)
    resp[torch.arange(n, device=X.device), labels.long()] = 1.0
    return resp, centroids


def minibatch_k_means(loader, k, max_iters=50, tol=1e-3, device=None, ops=None):
    """Sculley's minibatch k-means with per-centre learning rates ``1/count``
    (deconv/gmm/util.py:36-75).  ``loader`` yields ``(X_b, ...)`` batches.  Returns
    ``(counts [k], centroids [k,D])``."""
    ops = ops or _real_ops
    first = next(iter(loader))[0]
    centroids = _from_rank0(first[:k].to(device=device, dtype=torch.float32).clone()).contiguous()
    counts = torch.ones(k, device=centroids.device)
    prev_norm = torch.tensor(0.0, device=centroids.device)
    stats = None
    for _ in range(max_iters):
        for batch in loader:
            X = batch[0].to(centroids.device)
            # counts and the shift sum_i (x_i - c_label) are both taken against the centroids at the
            # start of the batch, as in the reference (util.py:53-64)
            _, stats = ops.kmeans_assign(X, centroids, out_stats=stats)
            _allreduce(stats)
            ops.kmeans_update("minibatch", stats, centroids, counts)
        norm = torch.norm(centroids, dim=0).sum()
        if torch.abs(norm - prev_norm) < tol:
            break
        prev_norm = norm
    return counts, centroids

"""Initialisation helpers (reference: deconv/gmm/util.py).  Not on the hot path: plain
torch ops on the device, written without the reference's ``[B,K,D]`` mask tensors.  Under
``torch.distributed`` every rank passes its own shard and the counts / sums are all-reduced,
so all ranks end with identical centroids (ranks must make the same number of batches)."""
import torch

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


def k_means(X, k, max_iters=50, tol=1e-9, device=None, generator=None):
    """Lloyd's k-means from uniformly drawn centroids in the bounding box
    (deconv/gmm/util.py:4-33).  Returns ``(one-hot resp [N,k] float, centroids [k,D])``.
    Deliberate fix: an empty cluster keeps its previous centroid
    instead of becoming NaN (util.py:24), which made ``DeconvGMM.fit`` fail at iteration 0."""
    n, d = X.shape
    lo = _allreduce(X.min(dim=0)[0], "min")
    hi = _allreduce(X.max(dim=0)[0], "max")
    centroids = _from_rank0(torch.rand(k, d, device=X.device, generator=generator)) * (hi - lo) + lo
    prev = None
    for _ in range(max_iters):
        dist = torch.cdist(X, centroids)
        labels = dist.argmin(dim=1)
        counts = _allreduce(torch.bincount(labels, minlength=k).to(X.dtype))
        sums = _allreduce(torch.zeros(k, d, device=X.device, dtype=X.dtype).index_add_(0, labels, X))
        centroids = torch.where(counts[:, None] > 0, sums / counts.clamp(min=1)[:, None], centroids)
        total = _allreduce(dist.gather(1, labels[:, None]).sum())
        if prev is not None and torch.abs(total - prev) < tol:
            break
        prev = total
    resp = torch.zeros(n, k, device=X.device, dtype=X.dtype)
    resp[torch.arange(n, device=X.device), labels] = 1.0
    return resp, centroids


def minibatch_k_means(loader, k, max_iters=50, tol=1e-3, device=None):
    """Sculley's minibatch k-means with per-centre learning rates ``1/count``
    (deconv/gmm/util.py:36-75).  ``loader`` yields ``(X_b, ...)`` batches.  Returns
    ``(counts [k], centroids [k,D])``."""
    first = next(iter(loader))[0]
    centroids = _from_rank0(first[:k].to(device).clone())
    counts = torch.ones(k, device=centroids.device)
    prev_norm = torch.tensor(0.0, device=centroids.device)
    for _ in range(max_iters):
        for batch in loader:
            X = batch[0].to(centroids.device)
            labels = torch.cdist(X, centroids).argmin(dim=1)
            counts = counts + _allreduce(torch.bincount(labels, minlength=k).to(counts.dtype))
            eta = 1.0 / counts
            # sum over the batch of eta_c (x - centroid_c) for the points assigned to c,
            # all taken against the centroids at the start of the batch as in the reference
            shift = _allreduce(torch.zeros_like(centroids).index_add_(0, labels, X - centroids[labels]))
            centroids = centroids + eta[:, None] * shift
        norm = torch.norm(centroids, dim=0).sum()
        if torch.abs(norm - prev_norm) < tol:
            break
        prev_norm = norm
    return counts, centroids

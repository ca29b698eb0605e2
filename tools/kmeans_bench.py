"""Times the fused k-means assignment kernel (dgmm_kmeans_assign_f32) on the config shapes, next to
the torch composition it replaced (cdist + argmin + bincount + index_add) on the same GPU.
One JSON line per shape."""
import json
import sys
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from density_deconvolution_b200 import ops

dev = torch.device("cuda", 0)
peak = ops.fp32_probe(dev)
for (N, D, K) in ((10_000_000, 10, 64), (12_500_000, 9, 128), (12_500_000, 16, 128), (2_000_000, 32, 256)):
    g = torch.Generator(device=dev).manual_seed(1)
    C = 5.0 * torch.randn(K, D, device=dev, generator=g)
    X = C[torch.randint(0, K, (N,), device=dev, generator=g)] + torch.randn(N, D, device=dev, generator=g)
    stats = torch.empty(ops.kmeans_stats_len(D, K), device=dev)

    def ours():
        ops.kmeans_assign(X, C, out_stats=stats)

    def torch_way(chunk=1 << 20):
        counts = torch.zeros(K, device=dev)
        sums = torch.zeros(K, D, device=dev)
        for lo in range(0, N, chunk):
            x = X[lo:lo + chunk]
            lab = torch.cdist(x, C).argmin(dim=1)
            counts += torch.bincount(lab, minlength=K)
            sums.index_add_(0, lab, x)
        return counts, sums

    res = {}
    for name, fn, reps in (("kernel", ours, 10), ("torch", torch_way, 3)):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        res[name] = e0.elapsed_time(e1) / reps
    t = res["kernel"] * 1e-3
    # algorithmic work of the assignment: K*D (sub + fma = 3 flop) per point; bytes: 4*D per point
    print(json.dumps(dict(N=N, D=D, K=K, kernel_ms=res["kernel"], torch_ms=res["torch"],
                          tflops=N * K * D * 3 / t / 1e12, frac_fp32_peak=N * K * D * 3 / t / peak,
                          issue_frac=N * K * D * 2 / t / (peak / 2), gbs=N * D * 4 / t / 1e9)), flush=True)

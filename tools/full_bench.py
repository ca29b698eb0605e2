"""Times the materialising E-step (ops.estep_full) at bench shapes: python tools/full_bench.py "D K N" ..."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from density_deconvolution_b200 import ops
from density_deconvolution_b200.synthetic import make_mixture, make_points
dev = torch.device("cuda", 0)
for cfg in sys.argv[1:]:
    D, K, N = (int(v) for v in cfg.split())
    alpha, mu, Sigma, q = make_mixture(D, K, seed=0, device=dev)
    X, S = make_points(N, alpha, mu, q, seed=1)
    la = alpha.log()
    P = ops.pack_noise(S)
    res = {}
    for name, (Sx, layout) in (("full", (S, "full")), ("packed", (P, "packed"))):
        ops.estep_full(X, Sx, la, mu, Sigma, noise=layout); torch.cuda.synchronize()
        best = 1e9
        for _ in range(5):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); out = ops.estep_full(X, Sx, la, mu, Sigma, noise=layout); b.record(); torch.cuda.synchronize()
            best = min(best, a.elapsed_time(b)); del out
        nbytes = N * K * (1 + D + D * D) * 4
        res[name] = dict(ms=best, gbs=nbytes / best / 1e6)
    print(json.dumps(dict(D=D, K=K, N=N, **res)))

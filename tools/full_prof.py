"""Runs the materialising E-step a few times at a bench shape (for ncu captures of full_kernel)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from density_deconvolution_b200 import ops
from density_deconvolution_b200.synthetic import make_mixture, make_points
D, K, N = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
dev = torch.device("cuda", 0)
alpha, mu, Sigma, q = make_mixture(D, K, seed=0, device=dev)
X, S = make_points(N, alpha, mu, q, seed=1)
la = alpha.log()
for _ in range(3):
    out = ops.estep_full(X, S, la, mu, Sigma)
    torch.cuda.synchronize()
    del out
print("ok")

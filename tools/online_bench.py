"""Times OnlineDeconvGMM.fit minibatch steps (statistics kernel + online M-step kernel) on one GPU."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from density_deconvolution_b200 import OnlineDeconvGMM
from density_deconvolution_b200.synthetic import make_mixture, make_points

dev = torch.device("cuda", 0)
for (D, K, B) in ((10, 64, 100), (10, 64, 4096), (16, 128, 4096)):
    alpha, mu, Sigma, q = make_mixture(D, K, seed=0, device=dev)
    n_steps = 200
    X, S = make_points(B * n_steps, alpha, mu, q, seed=1)
    for rep in range(2):
        m = OnlineDeconvGMM(K, D, epochs=1, batch_size=B, w=1e-3, device=dev)
        m.set_params(alpha, mu + 0.3, 1.2 * Sigma)
        torch.cuda.synchronize(); t0 = time.perf_counter()
        m.fit((X, S), init=False)
        torch.cuda.synchronize(); dt = time.perf_counter() - t0
    # the epoch ends with one score_batch over the table: time it separately and subtract
    torch.cuda.synchronize(); t0 = time.perf_counter(); m.score_batch((X, S)); torch.cuda.synchronize()
    ds = time.perf_counter() - t0
    print(json.dumps(dict(D=D, K=K, batch=B, steps=n_steps, us_per_step=(dt - ds) / n_steps * 1e6,
                          train_ll=m.train_ll_curve[-1] / (B * n_steps))), flush=True)

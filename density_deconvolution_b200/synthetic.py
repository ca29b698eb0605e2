"""Synthetic deconvolution-GMM problems in the reference's own style
(experiments/gmm/checks/data.py:5-10 of the reference):
``mu_j ~ U(-10,10)^D``, ``Sigma_j = q_j^T q_j`` with ``q_j ~ 2 N(0,1)``, full per-point noise
``S_i = Q_i^T Q_i + 1e-3 I`` with ``Q_i ~ 0.5 N(0,1)``; points drawn from the noisy mixture.
Generated directly on the target device, in chunks."""
import torch


def make_mixture(D, K, seed=0, device="cpu", spread=10.0, scale=2.0):
    gen = torch.Generator(device="cpu").manual_seed(seed)
    mu = (torch.rand(K, D, generator=gen) * 2 - 1) * spread
    q = scale * torch.randn(K, D, D, generator=gen)
    Sigma = q.transpose(1, 2) @ q
    alpha = torch.full((K,), 1.0 / K)
    return alpha.to(device), mu.to(device), Sigma.to(device), q.to(device)


def make_points(N, alpha, mu, q, seed=1234, noise="full", noise_scale=0.5, chunk=1 << 20):
    """Returns ``X [N,D]`` and ``S`` (``[N,D,D]`` or ``[N,D]`` for ``noise='diag'``), float32."""
    device = mu.device
    K, D = mu.shape
    gen = torch.Generator(device=device).manual_seed(seed)
    X = torch.empty(N, D, device=device)
    S = torch.empty((N, D, D) if noise == "full" else (N, D), device=device)
    for lo in range(0, N, chunk):
        n = min(chunk, N - lo)
        comp = torch.multinomial(alpha.float(), n, replacement=True, generator=gen)
        e1 = torch.randn(n, 1, D, device=device, generator=gen)
        x = mu[comp] + (e1 @ q[comp])[:, 0, :]           # mu_c + q_c^T e1  ~ N(mu_c, Sigma_c)
        if noise == "full":
            Q = noise_scale * torch.randn(n, D, D, device=device, generator=gen)
            S[lo:lo + n] = Q.transpose(1, 2) @ Q
            S[lo:lo + n].diagonal(dim1=1, dim2=2).add_(1e-3)
            e2 = torch.randn(n, 1, D, device=device, generator=gen)
            x = x + (e2 @ Q)[:, 0, :] + (1e-3 ** 0.5) * torch.randn(n, D, device=device, generator=gen)
            del Q
        else:
            var = 0.1 * (0.5 + torch.rand(n, D, device=device, generator=gen))
            S[lo:lo + n] = var
            x = x + var.sqrt() * torch.randn(n, D, device=device, generator=gen)
        X[lo:lo + n] = x
    return X, S

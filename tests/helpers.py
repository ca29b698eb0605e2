import numpy as np


def synth(rng, N, D, K, noise_scale=0.5, spread=3.0):
    """Small seeded problems (numpy, float64), same recipe as tests/golden/make_golden.py."""
    mu = rng.uniform(-spread, spread, size=(K, D))
    q = rng.standard_normal((K, D, D))
    Sigma = np.einsum("kba,kbc->kac", q, q) + 0.1 * np.eye(D)
    alpha = rng.dirichlet(np.ones(K) * 3.0)
    qn = noise_scale * rng.standard_normal((N, D, D))
    S = np.einsum("nba,nbc->nac", qn, qn) + 1e-3 * np.eye(D)
    comp = rng.integers(0, K, size=N)
    X = mu[comp] + np.einsum("nab,nb->na", np.linalg.cholesky(Sigma[comp] + S), rng.standard_normal((N, D)))
    return X, S, alpha, mu, Sigma


def rel(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return float(np.max(np.abs(a - b)) / (np.max(np.abs(b)) + 1e-300))

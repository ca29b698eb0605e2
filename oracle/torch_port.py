"""Torch-CPU port of the reference's E-step / M-step, used ONLY as the timed CPU baseline
(``bench.py``: ``cpu_baseline`` leg and ``--impl reference``) and in tests.

TEST / MEASUREMENT INFRASTRUCTURE ONLY - never imported by ``density_deconvolution_b200``.

The reference is pure Python on torch and ``/root/reference`` does not exist on the GPU box, so
the baseline is this restatement of the same ATen call sequence the reference issues (batched
``cholesky`` -> ``cholesky_solve`` against the identity -> broadcast ``matmul`` chains ->
``logsumexp``; deconv/gmm/deconv_gmm.py:39-109), running on all host threads torch is given.
It is pinned against the golden vectors in ``tests/test_oracle_golden.py``."""
import math

import torch


def e_step(X, S, weights, means, covars):
    """deconv/gmm/deconv_gmm.py:39-79 -> ``(ll_sum, (log_resps, cond_means, cond_covars))``."""
    d = X.shape[1]
    T = covars.unsqueeze(0) + S.unsqueeze(1)                                   # :43  [n,k,d,d]
    chol = torch.linalg.cholesky(T)                                            # :45
    T_inv = torch.cholesky_solve(torch.eye(d, dtype=X.dtype).expand_as(T), chol)   # :48-50
    dev = X.unsqueeze(1) - means                                               # :52
    u = torch.matmul(T_inv, dev.unsqueeze(-1))                                 # :53
    maha = torch.matmul(dev.unsqueeze(-2), u)[..., 0, 0]                       # :55-58
    log_r = -0.5 * (maha + d * math.log(2 * math.pi))
    log_r = log_r - chol.diagonal(dim1=-2, dim2=-1).log().sum(-1)              # :60
    log_r = log_r + torch.log(weights.reshape(1, -1))                          # :62
    cond_means = means + torch.matmul(covars.unsqueeze(0), u)[..., 0]          # :64-67
    cond_covars = covars - torch.matmul(covars, torch.matmul(T_inv, covars))   # :69-75
    ll = torch.logsumexp(log_r, dim=1, keepdim=True)                           # :77
    return ll.sum(), (log_r - ll, cond_means, cond_covars)                     # :78-79


def m_step_sums(expectations, new_means=None):
    """The reductions of deconv/gmm/deconv_gmm.py:85-104 in the accumulate-over-minibatches form
    of deconv/gmm/batch_deconv_gmm.py:95-120: pass 1 (``new_means is None``) returns
    ``(sum_r [k], sum_r_b [k,d])``, pass 2 returns ``sum_r (B + (b - m)(b - m)^T) [k,d,d]``."""
    log_r, b, B = expectations
    r = torch.exp(log_r)
    if new_means is None:
        return r.sum(0), torch.einsum("nk,nkd->kd", r, b)
    dev = b - new_means
    return torch.einsum("nk,nkab->kab", r, B + dev.unsqueeze(-1) * dev.unsqueeze(-2))


def em_iteration(X, S, weights, means, covars, w, chunk=2048):
    """One exact EM iteration over ``(X, S)`` in chunks of ``chunk`` points, the way the reference
    has to run it when ``[n,k,d,d]`` does not fit (BatchDeconvGMM: two E-step passes per epoch,
    batch_deconv_gmm.py:48-75).  Returns ``(ll_sum, weights [k,1], means, covars)``."""
    n, d = X.shape
    k = means.shape[0]
    s0 = torch.zeros(k, dtype=X.dtype)
    s1 = torch.zeros(k, d, dtype=X.dtype)
    ll = 0.0
    for lo in range(0, n, chunk):
        l, e = e_step(X[lo:lo + chunk], S[lo:lo + chunk], weights, means, covars)
        a, b = m_step_sums(e)
        s0 += a
        s1 += b
        ll += float(l)
    new_means = s1 / s0[:, None]
    s2 = torch.zeros(k, d, d, dtype=X.dtype)
    for lo in range(0, n, chunk):
        _, e = e_step(X[lo:lo + chunk], S[lo:lo + chunk], weights, means, covars)
        s2 += m_step_sums(e, new_means)
    new_covars = (s2 + 2 * w * torch.eye(d, dtype=X.dtype)) / s0[:, None, None]
    return ll, (s0 / n)[:, None], new_means, new_covars


def em_iteration_single_pass(X, S, weights, means, covars, w, chunk=2048):
    """One EM iteration with ONE E-step per chunk (the cheapest way to run the reference's own
    ``DeconvGMM`` maths in chunks: the covariance scatter is accumulated about the old means and
    re-centred at the end).  This is the CPU baseline bench.py reports: it does not charge the
    reference for BatchDeconvGMM's second pass."""
    n, d = X.shape
    k = means.shape[0]
    s0 = torch.zeros(k, dtype=X.dtype)
    s1 = torch.zeros(k, d, dtype=X.dtype)
    s2 = torch.zeros(k, d, d, dtype=X.dtype)
    ll = 0.0
    for lo in range(0, n, chunk):
        l, e = e_step(X[lo:lo + chunk], S[lo:lo + chunk], weights, means, covars)
        a, b = m_step_sums(e)
        s2 += m_step_sums(e, means)
        s0 += a
        s1 += b
        ll += float(l)
    new_means = s1 / s0[:, None]
    shift = new_means - means
    new_covars = (s2 - s0[:, None, None] * shift.unsqueeze(-1) * shift.unsqueeze(-2)
                  + 2 * w * torch.eye(d, dtype=X.dtype)) / s0[:, None, None]
    return ll, (s0 / n)[:, None], new_means, new_covars


def loglik_sum(X, S, weights, means, covars, chunk=4096):
    """``score_batch`` (deconv/gmm/online_deconv_gmm.py:190-205): the reference runs the whole
    ``_e_step`` and discards the moments."""
    return sum(float(e_step(X[lo:lo + chunk], S[lo:lo + chunk], weights, means, covars)[0])
               for lo in range(0, X.shape[0], chunk))

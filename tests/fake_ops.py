"""TEST DOUBLE for ``density_deconvolution_b200.ops`` built on the CPU oracle.

The product has no CPU path.  The ``-m "not gpu"`` tests of the HOST logic (estimator loops,
statistics algebra, the all-reduce plumbing under gloo) monkeypatch
``density_deconvolution_b200.estimators._ops`` with this module so that the same estimator code
runs on CPU tensors; the numbers come from ``oracle/gmm_oracle.py`` (float64, cast to float32
at the boundary like the real kernels' outputs).  Never imported by the package itself."""
import math

import numpy as np
import torch

from oracle import gmm_oracle as orc


def stats_len(D, K):
    return K * (1 + D + D * D) + 4


def _np(t):
    return None if t is None else t.detach().cpu().numpy().astype(np.float64)


def _full_noise(X, S, noise):
    N, D = X.shape
    if noise == "full":
        return _np(S)
    if noise == "diag":
        s = _np(S)
        out = np.zeros((N, D, D))
        out[:, np.arange(D), np.arange(D)] = s
        return out
    if noise == "shared":
        return np.broadcast_to(_np(S), (N, D, D)).copy()
    if noise == "packed":
        ia, ib = np.tril_indices(D)
        out = np.zeros((N, D, D))
        out[:, ia, ib] = _np(S)
        out[:, ib, ia] = _np(S)
        return out
    return np.zeros((N, D, D))


def loglik(X, S, log_alpha, mu, Sigma, noise="full", want_sum=False):
    Sn = _full_noise(X, S, noise)
    N = X.shape[0]
    bad = 0
    try:
        ll = orc.loglik(_np(X), Sn, np.exp(_np(log_alpha)), _np(mu), _np(Sigma)) if N else np.zeros(0)
    except np.linalg.LinAlgError:
        ll, bad = np.full(N, -math.inf), 1
    out = torch.tensor(ll, dtype=torch.float32, device=X.device)
    if want_sum:
        return out, torch.tensor([float(np.sum(ll)), float(bad)], dtype=torch.float64, device=X.device)
    return out


def estep_stats(X, S, log_alpha, mu, Sigma, g=None, resp_floor=0.0, noise="full", want_ll=True, out_stats=None):
    N, D = X.shape
    K = mu.shape[0]
    Sn = _full_noise(X, S, noise)
    buf = np.zeros(stats_len(D, K))
    ll = np.zeros(0)
    if N:
        try:
            s0, s1, S2, ll = orc.suff_stats(_np(X), Sn, np.exp(_np(log_alpha)), _np(mu), _np(Sigma), g=_np(g),
                                            resp_floor=resp_floor)
            buf[:K] = s0
            buf[K:K + K * D] = s1.reshape(-1)
            buf[K + K * D:K + K * D + K * D * D] = S2.reshape(-1)
            tot = float(ll.sum())
            hi = float(np.float32(tot))
            buf[-4], buf[-3] = hi, tot - hi
            buf[-2] = float(_np(g).sum()) if g is not None else float(N)
        except np.linalg.LinAlgError:
            ll = np.full(N, -math.inf)
            buf[-4], buf[-1] = -math.inf, 1.0
    st = torch.tensor(buf, dtype=torch.float32, device=X.device)
    if out_stats is not None:
        out_stats.copy_(st)
        st = out_stats
    return (torch.tensor(ll, dtype=torch.float32, device=X.device) if want_ll else None), st


def estep_full(X, S, log_alpha, mu, Sigma, noise="full", want_means=True, want_covars=True):
    Sn = _full_noise(X, S, noise)
    N, D = X.shape
    K = mu.shape[0]
    dev = X.device
    ll_sum, exp, ll = orc.e_step(_np(X), Sn, np.exp(_np(log_alpha)), _np(mu), _np(Sigma))
    if exp is None:
        nan = torch.full((N, K), float("nan"), device=dev)
        return torch.full((N,), -math.inf, device=dev), nan, None, None, torch.ones(1, dtype=torch.int32, device=dev)
    f = lambda a: torch.tensor(a, dtype=torch.float32, device=dev)  # noqa: E731
    return (f(ll), f(exp[0]), f(exp[1]) if want_means else None, f(exp[2]) if want_covars else None,
            torch.zeros(1, dtype=torch.int32, device=dev))


def unpack_stats(stats, D, K):
    o = 0
    s0 = stats[o:o + K]; o += K
    s1 = stats[o:o + K * D].view(K, D); o += K * D
    S2 = stats[o:o + K * D * D].view(K, D, D); o += K * D * D
    return s0, s1, S2, stats[o:o + 4]


def mstep_update(stats, n_total, w, mu, Sigma, alpha, log_alpha):
    K, D = mu.shape
    s0, s1, S2, _ = unpack_stats(stats, D, K)
    wts, means, covars = orc.m_step_from_stats(_np(s0), _np(s1), _np(S2), _np(mu), _np(Sigma), n_total, w)
    mu.copy_(torch.tensor(means, dtype=torch.float32))
    Sigma.copy_(torch.tensor(covars, dtype=torch.float32))
    alpha.copy_(torch.tensor(wts[:, 0], dtype=torch.float32))
    log_alpha.copy_(torch.log(alpha))


# ---- k-means (ops.kmeans_*) ------------------------------------------------------------------
def kmeans_stats_len(D, K):
    return K * (1 + D) + 2


def kmeans_assign(X, centroids, want_labels=False, out_stats=None):
    K, D = centroids.shape
    labels, counts, sums, total = orc.kmeans_assign(_np(X), _np(centroids))
    hi = np.float32(total)
    buf = np.concatenate([counts, sums.reshape(-1), [float(hi), total - float(hi)]])
    stats = torch.tensor(buf, dtype=torch.float32, device=X.device)
    if out_stats is not None:
        out_stats.copy_(stats)
        stats = out_stats
    lab = torch.tensor(labels, dtype=torch.int32, device=X.device) if want_labels else None
    return lab, stats


def kmeans_update(mode, stats, centroids, counts_run=None):
    K, D = centroids.shape
    s = _np(stats)
    counts, sums = s[:K], s[K:K + K * D].reshape(K, D)
    if mode == "lloyd":
        new = orc.kmeans_lloyd_update(counts, sums, _np(centroids))
    else:
        new, n_new = orc.kmeans_minibatch_update(counts, sums, _np(centroids), _np(counts_run))
        counts_run.copy_(torch.tensor(n_new, dtype=torch.float32))
    centroids.copy_(torch.tensor(new, dtype=torch.float32))


def kmeans_total_distance(stats):
    return float(stats[-2].item()) + float(stats[-1].item())


# ---- minibatch SGD parameter side (ops.sgd_*) ----------------------------------------------------
def sgd_params(soft_weights, l_diag, l_lower, out_log_alpha=None, out_Sigma=None):
    K, D = l_diag.shape
    alpha, _, Sigma, _ = orc.sgd_to_mixture(_np(soft_weights), np.zeros((K, D)), _np(l_diag), _np(l_lower))
    la = torch.tensor(np.log(alpha), dtype=torch.float32, device=soft_weights.device)
    Sg = torch.tensor(Sigma, dtype=torch.float32, device=soft_weights.device)
    if out_log_alpha is not None:
        out_log_alpha.copy_(la); la = out_log_alpha
    if out_Sigma is not None:
        out_Sigma.copy_(Sg); Sg = out_Sigma
    return la, Sg


def sgd_adam_step(stats, n_batch, n_total, w, lr, betas, eps, step, params, exp_avg, exp_avg_sq, log_alpha, Sigma,
                  epoch_acc=None):
    K, D = params[1].shape
    s = _np(stats)
    s0, s1, S2 = s[:K], s[K:K + K * D].reshape(K, D), s[K + K * D:K + K * D + K * D * D].reshape(K, D, D)
    new_p, new_m, new_v = orc.sgd_adam_step((s0, s1, S2), [_np(p) for p in params],
                                            ([_np(m) for m in exp_avg], [_np(v) for v in exp_avg_sq]),
                                            n_batch, n_total, w, lr, betas[0], betas[1], eps, step)
    with torch.no_grad():
        for dst, src in zip(list(params) + list(exp_avg) + list(exp_avg_sq), new_p + new_m + new_v):
            dst.copy_(torch.tensor(src, dtype=torch.float32).reshape(dst.shape))
    sgd_params(params[0], params[2], params[3], out_log_alpha=log_alpha, out_Sigma=Sigma)
    if epoch_acc is not None:
        epoch_acc[0] += float(s[-4]) + float(s[-3])
        epoch_acc[1] += float(s[-1])


# ---- posterior sampling (ops.posterior_sample) ---------------------------------------------------
def posterior_sample(X, S, log_alpha, mu, Sigma, u, z, noise="full", want_comp=False):
    Sn = _full_noise(X, S, noise)
    out, comp, cdf = orc.posterior_sample(_np(X), Sn, np.exp(_np(log_alpha)), _np(mu), _np(Sigma), _np(u), _np(z))
    r = np.diff(np.concatenate([np.zeros((cdf.shape[0], 1)), cdf], axis=1), axis=1)
    dev = X.device
    return (torch.tensor(out, dtype=torch.float32, device=dev),
            torch.tensor(comp, dtype=torch.int32, device=dev) if want_comp else None,
            torch.tensor(np.log(np.maximum(r, 1e-300)), dtype=torch.float32, device=dev),
            torch.zeros(1, dtype=torch.int32, device=dev))


# ---- Online-EM M-step (ops.online_mstep): the [K,.] algebra of online_deconv_gmm.py:146-188 in float64 ----
def online_mstep(stats, resp_floor, step_size, w, global_batch, mu, Sigma, sum_resps, sum_cond_means, alpha, log_alpha,
                 crash_flag=None):
    from density_deconvolution_b200 import stats_algebra as alg
    K, D = mu.shape
    if (crash_flag is not None and int(crash_flag.item())) or float(stats[-1]) > 0:
        if crash_flag is not None:
            crash_flag.fill_(1)
        return
    s = stats.double()
    s0, s1, S2 = s[:K], s[K:K + K * D].view(K, D), s[K + K * D:K + K * D + K * D * D].view(K, D, D)
    m_old, Sg = mu.double(), Sigma.double()
    Sg = torch.tril(Sg) + torch.tril(Sg, -1).transpose(1, 2)
    wI = w * torch.eye(D, dtype=torch.float64)
    sr = (s0 + resp_floor)[:, None]
    scm = alg.sum_resp_means(s0, s1, m_old, Sg)
    bm = scm / sr
    bc = alg.centred_scatter(s0, s1, S2, m_old, Sg, bm) / sr[:, :, None]
    sum_old = sum_resps.double().view(K, 1)
    sr_new = (1 - step_size) * sum_old + step_size * sr
    scm_new = (1 - step_size) * sum_cond_means.double() + step_size * scm
    m_new = scm_new / sr_new
    cov = ((1 - step_size) * alg.moment_shift(Sg - wI, sum_old / sr_new, m_old, m_new)
           + step_size * alg.moment_shift(bc, sr / sr_new, bm, m_new) + wI)
    sum_resps.copy_(sr_new.view(sum_resps.shape).float())
    sum_cond_means.copy_(scm_new.float())
    mu.copy_(m_new.float())
    Sigma.copy_(cov.float())
    al = (sr_new / global_batch).view(-1)
    alpha.copy_(al.view(alpha.shape).float())
    log_alpha.copy_(al.log().float())


def posterior_logprob(X, S, log_alpha, mu, Sigma, V, noise="full"):
    Sn = _full_noise(X, S, noise)
    out = orc.posterior_log_prob(_np(X), Sn, np.exp(_np(log_alpha)), _np(mu), _np(Sigma), _np(V))
    return torch.tensor(out, dtype=torch.float32, device=X.device), torch.zeros(1, dtype=torch.int32, device=X.device)


def gather_points(X, S, idx, outX=None, outS=None):
    return X.index_select(0, idx), (S.index_select(0, idx) if S is not None else None)

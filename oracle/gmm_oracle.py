"""NumPy restatement of the reference's deconvolution-GMM mathematics.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).  Every function names the
reference lines it restates; paths are relative to the reference checkout.
All functions take/return ``numpy`` arrays and compute in the dtype given by
``dtype`` (float64 by default: the tie-breaker precision).

Notation: N points, D dims, K components; ``X [N,D]`` noisy data, ``S [N,D,D]``
per-point noise covariances, ``alpha [K]`` mixture weights, ``mu [K,D]``,
``Sigma [K,D,D]``; ``T_ij = Sigma_j + S_i``, ``d_ij = x_i - mu_j``,
``u_ij = T_ij^{-1} d_ij``.
"""
import math

import numpy as np

LOG_2PI = math.log(2.0 * math.pi)


def _logsumexp(a, axis):
    m = np.max(a, axis=axis, keepdims=True)
    m = np.where(np.isfinite(m), m, 0.0)
    return np.squeeze(m, axis=axis) + np.log(np.sum(np.exp(a - m), axis=axis))


def pair_terms(X, S, mu, Sigma, dtype=np.float64):
    """Per-(point, component) factorisation terms.

    Restates deconv/gmm/deconv_gmm.py:43-53 (``T``, ``T_chol``, ``T_inv``,
    ``diff``, ``T_inv_diff``).  Only the lower triangle of ``T`` is read, as
    LAPACK ``potrf('L')`` does for ``torch.cholesky``.
    Returns dict with ``T_inv [N,K,D,D]``, ``u [N,K,D]``, ``diff [N,K,D]``,
    ``maha [N,K]``, ``half_logdet [N,K]`` (= sum log diag chol(T)).
    Raises ``np.linalg.LinAlgError`` when some ``T_ij`` is not positive definite.
    """
    X = np.asarray(X, dtype=dtype)
    S = np.asarray(S, dtype=dtype)
    mu = np.asarray(mu, dtype=dtype)
    Sigma = np.asarray(Sigma, dtype=dtype)
    T = Sigma[None, :, :, :] + S[:, None, :, :]
    low = np.tril(T)
    T = low + np.swapaxes(np.tril(T, -1), -1, -2)
    L = np.linalg.cholesky(T)
    D = X.shape[1]
    eye = np.eye(D, dtype=dtype)
    L_inv = np.linalg.solve(L, np.broadcast_to(eye, L.shape))
    T_inv = np.swapaxes(L_inv, -1, -2) @ L_inv
    diff = X[:, None, :] - mu[None, :, :]
    u = (T_inv @ diff[..., None])[..., 0]
    maha = np.sum(diff * u, axis=-1)
    half_logdet = np.sum(np.log(np.diagonal(L, axis1=-2, axis2=-1)), axis=-1)
    return dict(T_inv=T_inv, u=u, diff=diff, maha=maha, half_logdet=half_logdet)


def log_joint(X, S, alpha, mu, Sigma, dtype=np.float64):
    """``log alpha_j + log N(x_i | mu_j, Sigma_j + S_i)`` -> ``[N,K]``.

    Restates deconv/gmm/deconv_gmm.py:54-62 (before normalisation) and
    deconv/gmm/sgd_deconv_gmm.py:19-24.
    """
    t = pair_terms(X, S, mu, Sigma, dtype)
    D = np.asarray(X).shape[1]
    lp = -0.5 * (t["maha"] + D * LOG_2PI) - t["half_logdet"]
    with np.errstate(divide="ignore"):
        lp = lp + np.log(np.asarray(alpha, dtype=dtype).reshape(-1))[None, :]
    return lp, t


def loglik(X, S, alpha, mu, Sigma, dtype=np.float64):
    """Per-point ``log sum_j alpha_j N(x_i | mu_j, Sigma_j + S_i)`` -> ``[N]``.

    deconv/gmm/deconv_gmm.py:77 / deconv/gmm/sgd_deconv_gmm.py:26.
    """
    lp, _ = log_joint(X, S, alpha, mu, Sigma, dtype)
    return _logsumexp(lp, axis=1)


def e_step(X, S, alpha, mu, Sigma, dtype=np.float64):
    """Restates ``DeconvGMM._e_step`` (deconv/gmm/deconv_gmm.py:39-79).

    Returns ``(ll_sum, (log_resps [N,K], cond_means [N,K,D], cond_covars
    [N,K,D,D]), ll_i [N])``.  A non-PD ``T`` gives ``(-inf, None, None)`` as the
    reference does at :46-47.
    """
    Sigma = np.asarray(Sigma, dtype=dtype)
    mu = np.asarray(mu, dtype=dtype)
    try:
        lp, t = log_joint(X, S, alpha, mu, Sigma, dtype)
    except np.linalg.LinAlgError:
        return -math.inf, None, None
    cond_means = mu[None, :, :] + np.einsum("kab,nkb->nka", Sigma, t["u"])
    cond_covars = Sigma[None] - np.einsum("kab,nkbc,kcd->nkad", Sigma, t["T_inv"], Sigma)
    ll = _logsumexp(lp, axis=1)
    log_resps = lp - ll[:, None]
    return float(np.sum(ll)), (log_resps, cond_means, cond_covars), ll


def m_step(expectations, w, dtype=np.float64):
    """Restates ``DeconvGMM._m_step`` (deconv/gmm/deconv_gmm.py:81-109).

    ``w`` is the scalar regulariser (the reference stores ``w * I``).  Returns
    ``(weights [K,1], means [K,D], covars [K,D,D])``.  The covariance uses the
    NEW means (:92) and adds ``2 w I`` to the un-normalised sum (:106-107).
    """
    log_resps, cond_means, cond_covars = [np.asarray(a, dtype=dtype) for a in expectations]
    n, K, D = cond_means.shape
    r = np.exp(log_resps)
    s0 = r.sum(axis=0)
    means = np.einsum("nk,nkd->kd", r, cond_means) / s0[:, None]
    dev = means[None] - cond_means
    scatter = cond_covars + dev[..., :, None] * dev[..., None, :]
    covars = np.einsum("nk,nkab->kab", r, scatter) + 2.0 * w * np.eye(D, dtype=dtype)
    covars = covars / s0[:, None, None]
    return (s0 / n)[:, None], means, covars


def suff_stats(X, S, alpha, mu, Sigma, g=None, resp_floor=0.0, dtype=np.float64):
    """The fused statistics of SURVEY.md Appendix C.

    ``omega_ij = g_i (r_ij + resp_floor)``;
    ``s0_j = sum_i omega_ij``; ``s1_j = sum_i omega_ij u_ij``;
    ``S2_j = sum_i omega_ij (u_ij u_ij^T - T_ij^{-1})``.
    Returns ``(s0 [K], s1 [K,D], S2 [K,D,D], ll_i [N])``.
    """
    lp, t = log_joint(X, S, alpha, mu, Sigma, dtype)
    ll = _logsumexp(lp, axis=1)
    om = np.exp(lp - ll[:, None]) + resp_floor
    if g is not None:
        om = om * np.asarray(g, dtype=dtype)[:, None]
    u = t["u"]
    s0 = om.sum(axis=0)
    s1 = np.einsum("nk,nkd->kd", om, u)
    S2 = np.einsum("nk,nkab->kab", om, u[..., :, None] * u[..., None, :] - t["T_inv"])
    return s0, s1, S2, ll


def suff_stats_scales(X, S, alpha, mu, Sigma, g=None, resp_floor=0.0, dtype=np.float64):
    """Per-component magnitudes of the sums behind ``suff_stats`` (test infrastructure: the natural
    error scale of a float32 accumulation of each component's statistics, also where the signed sum
    cancels): ``A1_j = sum_i |omega_ij| |u_ij|`` ``[K,D]``, ``A2_j = sum_i |omega_ij| (|u u^T| + |T^{-1}|)``
    ``[K,D,D]``."""
    lp, t = log_joint(X, S, alpha, mu, Sigma, dtype)
    ll = _logsumexp(lp, axis=1)
    om = np.abs(np.exp(lp - ll[:, None]) + resp_floor)
    if g is not None:
        om = om * np.abs(np.asarray(g, dtype=dtype))[:, None]
    u = np.abs(t["u"])
    A1 = np.einsum("nk,nkd->kd", om, u)
    A2 = np.einsum("nk,nkab->kab", om, u[..., :, None] * u[..., None, :] + np.abs(t["T_inv"]))
    return A1, A2


def m_step_from_stats(s0, s1, S2, mu, Sigma, n_total, w, dtype=np.float64):
    """EM update written on the fused statistics.

    Algebraically identical to :func:`m_step` (deconv/gmm/deconv_gmm.py:81-109):
    ``delta = Sigma s1 / s0``; ``mu' = mu + delta``;
    ``Sigma' = Sigma + Sigma (S2/s0) Sigma - delta delta^T + 2 w I / s0``;
    ``alpha' = s0 / n``.
    """
    s0 = np.asarray(s0, dtype=dtype)
    mu = np.asarray(mu, dtype=dtype)
    Sigma = np.asarray(Sigma, dtype=dtype)
    D = mu.shape[1]
    delta = np.einsum("kab,kb->ka", Sigma, np.asarray(s1, dtype=dtype)) / s0[:, None]
    P = np.einsum("kab,kbc,kcd->kad", Sigma, np.asarray(S2, dtype=dtype), Sigma) / s0[:, None, None]
    covars = Sigma + P - delta[:, :, None] * delta[:, None, :]
    covars = covars + (2.0 * w / s0)[:, None, None] * np.eye(D, dtype=dtype)
    return (s0 / n_total)[:, None], mu + delta, covars


def centred_scatter_from_stats(s0, s1, S2, mu, Sigma, centre):
    """``M(m) = sum_i omega (B + (b-m)(b-m)^T)`` for any centre ``m [K,D]``.

    SURVEY.md Appendix C; serves deconv/gmm/batch_deconv_gmm.py:106-120
    (``m`` = new means) and deconv/gmm/online_deconv_gmm.py:159-166 (``m`` =
    batch means).
    """
    v = np.einsum("kab,kb->ka", Sigma, s1)
    c = mu - centre
    M = s0[:, None, None] * (Sigma + c[:, :, None] * c[:, None, :])
    M = M + np.einsum("kab,kbc,kcd->kad", Sigma, S2, Sigma)
    M = M + c[:, :, None] * v[:, None, :] + v[:, :, None] * c[:, None, :]
    return M


# ----------------------------------------------------------------------------
# Online (stochastic-approximation) EM: deconv/gmm/online_deconv_gmm.py
# ----------------------------------------------------------------------------

def moment_shift(covar, scale, b, c):
    """Restates ``OnlineDeconvGMM._adjust`` (online_deconv_gmm.py:134-144):
    ``scale (C + b b^T) - c c^T`` written in the reference's symmetric form."""
    root = np.sqrt(scale)
    lo = root * b - c
    hi = root * b + c
    out = scale[:, :, None] * covar
    out = out + 0.5 * lo[:, :, None] * hi[:, None, :]
    out = out + 0.5 * hi[:, :, None] * lo[:, None, :]
    return out


def online_m_step(state, expectations, step_size, w, batch_size, eps, dtype=np.float64):
    """Restates ``OnlineDeconvGMM._m_step`` (online_deconv_gmm.py:146-188).

    ``state`` = dict(sum_resps [K,1], sum_cond_means [K,D], means, covars);
    ``eps`` = machine epsilon of the dtype the reference ran in (:150,:154).
    Returns the new state plus ``weights [K,1]``.
    """
    log_resps, cond_means, cond_covars = [np.asarray(a, dtype=dtype) for a in expectations]
    D = cond_means.shape[2]
    wI = w * np.eye(D, dtype=dtype)
    r = np.exp(log_resps)[:, :, None] + 10.0 * eps
    sum_r = r.sum(axis=0) + 10.0 * eps
    sum_m = (r * cond_means).sum(axis=0)
    batch_means = sum_m / sum_r
    dev = cond_means - batch_means
    scat = dev[..., :, None] * dev[..., None, :] + cond_covars
    batch_covars = (r[..., None] * scat).sum(axis=0) / sum_r[:, :, None]

    old_means = state["means"]
    old_sum = state["sum_resps"]
    new_sum = (1 - step_size) * old_sum + step_size * sum_r
    new_sum_m = (1 - step_size) * state["sum_cond_means"] + step_size * sum_m
    new_means = new_sum_m / new_sum
    covars = (
        (1 - step_size) * moment_shift(state["covars"] - wI, old_sum / new_sum, old_means, new_means)
        + step_size * moment_shift(batch_covars, sum_r / new_sum, batch_means, new_means)
        + wI
    )
    return dict(sum_resps=new_sum, sum_cond_means=new_sum_m, means=new_means, covars=covars,
                weights=new_sum / batch_size)


# ----------------------------------------------------------------------------
# Two-pass minibatch EM: deconv/gmm/batch_deconv_gmm.py
# ----------------------------------------------------------------------------

def batch_pass_1(X, S, alpha, mu, Sigma, dtype=np.float64):
    """``BatchDeconvGMM._step_1`` (batch_deconv_gmm.py:95-104):
    ``(ll_sum, sum_r [K,1], sum_r_b [K,D])``."""
    ll_sum, (log_r, b, _), _ = e_step(X, S, alpha, mu, Sigma, dtype)
    r = np.exp(log_r)
    return ll_sum, r.sum(axis=0)[:, None], np.einsum("nk,nkd->kd", r, b)


def batch_pass_2(X, S, alpha, mu, Sigma, new_means, dtype=np.float64):
    """``BatchDeconvGMM._step_2`` (batch_deconv_gmm.py:106-120) -> ``[K,D,D]``."""
    _, (log_r, b, B), _ = e_step(X, S, alpha, mu, Sigma, dtype)
    r = np.exp(log_r)
    dev = b - np.asarray(new_means, dtype=dtype)[None]
    return np.einsum("nk,nkab->kab", r, B + dev[..., :, None] * dev[..., None, :])


# ----------------------------------------------------------------------------
# SGD parametrisation: deconv/gmm/sgd_gmm.py:14-65, sgd_deconv_gmm.py:14-28
# ----------------------------------------------------------------------------

def strict_lower_index(D):
    """Row-major strictly-lower indices, the order of ``torch.tril_indices(D, D, -1)``
    used at deconv/gmm/sgd_gmm.py:34."""
    rows, cols = [], []
    for a in range(1, D):
        for b in range(a):
            rows.append(a)
            cols.append(b)
    return np.asarray(rows, dtype=np.int64), np.asarray(cols, dtype=np.int64)


def sgd_to_mixture(soft_weights, means, l_diag, l_lower, dtype=np.float64):
    """``softmax`` weights and ``Sigma = L L^T`` with ``L = strict-lower(l_lower) +
    diag(exp(l_diag))`` (deconv/gmm/sgd_gmm.py:38-47, sgd_deconv_gmm.py:17)."""
    sw = np.asarray(soft_weights, dtype=dtype)
    K, D = np.asarray(means).shape
    e = np.exp(sw - sw.max())
    alpha = e / e.sum()
    L = np.zeros((K, D, D), dtype=dtype)
    idx = np.arange(D)
    L[:, idx, idx] = np.exp(np.asarray(l_diag, dtype=dtype))
    rows, cols = strict_lower_index(D)
    if rows.size:
        L[:, rows, cols] = np.asarray(l_lower, dtype=dtype)
    Sigma = L @ np.swapaxes(L, -1, -2)
    return alpha, np.asarray(means, dtype=dtype), Sigma, L


def sgd_forward(X, S, soft_weights, means, l_diag, l_lower, dtype=np.float64):
    """``SGDDeconvGMMModule.forward`` (sgd_deconv_gmm.py:14-28) -> ``[N]``."""
    alpha, mu, Sigma, _ = sgd_to_mixture(soft_weights, means, l_diag, l_lower, dtype)
    return loglik(X, S, alpha, mu, Sigma, dtype)


def sgd_grads(X, S, soft_weights, means, l_diag, l_lower, g, dtype=np.float64):
    """Gradient of ``sum_i g_i ll_i`` w.r.t. the four SGD parameters, written on
    the fused statistics: what autograd gives through
    sgd_deconv_gmm.py:14-28 + sgd_gmm.py:38-47.
    Returns ``(d_soft_weights [K], d_means [K,D], d_l_diag [K,D], d_l_lower)``.
    """
    alpha, mu, Sigma, L = sgd_to_mixture(soft_weights, means, l_diag, l_lower, dtype)
    g = np.asarray(g, dtype=dtype)
    s0, s1, S2, _ = suff_stats(X, S, alpha, mu, Sigma, g=g, dtype=dtype)
    d_sw = s0 - g.sum() * alpha
    d_mu = s1
    G = 0.5 * S2                      # dF/dSigma, symmetric
    dL = 2.0 * np.einsum("kab,kbc->kac", G, L)
    D = mu.shape[1]
    idx = np.arange(D)
    d_diag = dL[:, idx, idx] * np.exp(np.asarray(l_diag, dtype=dtype))
    rows, cols = strict_lower_index(D)
    d_low = dL[:, rows, cols] if rows.size else np.zeros((mu.shape[0], 0), dtype=dtype)
    return d_sw, d_mu, d_diag, d_low


def sgd_reg_loss(l_diag, l_lower, means, n, n_total, w, dtype=np.float64):
    """``BaseSGDGMM.reg_loss`` (sgd_gmm.py:110-112):
    ``(n/n_total) * w * sum_{j,d} 1/Sigma_j[d,d]``."""
    K, D = np.asarray(means).shape
    _, _, Sigma, _ = sgd_to_mixture(np.zeros(K), means, l_diag, l_lower, dtype)
    diag = np.diagonal(Sigma, axis1=-2, axis2=-1)
    return float(((n / n_total) * w / diag).sum())


# ---------------------------------------------------------------------------------------------
# k-means initialisation (deconv/gmm/util.py)
# ---------------------------------------------------------------------------------------------
def kmeans_assign(X, centroids, dtype=np.float64):
    """Labels, counts, coordinate sums and total distance of one assignment pass.
    deconv/gmm/util.py:19-20 / :51-52 (``diffs.norm(dim=2).min(dim=1)[1]``; ties -> lowest index),
    :54 (bincount), :22-24 / :59-64 (the per-label sums behind the cluster means / masked shift),
    :27 (``distances[resp].sum()``)."""
    X = np.asarray(X, dtype=dtype)
    C = np.asarray(centroids, dtype=dtype)
    K = C.shape[0]
    dist = np.sqrt(((X[:, None, :] - C[None, :, :]) ** 2).sum(axis=2)) if len(X) else np.zeros((0, K), dtype=dtype)
    labels = dist.argmin(axis=1) if len(X) else np.zeros(0, dtype=np.int64)
    counts = np.bincount(labels, minlength=K).astype(dtype)
    sums = np.zeros_like(C)
    np.add.at(sums, labels, X)
    total = float(dist[np.arange(len(X)), labels].sum()) if len(X) else 0.0
    return labels, counts, sums, total


def kmeans_lloyd_update(counts, sums, centroids):
    """deconv/gmm/util.py:22-24 (cluster means); an empty cluster keeps its centroid (documented fix:
    the reference writes NaN there)."""
    C = np.array(centroids, dtype=np.float64)
    nz = counts > 0
    C[nz] = sums[nz] / counts[nz][:, None]
    return C


def kmeans_minibatch_update(counts, sums, centroids, counts_run):
    """deconv/gmm/util.py:54-64: ``counts += bincount``, ``eta = 1/counts``,
    ``centroids += sum_i eta_c (x_i - c)`` over the points assigned to c."""
    n_new = np.asarray(counts_run, dtype=np.float64) + counts
    C = np.asarray(centroids, dtype=np.float64)
    return C + (sums - counts[:, None] * C) / n_new[:, None], n_new


def sgd_loss_grads_from_stats(s0, s1, S2, soft_weights, means, l_diag, l_lower, n_batch, n_total, w,
                              dtype=np.float64):
    """Gradient of the fit loop's loss (deconv/gmm/sgd_gmm.py:173-181)
    ``-(1/n) sum_i ll_i + (n/n_total) w sum_{j,d} 1/Sigma_j[d,d]`` w.r.t. the four parameters, from
    the fused statistics of ``sum_i ll_i`` (``g = 1``): autograd through sgd_gmm.py:38-47,110-112."""
    alpha, mu, Sigma, L = sgd_to_mixture(soft_weights, means, l_diag, l_lower, dtype)
    K, D = mu.shape
    idx = np.arange(D)
    d_sw = -(s0 - alpha * s0.sum()) / n_batch
    d_mu = -s1 / n_batch
    G = -0.5 * S2 / n_batch
    G[:, idx, idx] -= (n_batch / n_total) * w / Sigma[:, idx, idx] ** 2
    dL = 2.0 * np.einsum("kab,kbc->kac", G, L)
    d_diag = dL[:, idx, idx] * np.exp(np.asarray(l_diag, dtype=dtype))
    rows, cols = strict_lower_index(D)
    d_low = dL[:, rows, cols] if rows.size else np.zeros((K, 0), dtype=dtype)
    return d_sw, d_mu, d_diag, d_low


def adam_update(p, g, m, v, lr, beta1, beta2, eps, step):
    """torch.optim.Adam (defaults; the optimiser of deconv/gmm/sgd_gmm.py:92-95), one tensor."""
    m = m + (g - m) * (1.0 - beta1)
    v = beta2 * v + (1.0 - beta2) * g * g
    bc1 = 1.0 - beta1 ** step
    bc2 = 1.0 - beta2 ** step
    denom = np.sqrt(v) / np.sqrt(bc2) + eps
    return p - (lr / bc1) * (m / denom), m, v


def sgd_adam_step(stats3, params, moments, n_batch, n_total, w, lr, beta1, beta2, eps, step):
    """One optimiser step of BaseSGDGMM.fit (sgd_gmm.py:167-182) from ``stats3 = (s0, s1, S2)``.
    ``params`` / ``moments = (exp_avg list, exp_avg_sq list)`` in the order soft_weights, means,
    l_diag, l_lower.  Returns ``(new_params, new_exp_avg, new_exp_avg_sq)``."""
    grads = sgd_loss_grads_from_stats(*stats3, *params, n_batch, n_total, w)
    new_p, new_m, new_v = [], [], []
    for p, g, m, v in zip(params, grads, moments[0], moments[1]):
        a, b, c = adam_update(np.asarray(p, np.float64), g, np.asarray(m, np.float64), np.asarray(v, np.float64),
                              lr, beta1, beta2, eps, step)
        new_p.append(a); new_m.append(b); new_v.append(c)
    return new_p, new_m, new_v


def posterior_sample(X, S, alpha, mu, Sigma, u, z, dtype=np.float64):
    """SGDDeconvGMM._sample_posterior (deconv/gmm/sgd_deconv_gmm.py:107-119) as a deterministic map
    of the caller's randomness: ``u [N,M]`` -> component by inverse CDF of ``r_i`` (what
    ``Categorical(logits=p_weights).sample`` draws), ``z [N,M,D]`` -> ``b_ij + cholesky(B_ij) z`` (what
    ``MultivariateNormal(loc, covariance_matrix).sample`` does with its standard normals:
    torch/distributions/multivariate_normal.py rsample = loc + scale_tril @ eps).
    Returns ``(samples [N,M,D], comp [N,M], cdf [N,K])``."""
    _, (log_r, b, B), _ = e_step(X, S, alpha, mu, Sigma, dtype)
    cdf = np.cumsum(np.exp(log_r), axis=1)
    N, M = np.asarray(u).shape
    K = cdf.shape[1]
    comp = np.minimum((np.asarray(u, dtype)[:, :, None] >= cdf[:, None, :]).sum(axis=2), K - 1)
    out = np.zeros((N, M, b.shape[2]), dtype=dtype)
    for i in range(N):
        for m in range(M):
            j = comp[i, m]
            out[i, m] = b[i, j] + np.linalg.cholesky(B[i, j]) @ np.asarray(z, dtype)[i, m]
    return out, comp, cdf


def posterior_log_prob(X, S, alpha, mu, Sigma, V, dtype=np.float64):
    """SGDDeconvGMM.posterior_log_prob (deconv/gmm/sgd_deconv_gmm.py:96-105) for ``V [N,M,D]``:
    ``logsumexp_j(log r_ij + log N(V[i,m] | b_ij, B_ij))`` -> ``[N,M]``."""
    _, (log_r, b, B), _ = e_step(X, S, alpha, mu, Sigma, dtype)
    V = np.asarray(V, dtype=dtype)
    N, M, D = V.shape
    L = np.linalg.cholesky(B)                                   # [N,K,D,D]
    diff = V[:, :, None, :] - b[:, None, :, :]                   # [N,M,K,D]
    y = np.linalg.solve(L[:, None], diff[..., None])[..., 0]
    maha = (y ** 2).sum(axis=-1)
    hld = np.log(np.diagonal(L, axis1=-2, axis2=-1)).sum(axis=-1)   # [N,K]
    lp = log_r[:, None, :] - 0.5 * D * np.log(2 * np.pi) - hld[:, None, :] - 0.5 * maha
    return _logsumexp(lp, axis=2)

"""[K,.]-sized algebra on the fused statistics: everything an estimator
needs after the kernel + all-reduce.  Plain torch on the parameter device; O(K D^3)."""
import torch


def sum_resp_means(s0, s1, mu, Sigma):
    """``sum_i omega_ij b_ij = s0 mu + Sigma s1`` (deconv/gmm/deconv_gmm.py:88,
    batch_deconv_gmm.py:102, online_deconv_gmm.py:156)."""
    return s0[:, None] * mu + torch.einsum("kab,kb->ka", Sigma, s1)


def centred_scatter(s0, s1, S2, mu, Sigma, centre):
    """``M(m) = sum_i omega (B + (b-m)(b-m)^T)`` for any centre ``m [K,D]``
    (batch_deconv_gmm.py:106-120, online_deconv_gmm.py:159-166, deconv_gmm.py:92-104)."""
    v = torch.einsum("kab,kb->ka", Sigma, s1)
    c = mu - centre
    M = s0[:, None, None] * (Sigma + c[:, :, None] * c[:, None, :])
    M = M + Sigma @ S2 @ Sigma
    return M + c[:, :, None] * v[:, None, :] + v[:, :, None] * c[:, None, :]


def moment_shift(covar, scale, b, c):
    """``scale (C + b b^T) - c c^T`` in the symmetric form of ``OnlineDeconvGMM._adjust``
    (online_deconv_gmm.py:134-144); ``scale [K,1]``."""
    root = torch.sqrt(scale)
    lo = root * b - c
    hi = root * b + c
    out = scale[:, :, None] * covar
    out = out + 0.5 * lo[:, :, None] * hi[:, None, :]
    return out + 0.5 * hi[:, :, None] * lo[:, None, :]


def sgd_param_grads(s0, s1, S2, sum_g, alpha):
    """Gradients of ``sum_i g_i ll_i`` w.r.t. ``(log_alpha_unnormalised, mu, Sigma)``:
    ``s0 - sum_g alpha``, ``s1``, ``S2 / 2`` (autograd of sgd_deconv_gmm.py:14-28)."""
    return s0 - sum_g * alpha, s1, 0.5 * S2

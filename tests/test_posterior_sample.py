"""Posterior sampling: ``dgmm_posterior_sample_f32`` against the oracle (exact map of
the caller's uniforms / normals) and against Monte-Carlo moments of the UNMODIFIED reference's
``SGDDeconvGMM._sample_posterior`` (``make_golden.py posterior_sample``)."""
import numpy as np
import pytest
import torch

from conftest import golden
from helpers import synth
from oracle import gmm_oracle as orc


@pytest.fixture(params=["oracle", pytest.param("cuda", marks=pytest.mark.gpu)])
def device(request, monkeypatch):
    from density_deconvolution_b200 import estimators
    if request.param == "oracle":
        import fake_ops
        monkeypatch.setattr(estimators, "_ops", fake_ops)
        return torch.device("cpu")
    return torch.device("cuda", 0)


def t32(a, device):
    return torch.tensor(np.asarray(a), dtype=torch.float32, device=device)


def test_oracle_moments_match_reference_posterior():
    """The oracle's mixture moments equal the reference's posterior_params; its draws have them."""
    z = golden("ref_posterior_sample.npz")
    alpha, mu, Sigma, _ = orc.sgd_to_mixture(z["soft_weights"], z["means"], z["l_diag"], z["l_lower"])
    _, (log_r, b, B), _ = orc.e_step(z["X"], z["S"], alpha, mu, Sigma)
    np.testing.assert_allclose(np.exp(log_r), np.exp(z["post_log_w"]), atol=1e-10)
    np.testing.assert_allclose(b, z["post_means"], atol=1e-9)
    np.testing.assert_allclose(B, z["post_covars"], atol=1e-9)
    rng = np.random.default_rng(0)
    N, D, M = z["X"].shape[0], z["X"].shape[1], 40000
    smp, comp, _ = orc.posterior_sample(z["X"], z["S"], alpha, mu, Sigma, rng.uniform(size=(N, M)),
                                        rng.standard_normal((N, M, D)))
    _check_moments(smp, z, M)


def _check_moments(smp, z, M):
    """Sample mean / covariance against the reference's Monte-Carlo moments (both carry MC error)."""
    N = smp.shape[0]
    for i in range(N):
        cov_ref = z["sample_cov"][i]
        se = np.sqrt(np.diag(cov_ref) * (1.0 / M + 1.0 / float(z["M"])))
        assert np.all(np.abs(smp[i].mean(axis=0) - z["sample_mean"][i]) < 5 * se), i
        cov = np.cov(smp[i].T)
        assert np.max(np.abs(cov - cov_ref)) < 0.06 * np.max(np.abs(cov_ref)) + 5e-3, i


def test_estimator_sample_posterior_matches_reference_distribution(device):
    from density_deconvolution_b200 import SGDDeconvGMM
    z = golden("ref_posterior_sample.npz")
    K, D = z["means"].shape
    m = SGDDeconvGMM(K, D, w=1e-3, batch_size=16, device=device)
    with torch.no_grad():
        for p in ("soft_weights", "means", "l_diag", "l_lower"):
            getattr(m.module, p).copy_(t32(z[p], device))
    torch.manual_seed(5)
    M = 40000
    smp = m.sample_posterior((t32(z["X"], device), t32(z["S"], device)), M)
    assert smp.shape == (z["X"].shape[0], M, D) and smp.device.type == "cpu"
    _check_moments(smp.numpy().astype(np.float64), z, M)


def _check_exact(N, D, K, M, seed, noise="full"):
    from density_deconvolution_b200 import ops
    dev = torch.device("cuda", 0)
    rng = np.random.default_rng(seed)
    X, S, alpha, mu, Sigma = synth(rng, N, D, K)
    if noise == "diag":
        S_dev = np.stack([np.diag(s) for s in S])
        S = np.stack([np.diag(np.diag(s)) for s in S])
    else:
        S_dev = S
    u = rng.uniform(size=(N, M))
    zz = rng.standard_normal((N, M, D))
    out, comp, log_r, status = ops.posterior_sample(t32(X, dev), t32(S_dev, dev), t32(np.log(alpha), dev), t32(mu, dev),
                                                    t32(Sigma, dev), t32(u, dev), t32(zz, dev), noise=noise,
                                                    want_comp=True)
    assert status.item() == 0
    ref, comp_ref, cdf = orc.posterior_sample(X, S, alpha, mu, Sigma, u.astype(np.float32), zz.astype(np.float32))
    comp = comp.cpu().numpy()
    # the picked component is integer work: exact, except where u sits on a CDF step to float32 rounding
    lo = np.concatenate([np.zeros((N, 1)), cdf[:, :-1]], axis=1)
    for i, m in zip(*np.nonzero(comp != comp_ref)):
        edge = min(abs(u[i, m] - cdf[i, min(comp[i, m], comp_ref[i, m])]), abs(u[i, m] - lo[i, max(comp[i, m], comp_ref[i, m])]))
        assert edge < 1e-5, (i, m, comp[i, m], comp_ref[i, m], edge)
    same = comp == comp_ref
    assert same.mean() > 0.99
    got = out.cpu().numpy().astype(np.float64)
    scale = max(1.0, np.abs(ref).max())
    assert np.max(np.abs(got[same] - ref[same])) < 1e-3 * scale
    r = np.exp(log_r.cpu().numpy().astype(np.float64))
    np.testing.assert_allclose(np.cumsum(r, axis=1), cdf, atol=1e-4)


@pytest.mark.gpu
@pytest.mark.parametrize("D", [1, 2, 3, 5, 8, 10, 13, 16, 20, 32])
def test_sample_kernel_dims(D):
    _check_exact(N=37, D=D, K=5, M=7, seed=D)


@pytest.mark.gpu
@pytest.mark.parametrize("K,noise", [(1, "full"), (64, "full"), (128, "diag"), (256, "full")])
def test_sample_kernel_components(K, noise):
    _check_exact(N=29, D=4, K=K, M=5, seed=K, noise=noise)


@pytest.mark.gpu
def test_sample_kernel_full_size_moments():
    """N = 2e5 points, M = 4 (no [N,K,D,D] tensor could exist): the draws' global mean equals the mean of
    the posterior means sum_j r_ij b_ij, computed independently from the fused statistics."""
    from density_deconvolution_b200 import ops
    from density_deconvolution_b200.synthetic import make_mixture, make_points
    dev = torch.device("cuda", 0)
    D, K, N, M = 10, 64, 200_000, 4
    alpha, mu, Sigma, q = make_mixture(D, K, seed=0, device=dev)
    X, S = make_points(N, alpha, mu, q, seed=3)
    g = torch.Generator(device=dev).manual_seed(1)
    u = torch.rand(N, M, device=dev, generator=g)
    z = torch.randn(N, M, D, device=dev, generator=g)
    out, comp, log_r, status = ops.posterior_sample(X, S, alpha.log(), mu, Sigma, u, z, want_comp=True)
    assert status.item() == 0 and torch.isfinite(out).all()
    # E[v | data] summed over points = sum_j (s0_j mu_j + Sigma_j s1_j)
    _, stats = ops.estep_stats(X, S, alpha.log(), mu, Sigma, want_ll=False)
    s0, s1, _, _ = ops.unpack_stats(stats, D, K)
    want = (s0[:, None] * mu + torch.einsum("kab,kb->ka", Sigma, s1)).sum(dim=0).double() / N
    got = out.double().mean(dim=(0, 1))
    sd = out.double().std(dim=(0, 1)) / (N * M) ** 0.5
    assert torch.all((got - want).abs() < 6 * sd + 1e-3 * want.abs()), (got, want)
    # component frequencies follow the responsibilities
    freq = torch.bincount(comp.flatten().long(), minlength=K).double() / (N * M)
    assert torch.all((freq - s0.double() / N).abs() < 6 * (freq / (N * M)).sqrt() + 1e-4)


# ---- posterior_log_prob (sgd_deconv_gmm.py:96-105) -------------------------------------------------
def test_oracle_posterior_log_prob_matches_reference():
    z = golden("ref_posterior_sample.npz")
    alpha, mu, Sigma, _ = orc.sgd_to_mixture(z["soft_weights"], z["means"], z["l_diag"], z["l_lower"])
    got = orc.posterior_log_prob(z["X"], z["S"], alpha, mu, Sigma, z["query3"])
    np.testing.assert_allclose(got, z["logprob3"], rtol=1e-9, atol=1e-9)


def test_estimator_posterior_log_prob_matches_reference(device):
    from density_deconvolution_b200 import SGDDeconvGMM
    z = golden("ref_posterior_sample.npz")
    K, D = z["means"].shape
    m = SGDDeconvGMM(K, D, w=1e-3, batch_size=16, device=device)
    with torch.no_grad():
        for p in ("soft_weights", "means", "l_diag", "l_lower"):
            getattr(m.module, p).copy_(t32(z[p], device))
    ctx = (t32(z["X"], device), t32(z["S"], device))
    lp3 = m.posterior_log_prob(t32(z["query3"], device), ctx).cpu().numpy()
    assert lp3.shape == z["logprob3"].shape
    np.testing.assert_allclose(lp3, z["logprob3"], rtol=2e-4, atol=2e-4)
    lp2 = m.posterior_log_prob(t32(z["query2"], device), ctx).cpu().numpy()
    assert lp2.shape == z["logprob2"].shape                      # [Q, N]: every query against every context
    np.testing.assert_allclose(lp2, z["logprob2"], rtol=2e-4, atol=2e-3)


@pytest.mark.gpu
@pytest.mark.parametrize("D,K", [(1, 3), (3, 4), (10, 64), (16, 8), (20, 5)])
def test_logprob_kernel_against_oracle(D, K):
    from density_deconvolution_b200 import ops
    dev = torch.device("cuda", 0)
    rng = np.random.default_rng(D * 10 + K)
    N, M = 23, 4
    X, S, alpha, mu, Sigma = synth(rng, N, D, K)
    V = X[:, None, :] + rng.standard_normal((N, M, D))
    out, status = ops.posterior_logprob(t32(X, dev), t32(S, dev), t32(np.log(alpha), dev), t32(mu, dev), t32(Sigma, dev),
                                        t32(V, dev))
    assert status.item() == 0
    want = orc.posterior_log_prob(X, S, alpha, mu, Sigma, V.astype(np.float32))
    np.testing.assert_allclose(out.cpu().numpy(), want, rtol=5e-4, atol=5e-3)

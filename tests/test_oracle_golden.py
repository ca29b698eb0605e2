"""Pins oracle/gmm_oracle.py to the reference: golden vectors made by importing the
reference (tests/golden/make_golden.py) and scipy closed forms."""
import math

import numpy as np
import pytest

from oracle import gmm_oracle as orc
from conftest import golden


def rel(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return float(np.max(np.abs(a - b)) / (np.max(np.abs(b)) + 1e-300))


@pytest.mark.parametrize("name,tol", [
    ("ref_em_d3_k4_f64.npz", 1e-10),
    ("ref_em_d10_k6_f64.npz", 1e-10),
    ("ref_em_d16_k5_f64.npz", 1e-9),
    ("ref_em_d10_k6_f32.npz", 2e-4),
])
def test_e_step_and_m_step_match_reference(name, tol):
    g = golden(name)
    ll, (log_r, b, B), ll_i = orc.e_step(g["X"], g["S"], g["alpha"], g["mu"], g["Sigma"])
    assert abs(ll - float(g["ll_sum"])) <= tol * abs(float(g["ll_sum"]))
    assert np.max(np.abs(np.exp(log_r) - np.exp(g["log_resps"]))) < max(tol, 1e-12) * 10
    assert rel(b, g["cond_means"]) < tol * 10
    assert rel(B, g["cond_covars"]) < tol * 100
    wts, mu1, Sg1 = orc.m_step((log_r, b, B), float(g["w"]))
    assert rel(wts, g["m1_weights"]) < tol * 10
    assert rel(mu1, g["m1_means"]) < tol * 10
    assert rel(Sg1, g["m1_covars"]) < tol * 100


@pytest.mark.parametrize("name", ["ref_em_d3_k4_f64.npz", "ref_em_d10_k6_f64.npz", "ref_em_d16_k5_f64.npz"])
def test_fused_statistics_reproduce_reference_m_step(name):
    """(s0,s1,S2) -> the reference M-step, to rounding."""
    g = golden(name)
    N = g["X"].shape[0]
    s0, s1, S2, ll_i = orc.suff_stats(g["X"], g["S"], g["alpha"], g["mu"], g["Sigma"])
    wts, mu1, Sg1 = orc.m_step_from_stats(s0, s1, S2, g["mu"], g["Sigma"], N, float(g["w"]))
    assert rel(wts, g["m1_weights"]) < 1e-10
    assert rel(mu1, g["m1_means"]) < 1e-10
    assert rel(Sg1, g["m1_covars"]) < 1e-9
    assert abs(ll_i.sum() - float(g["ll_sum"])) < 1e-9 * abs(float(g["ll_sum"]))


def test_multi_iteration_em_tracks_reference():
    g = golden("ref_em_d10_k6_f64.npz")
    X, S = g["X"], g["S"]
    alpha, mu, Sigma = g["alpha"], g["mu"], g["Sigma"]
    N = X.shape[0]
    for it in range(1, 4):
        s0, s1, S2, _ = orc.suff_stats(X, S, alpha, mu, Sigma)
        wts, mu, Sigma = orc.m_step_from_stats(s0, s1, S2, mu, Sigma, N, float(g["w"]))
        alpha = wts[:, 0]
        if it == 3:
            assert rel(mu, g["it3_means"]) < 1e-8
            assert rel(Sigma, g["it3_covars"]) < 1e-7
            assert rel(alpha, g["it3_weights"][:, 0]) < 1e-8
            assert abs(orc.loglik(X, S, alpha, mu, Sigma).sum() - float(g["it3_ll"])) < 1e-7


@pytest.mark.parametrize("name,tol", [("ref_sgd_d4_k5_f64.npz", 1e-9), ("ref_sgd_d7_k3_f32.npz", 5e-4)])
def test_sgd_forward_and_gradients_match_reference_autograd(name, tol):
    g = golden(name)
    args = (g["X"], g["S"], g["soft_weights"], g["means"], g["l_diag"], g["l_lower"])
    ll = orc.sgd_forward(*args)
    assert rel(ll, g["ll"]) < tol
    d_sw, d_mu, d_diag, d_low = orc.sgd_grads(*args, g=g["g"])
    scale = max(np.abs(g["d_means"]).max(), 1.0)
    assert np.max(np.abs(d_sw - g["d_soft_weights"])) < tol * scale * 10
    assert np.max(np.abs(d_mu - g["d_means"])) < tol * scale * 10
    assert np.max(np.abs(d_diag - g["d_l_diag"])) < tol * scale * 100
    assert np.max(np.abs(d_low - g["d_l_lower"])) < tol * scale * 100
    _, _, Sigma, _ = orc.sgd_to_mixture(g["soft_weights"], g["means"], g["l_diag"], g["l_lower"])
    assert rel(Sigma, g["covars"]) < max(tol, 1e-12)
    assert abs(orc.sgd_reg_loss(g["l_diag"], g["l_lower"], g["means"], 16, 64, 1e-3) - float(g["reg_loss"])) \
        < max(tol, 1e-12) * abs(float(g["reg_loss"])) * 10
    # posterior_params (sgd_deconv_gmm.py:71-94) == e_step from the SGD parametrisation
    alpha, mu, Sigma, _ = orc.sgd_to_mixture(g["soft_weights"], g["means"], g["l_diag"], g["l_lower"])
    _, (log_r, b, B), _ = orc.e_step(g["X"], g["S"], alpha, mu, Sigma)
    assert np.max(np.abs(np.exp(log_r) - np.exp(g["post_log_w"]))) < tol * 10
    assert rel(b, g["post_means"]) < tol * 10
    assert rel(B, g["post_covars"]) < tol * 100


@pytest.mark.parametrize("name,tol", [("ref_online_batch_d3_k4_f64.npz", 1e-10),
                                      ("ref_online_batch_d5_k3_f32.npz", 5e-4)])
def test_online_and_batch_em_match_reference(name, tol):
    g = golden(name)
    X, S, alpha, mu, Sigma = g["X"], g["S"], g["alpha"], g["mu"], g["Sigma"]
    N = X.shape[0]
    w = float(g["w"])
    state = dict(sum_resps=alpha[:, None] * N, sum_cond_means=mu * (alpha[:, None] * N),
                 means=mu.copy(), covars=Sigma.copy())
    ll, e, _ = orc.e_step(X, S, alpha, mu, Sigma)
    assert abs(ll - float(g["ll_sum"])) < tol * abs(ll)
    state = orc.online_m_step(state, e, float(g["step_size"]), w, N, float(g["eps"]))
    assert rel(state["weights"], g["on_weights"]) < tol * 10
    assert rel(state["means"], g["on_means"]) < tol * 10
    assert rel(state["covars"], g["on_covars"]) < tol * 100
    ll2, e2, _ = orc.e_step(X, S, state["weights"][:, 0], state["means"], state["covars"])
    state = orc.online_m_step(state, e2, float(g["step_size"]), w, N, float(g["eps"]))
    assert rel(state["covars"], g["on2_covars"]) < tol * 100
    # two-pass batch EM over two half-batches
    half = N // 2
    sr = np.zeros((alpha.size, 1))
    sm = np.zeros_like(mu)
    tl = 0.0
    for sl in (slice(0, half), slice(half, N)):
        a, b, c = orc.batch_pass_1(X[sl], S[sl], alpha, mu, Sigma)
        tl += a
        sr += b
        sm += c
    new_means = sm / sr
    sc = sum(orc.batch_pass_2(X[sl], S[sl], alpha, mu, Sigma, new_means) for sl in (slice(0, half), slice(half, N)))
    assert abs(tl - float(g["bt_ll"])) < tol * abs(tl)
    assert rel(sr, g["bt_sum_resps"]) < tol * 10
    assert rel(sm, g["bt_sum_cond_means"]) < tol * 10
    assert rel(sc, g["bt_sum_cond_covars"]) < tol * 100
    # ... and the same quantities from the fused statistics (Appendix C)
    s0, s1, S2, _ = orc.suff_stats(X, S, alpha, mu, Sigma)
    M = orc.centred_scatter_from_stats(s0, s1, S2, mu, Sigma, new_means)
    assert rel(M, g["bt_sum_cond_covars"]) < max(tol, 1e-9) * 100


def test_online_floor_is_a_weight_floor():
    """online_deconv_gmm.py:150-166 from fused statistics with resp_floor."""
    g = golden("ref_online_batch_d3_k4_f64.npz")
    X, S, alpha, mu, Sigma = g["X"], g["S"], g["alpha"], g["mu"], g["Sigma"]
    eps = float(g["eps"])
    _, (log_r, b, B), _ = orc.e_step(X, S, alpha, mu, Sigma)
    r = np.exp(log_r)[:, :, None] + 10 * eps
    sum_r = r.sum(0) + 10 * eps
    sum_m = (r * b).sum(0)
    bm = sum_m / sum_r
    dev = b - bm
    bc = (r[..., None] * (dev[..., :, None] * dev[..., None, :] + B)).sum(0)
    s0, s1, S2, _ = orc.suff_stats(X, S, alpha, mu, Sigma, resp_floor=10 * eps)
    assert rel(s0 + 10 * eps, sum_r[:, 0]) < 1e-12
    assert rel(s0[:, None] * mu + np.einsum("kab,kb->ka", Sigma, s1), sum_m) < 1e-10
    assert rel(orc.centred_scatter_from_stats(s0, s1, S2, mu, Sigma, bm), bc) < 1e-9


def test_closed_form_mixture_1():
    """SURVEY.md 8c item 1 == deconv/utils/compute_2d_log_likelihood.py:24-34 convolved with S_i."""
    from scipy.stats import multivariate_normal as mn
    alpha = np.full(3, 1 / 3)
    mu = np.array([[0.0, 0.0], [2.0, 3.0], [2.0, -3.0]])
    Sigma = np.array([np.diag([0.1, 1.5]), np.diag([1.0, 0.1]), np.diag([1.0, 0.1])])
    X = np.array([[0.5, 0.25], [2.0, 2.5], [-1.0, 4.0]])
    S = np.array([[[0.1, 0], [0, 1.0]], [[0.3, 0.1], [0.1, 0.2]], [[2.0, -0.5], [-0.5, 1.0]]])
    want = np.array([math.log(sum(alpha[j] * mn.pdf(X[i], mu[j], Sigma[j] + S[i]) for j in range(3)))
                     for i in range(3)])
    got = orc.loglik(X, S, alpha, mu, Sigma)
    assert np.max(np.abs(got - want)) < 1e-12
    assert abs(got.sum() - (-11.056704960174)) < 1e-9
    _, (log_r, _, _), _ = orc.e_step(X, S, alpha, mu, Sigma)
    assert np.max(np.abs(np.exp(log_r[0]) - [0.98265399135522, 0.01381349513774, 0.00353251350704])) < 1e-12


def test_data_gen_fixture_and_robustness():
    g = golden("ref_data_gen.npz")
    S = np.broadcast_to(g["S"], (g["x_test"].shape[0], 2, 2))
    ll = orc.sgd_forward(g["x_test"], S, g["soft_weights"], g["means"], g["l_diag"], g["l_lower"])
    assert np.max(np.abs(ll - g["ll_test"])) < 2e-5
    assert abs(float(g["mean_ll_test"]) - (-3.6001153)) < 1e-6
    r = golden("ref_robust_f32.npz")
    ll, (log_r, b, B), _ = orc.e_step(r["X"], r["S"], r["alpha"], r["mu"], r["Sigma"])
    assert abs(ll - float(r["ll_sum"])) < 1e-4 * abs(ll)
    assert np.isfinite(B).all()


def test_boston_config_trace():
    g = golden("ref_boston_k10_f64.npz")
    X = g["X"]
    N, D = X.shape
    K = 10
    S = np.zeros((N, D, D))
    S[:, np.arange(D), np.arange(D)] = g["noise_var"]
    alpha = np.full(K, 1.0 / K)
    mu = X[g["rows"]].copy()
    Sigma = np.tile(np.eye(D), (K, 1, 1))
    for it in range(1, 6):
        s0, s1, S2, ll_i = orc.suff_stats(X, S, alpha, mu, Sigma)
        assert abs(ll_i.sum() - g["ll_trace"][it - 1]) < 1e-8 * abs(g["ll_trace"][it - 1])
        wts, mu, Sigma = orc.m_step_from_stats(s0, s1, S2, mu, Sigma, N, float(g["w"]))
        alpha = wts[:, 0]
    assert rel(mu, g["it5_means"]) < 1e-7
    assert rel(Sigma, g["it5_covars"]) < 1e-6


def test_non_pd_returns_minus_inf():
    X = np.zeros((2, 2))
    S = np.tile(np.diag([1.0, -5.0]), (2, 1, 1))
    ll, e, _ = orc.e_step(X, S, np.ones(1), np.zeros((1, 2)), np.eye(2)[None])
    assert ll == -math.inf and e is None


# ---- real tables (BASELINE.json configs[2], configs[3]): the oracle tracks the reference's EM iterations ------------
def _real_noise(seed, N, D):
    """Same draw as tests/golden/make_golden.py::real_noise (S only)."""
    rng = np.random.default_rng(seed)
    Q = (0.5 / np.sqrt(D)) * rng.standard_normal((N, D, D))
    return np.einsum("nba,nbc->nac", Q, Q) + 1e-3 * np.eye(D)


def _tril_unpack(P, D):
    ia, ib = np.tril_indices(D)
    C = np.zeros(P.shape[:-1] + (D, D))
    C[..., ia, ib] = P
    C[..., ib, ia] = P
    return C


@pytest.mark.parametrize("name,iters", [("ref_wine_k128_f64.npz", 5), ("ref_ionosphere_k256_f64.npz", 5)])
def test_oracle_tracks_reference_em_on_the_real_tables(name, iters):
    """Injected init in the reference's style (alpha = 1/K, mu = K data rows, Sigma = I), full per-point noise, w = 1e-3:
    log-likelihood trace and parameters after 1 and 5 iterations of the float64 reference (stored as float32)."""
    g = golden(name)
    X = g["X64"]
    N, D = X.shape
    K = int(g["K"])
    S = _real_noise(int(g["seed"]), N, D)
    alpha, mu, Sigma = np.full(K, 1.0 / K), X[g["rows"]].copy(), np.tile(np.eye(D), (K, 1, 1))
    for it in range(1, iters + 1):
        s0, s1, S2, ll = orc.suff_stats(X, S, alpha, mu, Sigma)
        assert abs(ll.sum() - g["ll_trace"][it - 1]) <= 1e-9 * abs(g["ll_trace"][it - 1])
        w_new, mu, Sigma = orc.m_step_from_stats(s0, s1, S2, mu, Sigma, N, float(g["w"]))
        alpha = w_new[:, 0]
        if "it%d_means" % it in g.files:
            assert rel(alpha, g["it%d_weights" % it]) < 1e-6
            assert rel(mu, g["it%d_means" % it]) < 1e-6
            assert rel(Sigma, _tril_unpack(g["it%d_covars_tril" % it].astype(np.float64), D)) < 1e-6

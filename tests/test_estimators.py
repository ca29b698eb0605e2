"""Estimator-level parity with the reference's classes, on the golden vectors that
``tests/golden/make_golden.py`` recorded from the UNMODIFIED reference.

Every test runs twice: with the CPU test double (``fake_ops``: host logic only, part of the
``-m "not gpu"`` suite) and through the real C ABI on ``cuda:0`` (``-m gpu``).
Tolerances (BASELINE.json north_star): 1e-4 relative on log-likelihoods, 1e-3 on parameters."""
import math

import numpy as np
import pytest
import torch

from conftest import golden
from helpers import synth

LL_RTOL = 1e-4
P_RTOL = 1e-3


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


def close(got, want, rtol=P_RTOL):
    got = got.detach().cpu().numpy().astype(np.float64) if torch.is_tensor(got) else np.asarray(got, np.float64)
    want = want.detach().cpu().numpy() if torch.is_tensor(want) else want
    want = np.asarray(want, dtype=np.float64)
    assert got.shape == want.shape, (got.shape, want.shape)
    err = np.max(np.abs(got - want))
    assert err <= rtol * max(np.max(np.abs(want)), 1e-30), (err, np.max(np.abs(want)))


def em_model(cls, z, device, **kw):
    from density_deconvolution_b200 import estimators
    K, D = z["mu"].shape
    m = getattr(estimators, cls)(K, D, w=float(z["w"]), device=device, **kw)
    m.set_params(z["alpha"], z["mu"], z["Sigma"])
    return m, (t32(z["X"], device), t32(z["S"], device))


EM_CASES = ["ref_em_d3_k4_f64.npz", "ref_em_d10_k6_f64.npz", "ref_em_d10_k6_f32.npz", "ref_em_d16_k5_f64.npz"]


@pytest.mark.parametrize("name", EM_CASES)
def test_deconv_gmm_fit_iterations(device, name):
    z = golden(name)
    for it in (1, 3):
        if "it%d_means" % it not in z.files:
            continue
        m, data = em_model("DeconvGMM", z, device, epochs=it, tol=0.0)
        m.fit(data, init=False)
        assert m.n_iter_ == it
        assert m.weights.shape == (m.k, 1)
        close(m.weights, z["it%d_weights" % it])
        close(m.means, z["it%d_means" % it])
        close(m.covars, z["it%d_covars" % it])
        assert abs(m.log_prob_curve[0] - float(z["ll_sum"])) <= LL_RTOL * abs(float(z["ll_sum"]))
        ll = float(m.score(data).item())
        assert abs(ll - float(z["it%d_ll" % it])) <= 5 * LL_RTOL * abs(float(z["it%d_ll" % it]))


@pytest.mark.parametrize("name", EM_CASES)
def test_deconv_gmm_e_step_and_m_step_api(device, name):
    z = golden(name)
    m, data = em_model("DeconvGMM", z, device)
    ll, (log_r, b, B) = m._e_step(data)
    assert abs(float(ll) - float(z["ll_sum"])) <= LL_RTOL * abs(float(z["ll_sum"]))
    r_ref = np.exp(z["log_resps"].astype(np.float64))
    assert np.max(np.abs(np.exp(log_r.cpu().numpy().astype(np.float64)) - r_ref)) < 1e-4
    close(b, z["cond_means"])
    close(B, z["cond_covars"])
    close(m.predict(data), r_ref, rtol=1e-4)
    m._m_step(data, (log_r, b, B))
    close(m.weights, z["m1_weights"])
    close(m.means, z["m1_means"])
    close(m.covars, z["m1_covars"])


def test_boston_k10_diag_noise(device):
    """BASELINE.json configs[0]: boston (506 x 11), K = 10, per-point diagonal noise, injected
    init, 10 EM iterations; the diagonal layout [N,D] is passed straight to the kernel."""
    from density_deconvolution_b200 import DeconvGMM
    z = golden("ref_boston_k10_f64.npz")
    X, var, rows = z["X"], z["noise_var"], z["rows"]
    N, D = X.shape
    K = 10
    for it in (1, 5, 10):
        m = DeconvGMM(K, D, w=float(z["w"]), epochs=it, tol=0.0, device=device)
        m.set_params(np.full(K, 1.0 / K), X[rows], np.tile(np.eye(D), (K, 1, 1)))
        m.fit((t32(X, device), t32(var, device)), init=False)
        close(m.weights, z["it%d_weights" % it])
        close(m.means, z["it%d_means" % it])
        close(m.covars, z["it%d_covars" % it])
        trace = z["ll_trace"][:it]
        assert np.max(np.abs(np.asarray(m.log_prob_curve) - trace) / np.abs(trace)) <= LL_RTOL
    # the full [N,D,D] layout gives the same answer
    S = np.zeros((N, D, D))
    S[:, np.arange(D), np.arange(D)] = var
    m2 = DeconvGMM(K, D, w=float(z["w"]), epochs=10, tol=0.0, device=device)
    m2.set_params(np.full(K, 1.0 / K), X[rows], np.tile(np.eye(D), (K, 1, 1)))
    m2.fit((t32(X, device), t32(S, device)), init=False)
    close(m2.means, m.means, rtol=1e-5)


@pytest.mark.parametrize("name", ["ref_online_batch_d3_k4_f64.npz", "ref_online_batch_d5_k3_f32.npz"])
def test_online_m_step(device, name):
    from density_deconvolution_b200 import estimators
    z = golden(name)
    N = z["X"].shape[0]
    m, data = em_model("OnlineDeconvGMM", z, device, batch_size=N, step_size=float(z["step_size"]))
    m.resp_floor = 10 * float(z["eps"])
    m._reset_running_sums()
    ops = estimators._ops
    stats = None
    for tag, llkey in (("on", "ll_sum"), ("on2", "ll2_sum")):
        mix = m._mixture()
        _, stats = ops.estep_stats(data[0], data[1], mix.log_alpha, mix.mu, mix.Sigma, resp_floor=m.resp_floor)
        tail = stats[-4:].tolist()
        assert abs(tail[0] + tail[1] - float(z[llkey])) <= LL_RTOL * abs(float(z[llkey]))
        m._m_step_stats(stats, m.step_size)
        close(m.weights, z[tag + "_weights"])
        close(m.means, z[tag + "_means"])
        close(m.covars, z[tag + "_covars"])
        if tag == "on":
            close(m.sum_resps, z["on_sum_resps"])
            close(m.sum_cond_means, z["on_sum_cond_means"])
    # reference-signature _m_step on materialised expectations gives the same first step
    m2, _ = em_model("OnlineDeconvGMM", z, device, batch_size=N, step_size=float(z["step_size"]))
    m2.resp_floor = 10 * float(z["eps"])
    m2._reset_running_sums()
    _, e = m2._e_step(data)
    m2._m_step(e, N, m2.step_size)
    close(m2.means, z["on_means"])
    close(m2.covars, z["on_covars"])


@pytest.mark.parametrize("name", ["ref_online_batch_d3_k4_f64.npz", "ref_online_batch_d5_k3_f32.npz"])
def test_batch_epoch(device, name):
    z = golden(name)
    N = z["X"].shape[0]
    m, data = em_model("BatchDeconvGMM", z, device, batch_size=N // 2, epochs=1)
    # the reference's two passes, on the two halves
    half = N // 2
    parts = [(data[0][:half], data[1][:half]), (data[0][half:], data[1][half:])]
    ll, sr, sm = 0.0, 0.0, 0.0
    for p in parts:
        lp, (a, b) = m._step_1(p)
        ll, sr, sm = ll + float(lp), sr + a, sm + b
    assert abs(ll - float(z["bt_ll"])) <= LL_RTOL * abs(float(z["bt_ll"]))
    close(sr, z["bt_sum_resps"])
    close(sm, z["bt_sum_cond_means"])
    sc = sum(m._step_2(p, sm / sr) for p in parts)
    close(sc, z["bt_sum_cond_covars"])
    # one fused epoch
    m.fit(data, init=False)
    close(m.weights, z["bt_weights"])
    close(m.means, z["bt_means"])
    close(m.covars, z["bt_covars"])
    assert abs(m.train_ll_curve[0] - float(z["bt_ll"])) <= LL_RTOL * abs(float(z["bt_ll"]))


def sgd_model(z, device, **kw):
    from density_deconvolution_b200 import SGDDeconvGMM
    K, D = z["means"].shape
    m = SGDDeconvGMM(K, D, w=1e-3, batch_size=16, device=device, **kw)
    with torch.no_grad():
        m.module.soft_weights.copy_(t32(z["soft_weights"], device))
        m.module.means.copy_(t32(z["means"], device))
        m.module.l_diag.copy_(t32(z["l_diag"], device))
        m.module.l_lower.copy_(t32(z["l_lower"], device))
    return m


@pytest.mark.parametrize("name", ["ref_sgd_d4_k5_f64.npz", "ref_sgd_d7_k3_f32.npz"])
def test_sgd_forward_backward(device, name):
    z = golden(name)
    m = sgd_model(z, device)
    data = (t32(z["X"], device), t32(z["S"], device))
    close(m.covars, z["covars"], rtol=1e-5)
    assert abs(m.reg_loss(16, 64).item() - float(z["reg_loss"])) <= 1e-5 * abs(float(z["reg_loss"]))
    ll = m.module(data)
    assert ll.shape == (z["X"].shape[0],)
    assert np.max(np.abs(ll.detach().cpu().numpy() - z["ll"])) <= 2e-4 * np.max(np.abs(z["ll"]))
    (ll * t32(z["g"], device)).sum().backward()
    for p in ("soft_weights", "means", "l_diag", "l_lower"):
        close(getattr(m.module, p).grad, z["d_" + p])
    # the fused objective of the fit loop (sgd_gmm.py:173-181)
    for p in m.module.parameters():
        p.grad = None
    N = z["X"].shape[0]
    loss = -m.module.log_prob_sum(data) / N + m.reg_loss(N, 4 * N)
    assert abs(loss.item() - float(z["loss"])) <= LL_RTOL * abs(float(z["loss"]))
    loss.backward()
    for p in ("soft_weights", "means", "l_diag", "l_lower"):
        close(getattr(m.module, p).grad, z["loss_d_" + p])
    pw, pm, pc = m.posterior_params(data)
    assert np.max(np.abs(np.exp(pw.cpu().numpy()) - np.exp(z["post_log_w"]))) < 1e-4
    close(pm, z["post_means"])
    close(pc, z["post_covars"])
    assert abs(m.score_batch(data) - z["ll"].sum()) <= LL_RTOL * abs(z["ll"].sum())
    assert m.score(data).shape == ll.shape


def test_data_gen_fixture_shared_noise(device):
    """deconv/utils/data_gen.py fixture: one noise matrix for every point (layout [D,D])."""
    z = golden("ref_data_gen.npz")
    m = sgd_model(z, device)
    n = z["x_test"].shape[0]
    x = t32(z["x_test"], device)
    S_full = t32(np.broadcast_to(z["S"], (n, 2, 2)).copy(), device)
    with torch.no_grad():
        ll_full = m.module((x, S_full))
        ll_shared = m.module((x, t32(z["S"], device)))
    assert np.max(np.abs(ll_full.cpu().numpy() - z["ll_test"])) < 2e-4
    assert np.max(np.abs(ll_shared.cpu().numpy() - z["ll_test"])) < 2e-4
    pw, pm, pc = m.posterior_params((x, S_full))
    assert np.max(np.abs(np.exp(pw.cpu().numpy()) - np.exp(z["post_log_w"]))) < 1e-4
    close(pm, z["post_means"])
    close(pc, z["post_covars"])


def test_extreme_noise_stays_finite(device):
    z = golden("ref_robust_f32.npz")
    m, data = em_model_robust(z, device)
    ll, (log_r, b, B) = m._e_step(data)
    assert abs(float(ll) - float(z["ll_sum"])) <= LL_RTOL * abs(float(z["ll_sum"]))
    assert torch.isfinite(B).all() and torch.isfinite(b).all()
    assert np.max(np.abs(np.exp(log_r.cpu().numpy()) - np.exp(z["log_resps"]))) < 1e-4


def em_model_robust(z, device):
    from density_deconvolution_b200 import DeconvGMM
    K, D = z["mu"].shape
    m = DeconvGMM(K, D, device=device)
    m.set_params(z["alpha"], z["mu"], z["Sigma"])
    return m, (t32(z["X"], device), t32(z["S"], device))


def test_non_positive_definite_semantics(device):
    """EM: ``(-inf, None)`` and a clean stop (deconv_gmm.py:46-47, base.py:48-49).
    SGD: an exception, as torch's MultivariateNormal validation raises in the reference."""
    from density_deconvolution_b200 import DeconvGMM, SGDDeconvGMM
    rng = np.random.default_rng(3)
    X, S, alpha, mu, Sigma = synth(rng, 20, 3, 2)
    S[7] = -5.0 * np.eye(3)
    m = DeconvGMM(2, 3, epochs=5, device=device).set_params(alpha, mu, Sigma)
    data = (t32(X, device), t32(S, device))
    ll, e = m._e_step(data)
    assert float(ll) == -math.inf and e is None
    before = m.means.clone()
    m.fit(data, init=False)
    assert m.n_iter_ == 0 and torch.equal(m.means, before)
    assert float(m.score(data)) == -math.inf
    s = SGDDeconvGMM(2, 3, batch_size=20, epochs=1, device=device)
    with pytest.raises(ValueError):
        s.posterior_params(data)
    with pytest.raises(ValueError):
        s.fit(data, init=False)
    assert s.score_batch(data) == -math.inf


def test_fit_from_kmeans_improves_likelihood(device):
    from density_deconvolution_b200 import DeconvGMM, OnlineDeconvGMM, BatchDeconvGMM, SGDDeconvGMM, SGDDeconvDataset
    torch.manual_seed(0)
    rng = np.random.default_rng(5)
    X, S, alpha, mu, Sigma = synth(rng, 240, 2, 3, spread=6.0)
    data = (t32(X, device), t32(S, device))
    m = DeconvGMM(3, 2, epochs=8, w=1e-3, device=device)
    m.fit(data)
    assert all(b >= a - 1e-3 * abs(a) for a, b in zip(m.log_prob_curve, m.log_prob_curve[1:]))
    assert abs(m.weights.sum().item() - 1.0) < 1e-4
    ds = SGDDeconvDataset(*data)
    on = OnlineDeconvGMM(3, 2, epochs=3, batch_size=60, w=1e-3, k_means_iters=3, device=device).fit(ds)
    assert len(on.train_ll_curve) == 3 and math.isfinite(on.train_ll_curve[-1])
    bt = BatchDeconvGMM(3, 2, epochs=4, batch_size=60, w=1e-3, k_means_iters=3, device=device).fit(ds, val_data=ds)
    assert bt.train_ll_curve[-1] >= bt.train_ll_curve[0]
    assert len(bt.val_ll_curve) == 4
    sg = SGDDeconvGMM(3, 2, epochs=4, batch_size=60, lr=5e-2, k_means_iters=3, device=device).fit(ds, val_data=ds)
    assert sg.train_loss_curve[-1] > sg.train_loss_curve[0]
    assert sg.sample_prior(7).shape == (7, 2)
    assert sg.sample_posterior((data[0][:5], data[1][:5]), 4).shape == (5, 4, 2)
    assert sg.posterior_log_prob(data[0][:5], (data[0][:5], data[1][:5])).shape == (5, 5)   # [len(x), N]
    state = sg.module.state_dict()
    assert set(state) == {"soft_weights", "means", "l_diag", "l_lower"}

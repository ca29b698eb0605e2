"""End-to-end ``fit`` parity with the UNMODIFIED reference's own loops (``make_golden.py fit_api``):
``SGDGMM.fit`` (deconv/gmm/sgd_gmm.py:264-272, data[0] only) and the ``_tim`` fork
(deconv/gmm/sgd_gmm_tim.py:112-227, deconv/gmm/sgd_deconv_gmm_tim.py), both INCLUDING their minibatch-k-means
initialisation; plus the noise layouts the batch iterator must carry (one shared matrix, no noise).

Every test runs with the CPU test double (``fake_ops``) and through the C ABI on ``cuda:0`` (``-m gpu``).
Tolerances (north_star): 1e-4 relative on mean log-likelihoods, 1e-3 on fitted parameters."""
import logging
import math

import numpy as np
import pytest
import torch

from conftest import golden

PARAMS = ("soft_weights", "means", "l_diag", "l_lower")


@pytest.fixture(params=["oracle", pytest.param("cuda", marks=pytest.mark.gpu)])
def device(request, monkeypatch):
    from density_deconvolution_b200 import estimators, util
    if request.param == "oracle":
        import fake_ops
        monkeypatch.setattr(estimators, "_ops", fake_ops)
        monkeypatch.setattr(util, "_real_ops", fake_ops)
        return torch.device("cpu")
    return torch.device("cuda", 0)


def t32(a, device):
    return torch.tensor(np.asarray(a), dtype=torch.float32, device=device)


def rel(got, want):
    got = got.detach().cpu().numpy().astype(np.float64) if torch.is_tensor(got) else np.asarray(got, np.float64)
    want = np.asarray(want, np.float64)
    return np.max(np.abs(got - want)) / max(np.max(np.abs(want)), 1e-30)


def test_sgdgmm_fit_matches_reference(device):
    """``SGDGMM.fit`` must run (data[0] only, no noise tensor anywhere) and track the reference."""
    from density_deconvolution_b200 import SGDGMM
    z = golden("ref_fit_api.npz")
    X, Xv = t32(z["X"], device), t32(z["Xv"], device)
    for fused in (True, False):
        m = SGDGMM(3, 3, epochs=5, lr=2e-2, batch_size=16, w=1e-3, device=device)
        m.k_means_iters = 3
        m.shuffle = False
        m.fused_step = fused
        m.fit(torch.utils.data.TensorDataset(X), val_data=torch.utils.data.TensorDataset(Xv))
        np.testing.assert_allclose(m.train_loss_curve, z["plain_train_curve"], rtol=1e-4)
        np.testing.assert_allclose(m.val_loss_curve, z["plain_val_curve"], rtol=1e-4)
        for p in PARAMS:
            assert rel(getattr(m.module, p), z["plain_" + p]) < 1e-3, (p, fused)
        sb = m.score_batch(torch.utils.data.TensorDataset(X))
        assert abs(sb - float(z["plain_score_batch"])) <= 1e-4 * abs(float(z["plain_score_batch"]))
        assert m.score([X]).shape == (X.shape[0],)


def test_tim_fork_fit_matches_reference(device):
    from density_deconvolution_b200 import SGDDeconvGMMTim, SGDDeconvDataset
    z = golden("ref_fit_api.npz")
    ds = SGDDeconvDataset(t32(z["X"], device), t32(z["S"], device))
    dv = SGDDeconvDataset(t32(z["Xv"], device), t32(z["Sv"], device))
    msgs = []

    class Sink(logging.Handler):
        def emit(self, rec):
            msgs.append(rec.getMessage())
    logger = logging.getLogger("dgmm_test_tim_%s" % device.type)
    logger.setLevel(logging.INFO)
    logger.addHandler(Sink())
    t = SGDDeconvGMMTim(3, 3, epochs=4, lr=2e-2, batch_size=16, w=1e-3, k_means_iters=3, device=device)
    t.shuffle = False
    assert t.max_no_improvement == 30                                      # sgd_gmm_tim.py:68
    ret = t.fit(ds, logger, val_data=dv, verbose=True, interval=2)
    assert ret is None and bool(z["tim_returned_none"])
    np.testing.assert_allclose(t.train_loss_curve, z["tim_train_curve"], rtol=1e-4)   # the LOSS: -mean ll
    np.testing.assert_allclose(t.val_loss_curve, z["tim_val_curve"], rtol=1e-4)
    assert len(msgs) == int(z["tim_n_messages"]) and msgs[0].startswith("Epoch 0, Train Loss: 5.69")
    assert abs(t.optimiser.param_groups[0]["lr"] - float(z["tim_lr_final"])) < 1e-12   # scheduler never stepped
    for p in PARAMS:
        assert rel(getattr(t.module, p), z["tim_" + p]) < 1e-3, p
    # es_tl early stop (sgd_gmm_tim.py:187-189)
    t2 = SGDDeconvGMMTim(3, 3, epochs=50, lr=2e-2, batch_size=16, w=1e-3, k_means_iters=3, device=device)
    t2.shuffle = False
    t2.fit(ds, logger, es_tl=float(z["tim_es_tl"]))
    assert len(t2.train_loss_curve) == int(z["tim_es_epochs"])
    # log_prob_prior == the noise-free mixture density (sgd_deconv_gmm_tim.py:81-97)
    lp = t.log_prob_prior(ds.X[:7])
    mix = torch.distributions.MixtureSameFamily(
        torch.distributions.Categorical(probs=t.weights.cpu().double()),
        torch.distributions.MultivariateNormal(t.means.cpu().double(), covariance_matrix=t.covars.cpu().double()))
    assert rel(lp, mix.log_prob(ds.X[:7].cpu().double()).numpy()) < 1e-4


def test_shared_and_absent_noise_through_the_fit_loops(device):
    """one ``[D,D]`` noise matrix for all points (deconv/utils/data_gen.py:50-53) and ``S = None`` must
    run through the minibatch estimators and agree with the same fit on the repeated / zero ``[N,D,D]`` tensor."""
    from density_deconvolution_b200 import OnlineDeconvGMM, SGDDeconvGMM
    z = golden("ref_fit_api.npz")
    X = t32(z["X"], device)
    N, D = X.shape
    S1 = torch.tensor([[0.3, 0.05, 0.0], [0.05, 0.2, 0.02], [0.0, 0.02, 0.4]], device=device)
    for S_small, S_big in ((S1, S1.expand(N, D, D).contiguous()), (None, torch.zeros(N, D, D, device=device))):
        fits = []
        for S in (S_small, S_big):
            m = SGDDeconvGMM(3, D, epochs=2, lr=2e-2, batch_size=16, w=1e-3, k_means_iters=2, device=device)
            m.shuffle = False
            m.fit((X, S))
            fits.append(m)
            assert math.isfinite(m.score_batch((X, S)))
        for p in PARAMS:
            assert rel(getattr(fits[0].module, p), getattr(fits[1].module, p).detach().cpu().numpy()) < 1e-3, p
        np.testing.assert_allclose(fits[0].train_loss_curve, fits[1].train_loss_curve, rtol=1e-4)
    torch.manual_seed(3)
    on = OnlineDeconvGMM(3, D, epochs=2, batch_size=24, w=1e-3, k_means_iters=2, device=device).fit((X, S1))
    assert len(on.train_ll_curve) == 2 and math.isfinite(on.train_ll_curve[-1])


def test_score_is_float64(device):
    """``score`` keeps the kernel's float64 sum (a float32 cast makes consecutive epochs compare equal
    once |ll| > 1.7e7 and fires a spurious 'converged')."""
    from density_deconvolution_b200 import DeconvGMM
    z = golden("ref_fit_api.npz")
    m = DeconvGMM(3, 3, device=device)
    m.set_params(np.full(3, 1 / 3), z["X"][:3], np.tile(np.eye(3), (3, 1, 1)))
    s = m.score((t32(z["X"], device), t32(z["S"], device)))
    assert s.dtype == torch.float64 and math.isfinite(s.item())


def test_init_from_labels_has_no_dense_intermediate(device):
    """the k-means M-step works from labels (index_add), equal to the one-hot formula."""
    from density_deconvolution_b200 import DeconvGMM
    z = golden("ref_fit_api.npz")
    X = t32(z["X"], device)
    N, D = X.shape
    labels = torch.arange(N, device=device) % 3
    m = DeconvGMM(3, D, w=1e-3, device=device)
    m._init_from_labels(X, labels)
    R = torch.nn.functional.one_hot(labels, 3).double().cpu()
    Xd = X.double().cpu()
    s0 = R.sum(0)
    mu = (R.t() @ Xd) / s0[:, None]
    dev = Xd[:, None, :] - mu[None]
    cov = (torch.einsum("nk,nka,nkb->kab", R, dev, dev) + (s0[:, None, None] + 2e-3) * torch.eye(D).double()) / s0[:, None, None]
    assert rel(m.means, mu.numpy()) < 1e-5 and rel(m.covars, cov.numpy()) < 1e-5
    assert rel(m.weights.view(-1), (s0 / N).numpy()) < 1e-6

"""Minibatch-SGD step: ``dgmm_sgd_params_f32`` / ``dgmm_sgd_adam_step_f32`` and the
fused fit loop of ``SGDDeconvGMM`` against golden trajectories recorded from the UNMODIFIED
reference (its module, reg_loss, torch Adam and MultiStepLR; ``make_golden.py sgd_fit``).

Tolerances (north_star): 1e-4 relative on the mean log-likelihood, 1e-3 on fitted parameters."""
import numpy as np
import pytest
import torch

from conftest import golden
from oracle import gmm_oracle as orc

PARAMS = ("soft_weights", "means", "l_diag", "l_lower")


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


def rel(got, want):
    got = got.detach().cpu().numpy().astype(np.float64) if torch.is_tensor(got) else np.asarray(got, np.float64)
    want = np.asarray(want, np.float64)
    return np.max(np.abs(got - want)) / max(np.max(np.abs(want)), 1e-30)


def model(z, device, fused, **kw):
    from density_deconvolution_b200 import SGDDeconvGMM
    K, D = z["means"].shape
    m = SGDDeconvGMM(K, D, w=1e-3, batch_size=16, lr=2e-2, lr_step=2, device=device, **kw)
    m.shuffle = False           # the golden trajectory takes the batches in order
    m.fused_step = fused
    with torch.no_grad():
        for p in PARAMS:
            getattr(m.module, p).copy_(t32(z[p], device))
    return m


@pytest.mark.parametrize("name", ["ref_sgd_fit_d5_k3_f32.npz", "ref_sgd_fit_d3_k4_f64.npz"])
@pytest.mark.parametrize("fused", [True, False])
def test_fit_trajectory_matches_reference(device, name, fused):
    z = golden(name)
    data = (t32(z["X"], device), t32(z["S"], device))
    curve = z["train_loss_curve"]
    for epochs in (1, 3, len(curve)):
        m = model(z, device, fused, epochs=epochs)
        m.tol = 0.0
        m.fit(data, init=False)
        assert len(m.train_loss_curve) == epochs
        np.testing.assert_allclose(m.train_loss_curve, curve[:epochs], rtol=1e-4)
        for p in PARAMS:
            assert rel(getattr(m.module, p), z["%s_after_%d" % (p, epochs)]) < 1e-3, (p, epochs)
    assert abs(m.optimiser.param_groups[0]["lr"] - float(z["lr_final"])) < 1e-12   # MultiStepLR ran
    # the optimiser's own state was advanced (fused path writes torch's state tensors)
    st = m.optimiser.state[m.module.means]
    assert int(st["step"].item()) == len(curve) * (len(z["X"]) // 16)
    assert st["exp_avg_sq"].abs().sum().item() > 0


def test_fused_and_autograd_routes_agree(device):
    z = golden("ref_sgd_fit_d5_k3_f32.npz")
    data = (t32(z["X"], device), t32(z["S"], device))
    a = model(z, device, True, epochs=2).fit(data, init=False)
    b = model(z, device, False, epochs=2).fit(data, init=False)
    for p in PARAMS:
        assert rel(getattr(a.module, p), getattr(b.module, p).detach().cpu().numpy()) < 2e-4, p


# ---- the two kernels against the oracle ---------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("D,K", [(1, 3), (2, 1), (3, 4), (10, 64), (16, 128), (32, 256), (9, 200), (7, 512), (4, 1000)])
def test_adam_step_kernel(D, K):
    from density_deconvolution_b200 import ops
    dev = torch.device("cuda", 0)
    rng = np.random.default_rng(D * 100 + K)
    NL = D * (D - 1) // 2
    prm = [rng.standard_normal(K), 3 * rng.standard_normal((K, D)), 0.3 * rng.standard_normal((K, D)),
           0.5 * rng.standard_normal((K, NL))]
    m0 = [0.1 * rng.standard_normal(p.shape) for p in prm]
    v0 = [0.01 * rng.uniform(0.1, 1, size=p.shape) for p in prm]
    n, n_total = 4096.0, 100000.0
    s0 = rng.uniform(1, 50, size=K)
    s1 = rng.standard_normal((K, D)) * 20
    A = rng.standard_normal((K, D, D)) * 10
    S2 = A + np.swapaxes(A, 1, 2)
    stats = np.concatenate([s0, s1.ravel(), S2.ravel(), [-1234.5, 0.25, n, 0.0]])
    tp = [torch.tensor(p, dtype=torch.float32, device=dev) for p in prm]
    tm = [torch.tensor(p, dtype=torch.float32, device=dev) for p in m0]
    tv = [torch.tensor(p, dtype=torch.float32, device=dev) for p in v0]
    tstats = torch.tensor(stats, dtype=torch.float32, device=dev)
    la, Sg = ops.sgd_params(tp[0], tp[2], tp[3])
    alpha, _, Sigma, _ = orc.sgd_to_mixture(prm[0], prm[1], prm[2], prm[3])
    assert rel(la, np.log(alpha)) < 1e-5 and rel(Sg, Sigma) < 1e-5
    acc = torch.tensor([10.0, 0.0], dtype=torch.float64, device=dev)
    ops.sgd_adam_step(tstats, n, n_total, 1e-3, 1e-2, (0.9, 0.999), 1e-8, 7, tp, tm, tv, la, Sg, acc)
    torch.cuda.synchronize()
    f32 = lambda a: [np.asarray(x, np.float32).astype(np.float64) for x in a]  # noqa: E731
    new_p, new_m, new_v = orc.sgd_adam_step((s0, s1, S2), f32(prm), (f32(m0), f32(v0)), n, n_total, 1e-3, 1e-2,
                                            0.9, 0.999, 1e-8, 7)
    for got, want in zip(tp + tm + tv, new_p + new_m + new_v):
        if want.size:
            assert rel(got, want.reshape(got.shape)) < 2e-4
    alpha, _, Sigma, _ = orc.sgd_to_mixture(*new_p)
    assert rel(la, np.log(alpha)) < 2e-4 and rel(Sg, Sigma) < 2e-4
    assert torch.equal(Sg, Sg.transpose(1, 2))
    assert abs(acc[0].item() - (10.0 - 1234.5 + 0.25)) < 1e-3 and acc[1].item() == 0.0

"""Real-table parity for BASELINE.json configs[2]: white + red wine (6497 x 9, data_small/*wine*), K = 128, full per-point
noise, the reference's own style of injected init (alpha = 1/K, mu = K distinct data rows, Sigma = I:
deconv/gmm/util.py:42, online_deconv_gmm.py:32-36).  Golden = the UNMODIFIED reference in float64
(tests/golden/make_golden.py wine): parameters after 1 / 5 / 20 EM iterations, the log-likelihood trace, and a 2-epoch SGD
trajectory (its module / reg_loss / Adam / MultiStepLR, batches of 500 in order).
Tolerances (north_star): 1e-4 relative on the mean log-likelihood, 1e-3 on fitted parameters."""
import numpy as np
import pytest
import torch

from conftest import golden
from test_team_kernels import fit_against_golden, check_fit, real_noise, dev, LL_RTOL

pytestmark = pytest.mark.gpu


def test_wine_k128_em_iterations_1_5_20():
    from density_deconvolution_b200 import ops as o
    o.set_tuning()
    g = golden("ref_wine_k128_f64.npz")
    D = g["X64"].shape[1]
    fitted = fit_against_golden(g, (1, 5, 20), dev())
    check_fit(g, fitted, D)
    trace = g["ll_trace"]
    got = fitted[1][3] + fitted[5][3] + fitted[20][3]        # ll at the start of every one of the 20 iterations
    assert len(got) == 20
    for a, b in zip(got, trace):
        assert abs(a - b) <= LL_RTOL * abs(b)


def test_wine_k128_packed_layout_gives_the_same_fit():
    from density_deconvolution_b200 import DeconvGMM, ops
    g = golden("ref_wine_k128_f64.npz")
    X = g["X64"]
    N, D = X.shape
    K = int(g["K"])
    S = real_noise(int(g["seed"]), N, D)
    tX = torch.tensor(X, dtype=torch.float32, device=dev())
    tP = ops.pack_noise(torch.tensor(S, dtype=torch.float32, device=dev()))
    m = DeconvGMM(K, D, epochs=5, w=float(g["w"]), tol=0.0, device=dev())
    m.set_params(np.full((K, 1), 1.0 / K, dtype=np.float32), X[g["rows"]].astype(np.float32),
                 np.tile(np.eye(D, dtype=np.float32), (K, 1, 1)))
    m.fit((tX, tP), init=False)
    fitted = {5: (m.weights.reshape(-1).cpu().numpy().astype(np.float64), m.means.cpu().numpy().astype(np.float64),
                  m.covars.cpu().numpy().astype(np.float64), list(m.log_prob_curve))}
    check_fit(g, fitted, D)


@pytest.mark.parametrize("fused", [True, False])
def test_wine_k128_sgd_two_epochs(fused):
    from density_deconvolution_b200 import SGDDeconvGMM
    g = golden("ref_wine_k128_f64.npz")
    X = g["X64"]
    N, D = X.shape
    K = int(g["K"])
    S = real_noise(int(g["seed"]), N, D)
    tX = torch.tensor(X, dtype=torch.float32, device=dev())
    tS = torch.tensor(S, dtype=torch.float32, device=dev())
    m = SGDDeconvGMM(K, D, epochs=int(g["sgd_epochs"]), batch_size=int(g["sgd_batch"]), lr=float(g["sgd_lr"]),
                     w=float(g["w"]), device=dev())
    m.shuffle = False
    m.fused_step = fused
    m.tol = 0.0
    with torch.no_grad():
        m.module.soft_weights.zero_()
        m.module.means.copy_(torch.tensor(X[g["rows"]], dtype=torch.float32))
        m.module.l_diag.zero_()
        m.module.l_lower.zero_()
    m.fit((tX, tS), init=False)
    np.testing.assert_allclose(m.train_loss_curve, g["sgd_curve"], rtol=LL_RTOL)
    for p in ("soft_weights", "means", "l_diag", "l_lower"):
        got = getattr(m.module, p).detach().cpu().numpy().astype(np.float64)
        want = g["sgd_" + p].astype(np.float64)
        assert np.max(np.abs(got - want)) <= 1e-3 * max(np.max(np.abs(want)), 1e-2), p

"""Online-EM M-step kernel (``dgmm_online_mstep_f32``) against the float64 ``[K,.]``
algebra of deconv/gmm/online_deconv_gmm.py:146-188 (the test double, itself pinned to the reference's
golden vectors by tests/test_estimators.py), incl. the sticky crash flag (:79-88)."""
import numpy as np
import pytest
import torch

import fake_ops


def _case(D, K, seed, nonpd=0.0):
    rng = np.random.default_rng(seed)
    A = rng.standard_normal((K, D, D))
    Sigma = A @ np.swapaxes(A, 1, 2) + 0.5 * np.eye(D)
    mu = 3 * rng.standard_normal((K, D))
    s0 = rng.uniform(0.0, 30.0, size=K)
    s0[0] = 0.0                                            # a component nobody claims in this minibatch
    s1 = rng.standard_normal((K, D)) * s0[:, None] * 0.3
    B = rng.standard_normal((K, D, D)) * 0.2
    S2 = -(s0[:, None, None] * np.linalg.inv(Sigma + np.eye(D))) + B + np.swapaxes(B, 1, 2)
    stats = np.concatenate([s0, s1.ravel(), S2.ravel(), [-100.0, 0.0, 500.0, nonpd]])
    SR = rng.uniform(1.0, 20.0, size=(K, 1))
    SCM = mu * SR + rng.standard_normal((K, D))
    return stats, mu, Sigma, SR, SCM


@pytest.mark.gpu
@pytest.mark.parametrize("D,K", [(1, 2), (2, 3), (5, 7), (10, 64), (16, 128), (32, 256)])
def test_online_mstep_kernel(D, K):
    from density_deconvolution_b200 import ops
    dev = torch.device("cuda", 0)
    stats, mu, Sigma, SR, SCM = _case(D, K, seed=D + K)
    f = 10 * torch.finfo(torch.float32).eps

    def run(mod, device):
        t = lambda a: torch.tensor(a, dtype=torch.float32, device=device).contiguous()  # noqa: E731
        st, m, Sg, sr, scm = t(stats), t(mu), t(Sigma), t(SR), t(SCM)
        al, la = torch.empty(K, device=device), torch.empty(K, device=device)
        flag = torch.zeros(1, dtype=torch.int32, device=device)
        mod.online_mstep(st, f, 0.1, 1e-3, 500.0, m, Sg, sr.view(-1), scm, al, la, flag)
        return [x.cpu().double().numpy() for x in (m, Sg, sr, scm, al, la)] + [int(flag.item())]

    got, want = run(ops, dev), run(fake_ops, torch.device("cpu"))
    assert got[-1] == 0 and want[-1] == 0
    for g, w, name in zip(got[:-1], want[:-1], ("mu", "Sigma", "sum_resps", "sum_cond_means", "alpha", "log_alpha")):
        assert np.max(np.abs(g - w)) <= 1e-4 * max(1.0, np.max(np.abs(w))), name
    assert np.array_equal(got[1], np.swapaxes(got[1], 1, 2))


@pytest.mark.gpu
def test_online_mstep_crash_flag_is_sticky():
    from density_deconvolution_b200 import ops
    dev = torch.device("cuda", 0)
    D, K = 4, 5
    bad, mu, Sigma, SR, SCM = _case(D, K, seed=1, nonpd=3.0)
    good = _case(D, K, seed=1)[0]
    t = lambda a: torch.tensor(a, dtype=torch.float32, device=dev).contiguous()  # noqa: E731
    m, Sg, sr, scm = t(mu), t(Sigma), t(SR), t(SCM)
    al, la = torch.zeros(K, device=dev), torch.zeros(K, device=dev)
    flag = torch.zeros(1, dtype=torch.int32, device=dev)
    before = [x.clone() for x in (m, Sg, sr, scm)]
    ops.online_mstep(t(bad), 1e-6, 0.1, 1e-3, 500.0, m, Sg, sr.view(-1), scm, al, la, flag)
    assert flag.item() == 1
    ops.online_mstep(t(good), 1e-6, 0.1, 1e-3, 500.0, m, Sg, sr.view(-1), scm, al, la, flag)   # still frozen
    assert flag.item() == 1
    for x, b in zip((m, Sg, sr, scm), before):
        assert torch.equal(x, b)
    flag.zero_()
    ops.online_mstep(t(good), 1e-6, 0.1, 1e-3, 500.0, m, Sg, sr.view(-1), scm, al, la, flag)
    assert flag.item() == 0 and not torch.equal(m, before[0])

"""The C-ABI library loads without a GPU and exports every symbol include/dgmm_b200.h declares;
the torch-CPU port used as the timed CPU baseline matches the reference's golden vectors; the
product path fails loudly without CUDA tensors.  GPU part: the host-buffer entry points."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

from conftest import ROOT, golden
from helpers import synth


def header_symbols():
    text = open(os.path.join(ROOT, "include", "dgmm_b200.h")).read()
    return sorted(set(re.findall(r"DGMM_API\s+[\w\s\*]+?\b(dgmm_\w+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from density_deconvolution_b200 import _lib
    syms = header_symbols()
    assert len(syms) >= 15
    assert sorted(_lib.SYMBOLS) == syms
    if not os.path.exists(_lib.LIB_PATH):
        import density_deconvolution_b200 as pkg
        pkg.build()
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for s in syms:
        assert hasattr(lib, s), s
    lib.dgmm_abi_version.restype = ctypes.c_int
    assert lib.dgmm_abi_version() == 2
    n = ctypes.c_int(0)
    dims = (ctypes.c_int * 32)()
    assert lib.dgmm_supported_dims(dims, ctypes.byref(n)) == 0
    assert n.value == 32 and list(dims) == list(range(1, 33))


def test_no_cpu_fallback():
    from density_deconvolution_b200 import ops, DgmmError, DeconvGMM
    X = torch.zeros(4, 2)
    S = torch.eye(2).repeat(4, 1, 1)
    with pytest.raises(DgmmError):
        ops.loglik(X, S, torch.zeros(3), torch.zeros(3, 2), torch.eye(2).repeat(3, 1, 1))
    with pytest.raises(DgmmError):
        DeconvGMM(3, 2, device="cpu").fit((X, S), init=False)


def test_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "density_deconvolution_b200")
    for name in os.listdir(pkg):
        if name.endswith(".py"):
            src = open(os.path.join(pkg, name)).read()
            assert not re.search(r"^\s*(from|import)\s+oracle", src, re.M), name


@pytest.mark.parametrize("name", ["ref_em_d10_k6_f64.npz", "ref_em_d3_k4_f64.npz", "ref_em_d16_k5_f64.npz"])
def test_torch_port_matches_reference(name):
    from oracle import torch_port as tp
    z = golden(name)
    f = lambda a: torch.tensor(np.asarray(a), dtype=torch.float64)  # noqa: E731
    X, S, al, mu, Sg = f(z["X"]), f(z["S"]), f(z["alpha"]), f(z["mu"]), f(z["Sigma"])
    ll, (log_r, b, B) = tp.e_step(X, S, al, mu, Sg)
    assert abs(float(ll) - float(z["ll_sum"])) < 1e-9 * abs(float(z["ll_sum"]))
    np.testing.assert_allclose(log_r.numpy(), z["log_resps"], atol=1e-9)
    np.testing.assert_allclose(b.numpy(), z["cond_means"], atol=1e-9)
    np.testing.assert_allclose(B.numpy(), z["cond_covars"], atol=1e-9)
    for fn in (tp.em_iteration, tp.em_iteration_single_pass):
        ll2, w, m, c = fn(X, S, al, mu, Sg, float(z["w"]), chunk=5)
        assert abs(ll2 - float(z["ll_sum"])) < 1e-9 * abs(float(z["ll_sum"]))
        np.testing.assert_allclose(w.numpy(), z["it1_weights"], atol=1e-10)
        np.testing.assert_allclose(m.numpy(), z["it1_means"], atol=1e-9)
        np.testing.assert_allclose(c.numpy(), z["it1_covars"], atol=1e-9)
    assert abs(tp.loglik_sum(X, S, al, mu, Sg, chunk=7) - float(z["ll_sum"])) < 1e-9 * abs(float(z["ll_sum"]))


# ---------------------------------------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("noise", ["full", "diag", "shared"])
def test_host_buffer_entry_points(noise):
    """dgmm_em_iteration_host / dgmm_estep_stats_host / dgmm_score_host on HOST buffers, several
    ragged chunks, against the oracle."""
    from density_deconvolution_b200 import ops
    from oracle import gmm_oracle as orc
    rng = np.random.default_rng(21)
    N, D, K = 1000, 5, 6
    X, S, alpha, mu, Sigma = synth(rng, N, D, K)
    if noise == "diag":
        d = np.stack([np.diag(s) for s in S])
        S_or, S_in = np.stack([np.diag(v) for v in d]), d
    elif noise == "shared":
        S_or, S_in = np.broadcast_to(S[0], S.shape).copy(), S[0]
    else:
        S_or, S_in = S, S
    f = lambda a: torch.tensor(np.ascontiguousarray(a), dtype=torch.float32)  # noqa: E731
    hX, hS, ha, hm, hSg = f(X), f(S_in), f(alpha), f(mu), f(Sigma)
    ctx = ops.HostContext(D, K, chunk_points=300, device=0)
    tot, bad, ll = ctx.score(hX, hS, ha, hm, hSg, noise=noise, want_ll=True)
    ll_ref = orc.loglik(X, S_or, alpha, mu, Sigma)
    assert bad == 0 and abs(tot - ll_ref.sum()) <= 1e-4 * abs(ll_ref.sum())
    assert np.max(np.abs(ll.numpy() - ll_ref)) < 2e-4 * np.max(np.abs(ll_ref))
    stats = ctx.estep_stats(hX, hS, ha, hm, hSg, noise=noise)
    s0, s1, S2, _ = orc.suff_stats(X, S_or, alpha, mu, Sigma)
    d0, d1, d2, tail = ops.unpack_stats(stats, D, K)
    assert np.max(np.abs(d0.numpy() - s0)) < 1e-3 * s0.max()
    assert np.max(np.abs(d2.numpy() - S2)) < 1e-3 * np.abs(S2).max()
    assert abs(tail[2].item() - N) < 1e-3 and tail[3].item() == 0
    ll_sum, bad = ctx.em_iteration(hX, hS, 1e-3, ha, hm, hSg, noise=noise)
    w_ref, mu_ref, Sg_ref = orc.m_step_from_stats(s0, s1, S2, mu, Sigma, N, 1e-3)
    assert bad == 0 and abs(ll_sum - ll_ref.sum()) <= 1e-4 * abs(ll_ref.sum())
    assert np.max(np.abs(ha.numpy() - w_ref[:, 0])) < 1e-3 * w_ref.max()
    assert np.max(np.abs(hm.numpy() - mu_ref)) < 1e-3 * np.abs(mu_ref).max()
    assert np.max(np.abs(hSg.numpy() - Sg_ref)) < 1e-3 * np.abs(Sg_ref).max()
    # a non-PD point: -inf, count reported, parameters untouched
    hS_bad = f(S.copy())
    hS_bad[17] = -4.0 * torch.eye(D)
    if noise == "full":
        before = hm.clone()
        ll_sum, bad = ctx.em_iteration(hX, hS_bad, 1e-3, ha, hm, hSg, noise="full")
        assert ll_sum == -np.inf and bad > 0 and torch.equal(before, hm)
    ctx.close()

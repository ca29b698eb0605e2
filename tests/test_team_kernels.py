"""GPU parity tests of the cooperative (8-lane team) kernels that serve D = 17..32
(csrc/dgmm_team.cuh) against the float64 oracle, through the C ABI.

Every dimension, component counts that give one / several / ragged component chunks, all noise
layouts, per-point weights and the responsibility floor, ragged and unaligned inputs, non-PD input,
and the real ionosphere table (BASELINE.json configs[3]: 351 x 32, K = 256, 5 EM iterations against
the float64 run of the unmodified reference).
"""
import math

import numpy as np
import pytest
import torch

from oracle import gmm_oracle as orc
from helpers import synth
from conftest import golden
from test_gpu_ops import run_case, to_dev, dev, LL_RTOL, STAT_RTOL

pytestmark = pytest.mark.gpu


@pytest.fixture()
def ops():
    from density_deconvolution_b200 import ops as o
    o.set_tuning(rc=3)          # force the team path (it is the default from D = 21 and for K < 64 above D = 16)
    yield o
    o.set_tuning()


@pytest.mark.parametrize("D", list(range(17, 33)))
def test_every_dimension(ops, D):
    for mode in (0, 1):
        assert ops.get_plan(mode, D, 7, 500)["sigp"] == 3
    run_case(ops, N=61, D=D, K=7, seed=900 + D, with_g=(D % 3 == 0))


@pytest.mark.parametrize("D,K", [(32, 1), (32, 4), (32, 5), (32, 31), (32, 32), (32, 33), (32, 100), (32, 256),
                                 (24, 64), (24, 129), (21, 40), (17, 200), (29, 96), (26, 37)])
def test_component_chunks(ops, D, K):
    plan = ops.get_plan(1, D, K, 1000)
    assert plan["GS"] % 4 == 0 and plan["GS"] * plan["NSLOT"] >= K and plan["TPB"] == 8 * plan["GS"]
    run_case(ops, N=40 if K > 64 else 90, D=D, K=K, seed=1000 + D + K)


@pytest.mark.parametrize("N", [1, 2, 3, 4, 5, 7, 8, 9, 13, 300])
def test_ragged_point_counts(ops, N):
    run_case(ops, N=N, D=25, K=6, seed=1100 + N)


@pytest.mark.parametrize("D,K,noise", [(32, 12, "diag"), (23, 40, "diag"), (18, 9, "diag"), (30, 64, "diag")])
def test_diag_noise(ops, D, K, noise):
    run_case(ops, N=70, D=D, K=K, seed=1200 + D + K, noise=noise, with_g=True)


def test_weights_and_floor(ops):
    run_case(ops, N=80, D=32, K=20, seed=1301, with_g=True)
    run_case(ops, N=80, D=27, K=20, seed=1302, floor=10 * float(np.finfo(np.float32).eps))


def test_default_path_selection():
    from density_deconvolution_b200 import ops as o
    o.set_tuning()
    for D, K, want in [(32, 256, 3), (21, 64, 3), (24, 128, 3), (17, 8, 3), (20, 32, 3), (16, 128, None), (20, 128, None)]:
        for mode in (0, 1):
            plan = o.get_plan(mode, D, K, 10000)
            if want is None:
                assert plan["sigp"] != 3 or plan["rc"] != 3, (D, K, plan)
                assert plan["rc"] in (0, 1), (D, K, plan)
            else:
                assert plan["rc"] == want, (D, K, plan)


def test_shared_and_none_layouts(ops):
    rng = np.random.default_rng(1400)
    D, K, N = 28, 9, 50
    X, S, alpha, mu, Sigma = synth(rng, N, D, K)
    S1 = S[0]
    tX, tS1, tmu, tSg = to_dev(X, S1, mu, Sigma)
    tla = torch.tensor(np.log(alpha), dtype=torch.float32, device=dev())
    for layout, S_dev, S_or in (("shared", tS1, np.broadcast_to(S1, (N, D, D))), ("none", None, np.zeros((N, D, D)))):
        ll = ops.loglik(tX, S_dev, tla, tmu, tSg, noise=layout).cpu().numpy().astype(np.float64)
        _, stats = ops.estep_stats(tX, S_dev, tla, tmu, tSg, noise=layout)
        s0, s1, S2, ll_o = orc.suff_stats(X, S_or, alpha, mu, Sigma)
        assert abs(ll.mean() - ll_o.mean()) <= LL_RTOL * abs(ll_o.mean())
        d0, d1, d2, tail = [t.cpu().numpy().astype(np.float64) for t in ops.unpack_stats(stats, D, K)]
        assert np.max(np.abs(d0 - s0)) <= STAT_RTOL * s0.max()
        assert np.max(np.abs(d2 - S2)) <= STAT_RTOL * np.abs(S2).max()
        assert np.max(np.abs(d1 - s1)) <= STAT_RTOL * np.abs(s1).max()


def test_unaligned_inputs_take_the_cooperative_load_path(ops):
    """Base pointers that are not 16-byte aligned cannot use the bulk copies."""
    rng = np.random.default_rng(1500)
    D, K, N = 21, 10, 37
    X, S, alpha, mu, Sigma = synth(rng, N, D, K)
    bufX = torch.empty(N * D + 1, dtype=torch.float32, device=dev())
    bufS = torch.empty(N * D * D + 1, dtype=torch.float32, device=dev())
    tX = bufX[1:].view(N, D)
    tS = bufS[1:].view(N, D, D)
    tX.copy_(torch.tensor(X, dtype=torch.float32))
    tS.copy_(torch.tensor(S, dtype=torch.float32))
    assert tX.data_ptr() % 16 != 0 and tS.data_ptr() % 16 != 0
    _, tmu, tSg = to_dev(X, mu, Sigma)
    tla = torch.tensor(np.log(alpha), dtype=torch.float32, device=dev())
    ll, stats = ops.estep_stats(tX, tS, tla, tmu, tSg)
    s0, s1, S2, ll_o = orc.suff_stats(X, S, alpha, mu, Sigma)
    assert np.max(np.abs(ll.cpu().numpy() - ll_o)) <= 2e-4 * np.abs(ll_o).max()
    d0, d1, d2, _ = [t.cpu().numpy().astype(np.float64) for t in ops.unpack_stats(stats, D, K)]
    assert np.max(np.abs(d0 - s0)) <= STAT_RTOL * s0.max()
    assert np.max(np.abs(d2 - S2)) <= STAT_RTOL * np.abs(S2).max()


def test_non_pd_pair_poisons_its_point_only(ops):
    rng = np.random.default_rng(1600)
    D, K, N = 32, 5, 12
    X, S, alpha, mu, Sigma = synth(rng, N, D, K)
    S_bad = S.copy()
    S_bad[3] = -50.0 * np.eye(D)            # Sigma_j + S_3 indefinite for every j
    tX, tS, tmu, tSg = to_dev(X, S_bad, mu, Sigma)
    tla = torch.tensor(np.log(alpha), dtype=torch.float32, device=dev())
    ll, sums = ops.loglik(tX, tS, tla, tmu, tSg, want_sum=True)
    ll = ll.cpu().numpy()
    assert ll[3] == -np.inf and np.all(np.isfinite(np.delete(ll, 3)))
    assert sums[1].item() >= 1
    ll2, stats = ops.estep_stats(tX, tS, tla, tmu, tSg)
    tail = stats[-4:].cpu().numpy()
    assert tail[3] >= 1 and not np.isfinite(tail[0] + tail[1])
    # the remaining points still give finite statistics equal to the oracle's on the good points
    keep = np.arange(N) != 3
    s0, s1, S2, _ = orc.suff_stats(X[keep], S[keep], alpha, mu, Sigma)
    d0, d1, d2, _ = [t.cpu().numpy().astype(np.float64) for t in ops.unpack_stats(stats, D, K)]
    assert np.all(np.isfinite(d2))
    assert np.max(np.abs(d0 - s0)) <= STAT_RTOL * s0.max()
    assert np.max(np.abs(d2 - S2)) <= STAT_RTOL * np.abs(S2).max()


def test_large_dynamic_range_noise(ops):
    """Missing-error fill of the Gaia preprocessing: variance 1e12 on some diagonal entries."""
    rng = np.random.default_rng(1700)
    D, K, N = 32, 6, 16
    X, S, alpha, mu, Sigma = synth(rng, N, D, K)
    for i in range(0, N, 2):
        S[i, i % D, i % D] += 1e12
        S[i, (i + 5) % D, (i + 5) % D] += 1e8
    tX, tS, tmu, tSg = to_dev(X, S, mu, Sigma)
    tla = torch.tensor(np.log(alpha), dtype=torch.float32, device=dev())
    ll, stats = ops.estep_stats(tX, tS, tla, tmu, tSg)
    s0, s1, S2, ll_o = orc.suff_stats(X, S, alpha, mu, Sigma)
    ll = ll.cpu().numpy().astype(np.float64)
    assert np.all(np.isfinite(ll))
    assert np.max(np.abs(ll - ll_o)) <= 2e-4 * np.abs(ll_o).max()
    d0, d1, d2, _ = [t.cpu().numpy().astype(np.float64) for t in ops.unpack_stats(stats, D, K)]
    assert np.max(np.abs(d0 - s0)) <= STAT_RTOL * s0.max()
    assert np.max(np.abs(d2 - S2)) <= 2 * STAT_RTOL * np.abs(S2).max()


def test_host_buffer_pipeline_on_the_team_path():
    """dgmm_em_iteration_host / dgmm_score_host at D > 16 (chunked upload, ll scratch inside the context)."""
    from density_deconvolution_b200 import ops as o
    o.set_tuning()
    rng = np.random.default_rng(1800)
    D, K, N = 24, 10, 300
    X, S, alpha, mu, Sigma = synth(rng, N, D, K)
    hX, hS = torch.tensor(X, dtype=torch.float32), torch.tensor(S, dtype=torch.float32)
    ha = torch.tensor(alpha, dtype=torch.float32)
    hm = torch.tensor(mu, dtype=torch.float32)
    hSg = torch.tensor(Sigma, dtype=torch.float32)
    ctx = o.HostContext(D, K, chunk_points=128, device=0)
    tot, bad = ctx.score(hX, hS, ha, hm, hSg)
    ll_o = orc.loglik(X, S, alpha, mu, Sigma)
    assert bad == 0 and abs(tot - ll_o.sum()) <= LL_RTOL * abs(ll_o.sum())
    ll_sum, bad = ctx.em_iteration(hX, hS, 1e-3, ha, hm, hSg)
    assert bad == 0 and abs(ll_sum - ll_o.sum()) <= LL_RTOL * abs(ll_o.sum())
    s0, s1, S2, _ = orc.suff_stats(X, S, alpha, mu, Sigma)
    w_o, m_o, c_o = orc.m_step_from_stats(s0, s1, S2, mu, Sigma, N, 1e-3)
    assert np.max(np.abs(ha.numpy() - w_o[:, 0])) <= 1e-3 * w_o.max()
    assert np.max(np.abs(hm.numpy() - m_o)) <= 1e-3 * np.abs(m_o).max()
    assert np.max(np.abs(hSg.numpy() - c_o)) <= 1e-3 * np.abs(c_o).max()
    ctx.close()


# ---------------------------------------------------------------------------------------------------
# real data: BASELINE.json configs[3]
# ---------------------------------------------------------------------------------------------------
def real_noise(seed, N, D):
    """Same draw as tests/golden/make_golden.py::real_noise (S only)."""
    rng = np.random.default_rng(seed)
    Q = (0.5 / np.sqrt(D)) * rng.standard_normal((N, D, D))
    return np.einsum("nba,nbc->nac", Q, Q) + 1e-3 * np.eye(D)


def tril_unpack(P, D):
    ia, ib = np.tril_indices(D)
    C = np.zeros(P.shape[:-1] + (D, D), dtype=np.float64)
    C[..., ia, ib] = P
    C[..., ib, ia] = P
    return C


def fit_against_golden(g, iters, device):
    from density_deconvolution_b200.estimators import DeconvGMM
    X = g["X64"]
    N, D = X.shape
    K = int(g["K"])
    S = real_noise(int(g["seed"]), N, D)
    rows = g["rows"]
    tX = torch.tensor(X, dtype=torch.float32, device=device)
    tS = torch.tensor(S, dtype=torch.float32, device=device)
    m = DeconvGMM(K, D, epochs=max(iters), w=float(g["w"]), tol=0.0, device=device)
    m.set_params(np.full((K, 1), 1.0 / K, dtype=np.float32), X[rows].astype(np.float32),
                 np.tile(np.eye(D, dtype=np.float32), (K, 1, 1)))
    done = 0
    out = {}
    for it in sorted(iters):
        m.epochs = it - done
        m.fit((tX, tS), init=False)
        done = it
        out[it] = (m.weights.reshape(-1).cpu().numpy().astype(np.float64), m.means.cpu().numpy().astype(np.float64),
                   m.covars.cpu().numpy().astype(np.float64), list(m.log_prob_curve))
    return out


def check_fit(g, fitted, D):
    """1e-3 on fitted parameters (north_star), per component on its own scale for the means / covariances."""
    for it, (w, mu, cov, curve) in fitted.items():
        w_ref = g["it%d_weights" % it].astype(np.float64)
        mu_ref = g["it%d_means" % it].astype(np.float64)
        cov_ref = tril_unpack(g["it%d_covars_tril" % it].astype(np.float64), D)
        assert np.max(np.abs(w - w_ref)) <= 1e-3 * w_ref.max(), (it, "weights")
        sd = np.sqrt(np.einsum("kaa->ka", cov_ref))                 # per-component scale of a mean shift
        assert np.max(np.abs(mu - mu_ref) / sd) <= 1e-3 * max(1.0, np.abs(mu_ref / sd).max()), (it, "means")
        scale = np.sqrt(np.einsum("kaa,kbb->kab", cov_ref, cov_ref))  # |C_ab| <= sqrt(C_aa C_bb)
        assert np.max(np.abs(cov - cov_ref) / scale) <= 1e-3, (it, "covars", float(np.max(np.abs(cov - cov_ref) / scale)))


def test_ionosphere_k256_five_em_iterations():
    """351 x 32 points, K = 256, full per-point noise, injected init (alpha = 1/K, mu = data rows, Sigma = I),
    w = 1e-3: parameters after 1 and 5 iterations and the log-likelihood trace of the float64 reference."""
    from density_deconvolution_b200 import ops as o
    o.set_tuning()
    g = golden("ref_ionosphere_k256_f64.npz")
    D = g["X64"].shape[1]
    fitted = fit_against_golden(g, (1, 5), dev())
    check_fit(g, fitted, D)
    curve = fitted[5][3]                         # ll at the start of each of the last 4 iterations
    trace = g["ll_trace"]
    assert abs(fitted[1][3][0] - trace[0]) <= LL_RTOL * abs(trace[0])
    for got, want in zip(curve, trace[1:5]):
        assert abs(got - want) <= LL_RTOL * abs(want)

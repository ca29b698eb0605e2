"""GPU parity tests of the C-ABI entry points against the CPU oracle (float64 truth).

Tolerances (BASELINE.json north_star): relative 1e-4 on the mean log-likelihood, 1e-3 on
fitted parameters.  Per-point / per-statistic checks below are tighter than that.
"""
import math

import numpy as np
import pytest
import torch

from oracle import gmm_oracle as orc
from helpers import synth, rel
from conftest import golden

pytestmark = pytest.mark.gpu

LL_RTOL = 1e-4      # on the mean log-likelihood
STAT_RTOL = 1e-3    # on statistics / parameters, relative to the largest entry


def dev():
    return torch.device("cuda", 0)


def to_dev(*arrs):
    return [torch.tensor(np.ascontiguousarray(a), dtype=torch.float32, device=dev()) for a in arrs]


def check_stats_per_component(d0, d1, d2, s0, s1, S2, A1, A2, rtol=STAT_RTOL):
    """Every component on its own scale: a small component must be right too (the global checks above
    are relative to the LARGEST entry).  A1 / A2 are the absolute sums behind s1 / S2 (their float32
    accumulation error scales with them, also where the signed sum cancels)."""
    assert np.all(np.abs(d0 - s0) <= rtol * s0 + 1e-30), np.max(np.abs(d0 - s0) / (s0 + 1e-300))
    e1 = np.abs(d1 - s1).max(axis=1) / (A1.max(axis=1) + 1e-300)
    e2 = np.abs(d2 - S2).max(axis=(1, 2)) / (A2.max(axis=(1, 2)) + 1e-300)
    assert e1.max() <= rtol, ("s1", int(e1.argmax()), float(e1.max()))
    assert e2.max() <= rtol, ("S2", int(e2.argmax()), float(e2.max()))


def run_case(ops, N, D, K, seed, noise="full", with_g=False, floor=0.0):
    rng = np.random.default_rng(seed)
    X, S, alpha, mu, Sigma = synth(rng, N, D, K)
    g = rng.uniform(0.5, 1.5, size=N) if with_g else None
    if noise == "diag":
        Sd = np.stack([np.diag(np.diag(s)) for s in S])
        S_or, S_dev = Sd, np.stack([np.diag(s) for s in S])
    elif noise == "packed":
        ia, ib = np.tril_indices(D)
        S_or, S_dev = S, S[:, ia, ib]
    else:
        S_or, S_dev = S, S
    tX, tS, tmu, tSg = to_dev(X, S_dev, mu, Sigma)
    tla = torch.tensor(np.log(alpha), dtype=torch.float32, device=dev())
    tg = torch.tensor(g, dtype=torch.float32, device=dev()) if with_g else None
    ll_d, sums = ops.loglik(tX, tS, tla, tmu, tSg, noise=noise, want_sum=True)
    ll2_d, stats = ops.estep_stats(tX, tS, tla, tmu, tSg, g=tg, resp_floor=floor, noise=noise)
    torch.cuda.synchronize()
    s0, s1, S2, ll = orc.suff_stats(X, S_or, alpha, mu, Sigma, g=g, resp_floor=floor)
    ll_d = ll_d.cpu().numpy().astype(np.float64)
    assert np.all(np.isfinite(ll_d))
    assert abs(ll_d.mean() - ll.mean()) <= LL_RTOL * abs(ll.mean())
    assert np.max(np.abs(ll_d - ll)) <= 2e-4 * max(1.0, np.max(np.abs(ll)))
    # the two passes sum the K densities in different orders (warp butterflies vs the per-tile pass of
    # the log-likelihood kernel): equal to float32 rounding of the logsumexp, not bit for bit
    np.testing.assert_allclose(ll2_d.cpu().numpy(), ll_d.astype(np.float32), rtol=2e-6, atol=2e-6,
                               err_msg="stats pass and loglik pass disagree")
    assert abs(sums[0].item() - ll.sum()) <= LL_RTOL * abs(ll.sum())
    assert sums[1].item() == 0
    d0, d1, d2, tail = [t.cpu().numpy().astype(np.float64) for t in ops.unpack_stats(stats, D, K)]
    assert rel(d0, s0) < STAT_RTOL
    assert np.max(np.abs(d1 - s1)) < STAT_RTOL * max(np.abs(s1).max(), np.abs(s0).max() * 1e-2)
    assert np.max(np.abs(d2 - S2)) < STAT_RTOL * max(np.abs(S2).max(), np.abs(s0).max() * 1e-2)
    assert np.max(np.abs(d2 - np.swapaxes(d2, 1, 2))) == 0.0
    assert abs((tail[0] + tail[1]) - ll.sum()) <= LL_RTOL * abs(ll.sum())
    assert abs(tail[2] - (g.sum() if with_g else N)) <= 1e-5 * N
    assert tail[3] == 0
    A1, A2 = orc.suff_stats_scales(X, S_or, alpha, mu, Sigma, g=g, resp_floor=floor)
    check_stats_per_component(d0, d1, d2, s0, s1, S2, A1, A2)
    return dict(X=X, S=S_or, alpha=alpha, mu=mu, Sigma=Sigma, stats=stats, s=(s0, s1, S2), ll=ll)


@pytest.fixture(scope="module")
def ops():
    from density_deconvolution_b200 import ops as o
    o.set_tuning()
    return o


@pytest.mark.parametrize("D", list(range(1, 17)))
def test_dims_unrolled(ops, D):
    run_case(ops, N=301, D=D, K=5, seed=100 + D)


@pytest.mark.parametrize("D", [17, 24, 32])
def test_dims_above_16_small_k(ops, D):
    """D > 16 with few components: the cooperative team kernels (tests/test_team_kernels.py has the sweep)."""
    run_case(ops, N=67, D=D, K=4, seed=200 + D)


@pytest.mark.parametrize("D,K,noise", [(17, 64, "full"), (18, 128, "diag"), (19, 100, "full"), (20, 128, "full"),
                                       (20, 64, "diag"), (17, 128, "diag"), (20, 256, "full"), (19, 200, "diag")])
def test_dims_unrolled_with_spills(ops, D, K, noise):
    """D = 17..20 with 64+ components: the fully unrolled kernels (D >= 21 is served by the cooperative kernels only)."""
    ops.set_tuning(rc=4)     # the thread-per-pair family (D >= 21 goes to the team kernels by default)
    try:
        for mode in (0, 1):
            plan = ops.get_plan(mode, D, K, 1000, noise=noise)
            assert plan["rc"] in (0, 1), plan   # an unrolled placement, not the cooperative path (rc == 3)
        run_case(ops, N=203 if K <= 128 else 75, D=D, K=K, seed=250 + D + K, noise=noise, with_g=(D % 2 == 0))
    finally:
        ops.set_tuning()


@pytest.mark.parametrize("K", [1, 2, 3, 8, 10, 31, 32, 33, 64, 100, 128, 256])
def test_component_counts(ops, K):
    run_case(ops, N=257 if K <= 64 else 96, D=4, K=K, seed=300 + K)


@pytest.mark.parametrize("N", [1, 3, 4, 5, 15, 16, 17, 63, 1000])
def test_ragged_point_counts(ops, N):
    run_case(ops, N=N, D=5, K=6, seed=400 + N)


def test_empty_input(ops):
    D, K = 3, 4
    tX = torch.empty(0, D, device=dev())
    tS = torch.empty(0, D, D, device=dev())
    tla = torch.zeros(K, device=dev())
    tmu = torch.zeros(K, D, device=dev())
    tSg = torch.eye(D, device=dev()).repeat(K, 1, 1)
    ll, stats = ops.estep_stats(tX, tS, tla, tmu, tSg)
    torch.cuda.synchronize()
    assert ll.numel() == 0
    assert torch.count_nonzero(stats).item() == 0


@pytest.mark.parametrize("D,K", [(10, 64), (16, 128), (9, 128), (32, 256)])
def test_config_shapes(ops, D, K):
    run_case(ops, N=64 if D == 32 else 192, D=D, K=K, seed=500 + D + K)


def test_weights_and_floor(ops):
    run_case(ops, N=200, D=6, K=7, seed=7, with_g=True)
    run_case(ops, N=200, D=6, K=7, seed=8, floor=10 * float(np.finfo(np.float32).eps))
    run_case(ops, N=130, D=10, K=64, seed=9, with_g=True)


def test_diag_noise_layout(ops):
    run_case(ops, N=150, D=11, K=10, seed=11, noise="diag")


def test_shared_and_none_layouts(ops):
    rng = np.random.default_rng(12)
    X, S, alpha, mu, Sigma = synth(rng, 90, 4, 5)
    S1 = S[0]
    tX, tS1, tmu, tSg = to_dev(X, S1, mu, Sigma)
    tla = torch.tensor(np.log(alpha), dtype=torch.float32, device=dev())
    ll = ops.loglik(tX, tS1, tla, tmu, tSg, noise="shared").cpu().numpy()
    want = orc.loglik(X, np.broadcast_to(S1, S.shape), alpha, mu, Sigma)
    assert np.max(np.abs(ll - want)) < 2e-4 * np.abs(want).max()
    ll0 = ops.loglik(tX, None, tla, tmu, tSg, noise="none").cpu().numpy()
    want0 = orc.loglik(X, np.zeros_like(S), alpha, mu, Sigma)
    assert np.max(np.abs(ll0 - want0)) < 2e-4 * np.abs(want0).max()


@pytest.mark.parametrize("sigp,accp,nslot", [(1, 0, 1), (1, 0, 2), (0, 0, 2), (0, 1, 2), (1, 1, 2), (1, 1, 4),
                                             (2, 2, 1), (2, 2, 2)])
def test_placement_variants_agree(ops, sigp, accp, nslot):
    try:
        ops.set_tuning(sigp=sigp, accp=accp, nslot=nslot)
        run_case(ops, N=333, D=10, K=64, seed=21)
        run_case(ops, N=100, D=5, K=8, seed=22)
    finally:
        ops.set_tuning()


@pytest.mark.parametrize("no_bulk,pt,nst", [(1, -1, -1), (0, 4, 2), (0, 32, 4)])
def test_staging_variants_agree(ops, no_bulk, pt, nst):
    try:
        ops.set_tuning(no_bulk=no_bulk, pt=pt, nst=nst)
        run_case(ops, N=1003, D=10, K=64, seed=23)
        run_case(ops, N=1003, D=7, K=3, seed=24)
    finally:
        ops.set_tuning()


def test_unaligned_base_pointers(ops):
    rng = np.random.default_rng(31)
    X, S, alpha, mu, Sigma = synth(rng, 201, 3, 4)
    tX, tS, tmu, tSg = to_dev(X, S, mu, Sigma)
    tla = torch.tensor(np.log(alpha), dtype=torch.float32, device=dev())
    ll_a = ops.loglik(tX[1:], tS[1:], tla, tmu, tSg)      # 12-byte / 36-byte offsets: not 16-byte aligned
    want = orc.loglik(X[1:], S[1:], alpha, mu, Sigma)
    assert np.max(np.abs(ll_a.cpu().numpy() - want)) < 2e-4 * np.abs(want).max()


def test_estep_full_matches_reference_golden(ops):
    for name in ("ref_em_d3_k4_f64.npz", "ref_em_d10_k6_f64.npz", "ref_em_d16_k5_f64.npz"):
        g = golden(name)
        tX, tS, tmu, tSg = to_dev(g["X"], g["S"], g["mu"], g["Sigma"])
        tla = torch.tensor(np.log(g["alpha"]), dtype=torch.float32, device=dev())
        ll, log_r, b, B, status = ops.estep_full(tX, tS, tla, tmu, tSg)
        torch.cuda.synchronize()
        assert status.item() == 0
        assert abs(ll.sum().item() - float(g["ll_sum"])) < LL_RTOL * abs(float(g["ll_sum"]))
        assert np.max(np.abs(np.exp(log_r.cpu().numpy()) - np.exp(g["log_resps"]))) < 1e-4
        assert rel(b.cpu().numpy(), g["cond_means"]) < 1e-4
        assert rel(B.cpu().numpy(), g["cond_covars"]) < 1e-3


def test_stats_and_mstep_match_reference_golden(ops):
    for name in ("ref_em_d3_k4_f64.npz", "ref_em_d10_k6_f64.npz", "ref_em_d16_k5_f64.npz"):
        g = golden(name)
        N, D = g["X"].shape
        K = g["alpha"].size
        tX, tS, tmu, tSg = to_dev(g["X"], g["S"], g["mu"], g["Sigma"])
        talpha = torch.tensor(g["alpha"], dtype=torch.float32, device=dev())
        tla = talpha.log()
        _, stats = ops.estep_stats(tX, tS, tla, tmu, tSg)
        ops.mstep_update(stats, N, float(g["w"]), tmu, tSg, talpha, tla)
        torch.cuda.synchronize()
        assert rel(talpha.cpu().numpy(), g["m1_weights"][:, 0]) < STAT_RTOL
        assert rel(tmu.cpu().numpy(), g["m1_means"]) < STAT_RTOL
        assert rel(tSg.cpu().numpy(), g["m1_covars"]) < STAT_RTOL
        assert np.max(np.abs(tla.cpu().numpy() - np.log(g["m1_weights"][:, 0]))) < 1e-4


def test_em_iterations_track_reference_boston(ops):
    """BASELINE.json configs[0]: boston 506x11, K=10, diagonal noise, injected init."""
    g = golden("ref_boston_k10_f64.npz")
    X = g["X"]
    N, D = X.shape
    K = 10
    tX, tS = to_dev(X, g["noise_var"])
    talpha = torch.full((K,), 1.0 / K, device=dev())
    tla = talpha.log()
    tmu = tX[torch.tensor(g["rows"], device=dev())].clone()
    tSg = torch.eye(D, device=dev()).repeat(K, 1, 1).contiguous()
    for it in range(1, 11):
        _, stats = ops.estep_stats(tX, tS, tla, tmu, tSg, noise="diag", want_ll=False)
        tail = stats[-4:].cpu().numpy().astype(np.float64)
        assert abs(tail[0] + tail[1] - g["ll_trace"][it - 1]) < LL_RTOL * abs(g["ll_trace"][it - 1])
        ops.mstep_update(stats, N, float(g["w"]), tmu, tSg, talpha, tla)
        if it in (1, 5, 10):
            assert rel(tmu.cpu().numpy(), g["it%d_means" % it]) < STAT_RTOL
            assert rel(tSg.cpu().numpy(), g["it%d_covars" % it]) < STAT_RTOL
            assert rel(talpha.cpu().numpy(), g["it%d_weights" % it][:, 0]) < STAT_RTOL


def test_robust_dynamic_range(ops):
    """S = diag(1e12, 1e-2) and 1e-4 I stay finite."""
    r = golden("ref_robust_f32.npz")
    tX, tS, tmu, tSg = to_dev(r["X"], r["S"], r["mu"], r["Sigma"])
    tla = torch.tensor(np.log(r["alpha"]), dtype=torch.float32, device=dev())
    ll, log_r, b, B, status = ops.estep_full(tX, tS, tla, tmu, tSg)
    assert status.item() == 0
    assert abs(ll.sum().item() - float(r["ll_sum"])) < 1e-4 * abs(float(r["ll_sum"]))
    assert torch.isfinite(B).all()
    assert np.max(np.abs(np.exp(log_r.cpu().numpy()) - np.exp(r["log_resps"]))) < 1e-4
    _, stats = ops.estep_stats(tX, tS, tla, tmu, tSg)
    assert torch.isfinite(stats).all()


def test_non_positive_definite_is_flagged(ops):
    D, K, N = 2, 3, 40
    rng = np.random.default_rng(5)
    X, S, alpha, mu, Sigma = synth(rng, N, D, K)
    S[7] = np.diag([1.0, -50.0])           # Sigma_j + S_7 is indefinite for every j
    tX, tS, tmu, tSg = to_dev(X, S, mu, Sigma)
    tla = torch.tensor(np.log(alpha), dtype=torch.float32, device=dev())
    ll, sums = ops.loglik(tX, tS, tla, tmu, tSg, want_sum=True)
    assert sums[1].item() == K
    assert ll[7].item() == -math.inf
    assert torch.isfinite(ll[torch.arange(N, device=dev()) != 7]).all()
    _, stats = ops.estep_stats(tX, tS, tla, tmu, tSg)
    assert stats[-1].item() == K
    assert torch.isfinite(stats[:-4]).all()   # the bad point does not poison the statistics
    keep = np.arange(N) != 7
    s0, s1, S2, _ = orc.suff_stats(X[keep], S[keep], alpha, mu, Sigma)
    d0, d1, d2, _ = [t.cpu().numpy() for t in ops.unpack_stats(stats, D, K)]
    assert rel(d0, s0) < STAT_RTOL and rel(d2, S2) < STAT_RTOL


def test_argument_errors(ops):
    from density_deconvolution_b200 import DgmmError
    tX = torch.zeros(4, 33, device=dev())
    with pytest.raises(DgmmError):
        ops.loglik(tX, torch.zeros(4, 33, 33, device=dev()), torch.zeros(2, device=dev()),
                   torch.zeros(2, 33, device=dev()), torch.zeros(2, 33, 33, device=dev()))
    with pytest.raises(DgmmError):
        ops.loglik(torch.zeros(4, 2), torch.zeros(4, 2, 2), torch.zeros(2), torch.zeros(2, 2), torch.zeros(2, 2, 2))


def test_large_n_properties(ops):
    """Size-independent properties at a size the oracle cannot hold: additivity over shards,
    determinism, and agreement of the two passes."""
    from density_deconvolution_b200.synthetic import make_mixture, make_points
    D, K, N = 10, 64, 400_000
    alpha, mu, Sigma, q = make_mixture(D, K, seed=0, device=dev())
    X, S = make_points(N, alpha, mu, q, seed=1)
    la = alpha.log()
    ll, st = ops.estep_stats(X, S, la, mu, Sigma)
    ll_b, st_b = ops.estep_stats(X, S, la, mu, Sigma)
    assert torch.equal(st, st_b) and torch.equal(ll, ll_b), "run-to-run deterministic"
    h = N // 2 + 3
    _, s1 = ops.estep_stats(X[:h], S[:h], la, mu, Sigma)
    _, s2 = ops.estep_stats(X[h:], S[h:], la, mu, Sigma)
    tot = (s1 + s2).double()
    ref = st.double()
    scale = ref[:K].abs().max()
    assert (tot[:-4] - ref[:-4]).abs().max() < 1e-4 * max(scale.item(), ref[:-4].abs().max().item())
    assert abs((tot[-4] + tot[-3]) - (ref[-4] + ref[-3])) < 1e-6 * abs(ref[-4].item())
    s0 = ref[:K]
    assert abs(s0.sum().item() - N) < 1e-4 * N          # responsibilities sum to one per point
    ll_pass = ops.loglik(X, S, la, mu, Sigma)
    assert torch.equal(ops.loglik(X, S, la, mu, Sigma), ll_pass), "log-likelihood pass is deterministic"
    assert torch.allclose(ll_pass, ll, rtol=2e-6, atol=2e-6)   # different summation order of the K densities
    # oracle on a slice
    idx = slice(1000, 1256)
    want = orc.loglik(X[idx].cpu().numpy(), S[idx].cpu().numpy(), alpha.cpu().numpy(), mu.cpu().numpy(),
                      Sigma.cpu().numpy())
    assert np.max(np.abs(ll[idx].cpu().numpy() - want)) < 2e-4 * np.abs(want).max()


@pytest.mark.parametrize("N,D,K", [(10_000_000, 10, 64), (12_500_000, 16, 128)])
def test_full_size_properties(ops, N, D, K):
    """BASELINE.json's FULL single-GPU sizes (configs[1]; the per-GPU shard of configs[4]), through properties that do
    not need the oracle to hold the table: bit-reproducibility, additivity over shards, responsibilities summing to one
    per point, agreement of the two passes and of the packed layout, EM monotonicity, and the oracle on two slices."""
    from density_deconvolution_b200.synthetic import make_mixture, make_points
    free = torch.cuda.mem_get_info(dev())[0]
    if free < 2.2 * N * 4 * (D + D * D):
        pytest.skip("not enough free device memory for this size")
    alpha, mu, Sigma, q = make_mixture(D, K, seed=0, device=dev())
    X, S = make_points(N, alpha, mu, q, seed=1)
    la = alpha.log()
    stats = torch.empty(ops.stats_len(D, K), dtype=torch.float32, device=dev())
    ll, st = ops.estep_stats(X, S, la, mu, Sigma)
    st = st.clone()
    _, st_b = ops.estep_stats(X, S, la, mu, Sigma, want_ll=False, out_stats=stats)
    assert torch.equal(st, st_b), "run-to-run deterministic at full size"
    ref = st.double()
    assert ref[-1].item() == 0 and ref[-2].item() == N
    assert abs(ref[:K].sum().item() - N) < 1e-4 * N
    # additivity over three unequal shards (what the multi-GPU all-reduce relies on)
    cuts = [0, N // 3 + 1, 2 * (N // 3) + 7, N]
    tot = torch.zeros_like(ref)
    for lo, hi in zip(cuts, cuts[1:]):
        _, s = ops.estep_stats(X[lo:hi], S[lo:hi], la, mu, Sigma, want_ll=False)
        tot += s.double()
    body = slice(0, K * (1 + D + D * D))
    assert (tot[body] - ref[body]).abs().max() < 2e-5 * ref[body].abs().max()
    assert abs((tot[-4] + tot[-3]) - (ref[-4] + ref[-3])) < 1e-7 * abs(ref[-4].item())
    # the two passes, and the double sum against the float64 sum of the per-point values
    ll_pass, sums = ops.loglik(X, S, la, mu, Sigma, want_sum=True)
    assert torch.allclose(ll_pass, ll, rtol=3e-6, atol=3e-6)
    assert abs(sums[0].item() - ll.double().sum().item()) < 1e-7 * abs(sums[0].item())
    assert abs((ref[-4] + ref[-3]).item() - ll.double().sum().item()) < 1e-7 * abs(sums[0].item())
    # oracle on two slices (start and ragged end)
    for idx in (slice(0, 192), slice(N - 131, N)):
        want = orc.loglik(X[idx].cpu().numpy(), S[idx].cpu().numpy(), alpha.cpu().numpy(), mu.cpu().numpy(),
                          Sigma.cpu().numpy())
        assert np.max(np.abs(ll[idx].cpu().numpy() - want)) < 2e-4 * np.abs(want).max()
    del ll_pass
    # packed layout on the first 2e6 points
    n2 = 2_000_000
    P = ops.pack_noise(S[:n2])
    _, sp = ops.estep_stats(X[:n2], P, la, mu, Sigma, noise="packed", want_ll=False)
    _, sf = ops.estep_stats(X[:n2], S[:n2], la, mu, Sigma, want_ll=False)
    assert (sp - sf).abs().max().item() <= 2e-5 * sf.abs().max().item()
    del P
    # EM from a perturbed start: the log-likelihood never decreases (4 iterations, full table)
    g = torch.Generator(device="cpu").manual_seed(3)
    m = (mu.cpu() + 0.3 * torch.randn(K, D, generator=g)).to(dev())
    Sg = (1.2 * Sigma).contiguous()
    a, la2 = alpha.clone(), alpha.log()
    trace = []
    for _ in range(4):
        ops.estep_stats(X, S, la2, m, Sg, want_ll=False, out_stats=stats)
        trace.append((stats[-4].double() + stats[-3].double()).item())
        assert stats[-1].item() == 0
        ops.mstep_update(stats, N, 1e-3, m, Sg, a, la2)
    assert all(b >= a_ - 1e-7 * abs(a_) for a_, b in zip(trace, trace[1:])), trace
    assert abs(a.sum().item() - 1.0) < 1e-4

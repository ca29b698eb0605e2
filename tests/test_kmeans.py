"""k-means initialisation: ``dgmm_kmeans_assign_f32`` / ``dgmm_kmeans_update_f32``
against the oracle and against golden vectors recorded from the UNMODIFIED reference's
``deconv/gmm/util.py`` (``tests/golden/make_golden.py kmeans``).

``-m "not gpu"``: the oracle against the golden vectors, and the host loops of ``util.py`` on the CPU
test double.  ``-m gpu``: the kernels through the C ABI.  Labels are integer work (exact, except where
two centroids are equidistant to within float32 rounding); counts are exact; coordinate sums, centroids
and the total distance are float32 (relative 1e-5 / 1e-4)."""
import numpy as np
import pytest
import torch

from conftest import golden
from oracle import gmm_oracle as orc


@pytest.fixture(params=["oracle", pytest.param("cuda", marks=pytest.mark.gpu)])
def backend(request):
    if request.param == "oracle":
        import fake_ops
        return torch.device("cpu"), fake_ops
    from density_deconvolution_b200 import ops
    return torch.device("cuda", 0), ops


# ---- oracle pinned to the reference ------------------------------------------------------------
@pytest.mark.parametrize("tag", ["f64", "f32"])
def test_oracle_minibatch_k_means_matches_reference(tag):
    G = golden("ref_kmeans.npz")
    X = G["X_" + tag].astype(np.float64)
    K = G["mb_%s_counts_1" % tag].shape[0]
    for iters in (1, 3):
        C, n = X[:K].copy(), np.ones(K)
        for _ in range(iters):
            for lo in range(0, len(X), 60):
                _, counts, sums, _ = orc.kmeans_assign(X[lo:lo + 60], C)
                C, n = orc.kmeans_minibatch_update(counts, sums, C, n)
        tol = 1e-12 if tag == "f64" else 2e-5
        np.testing.assert_allclose(n, G["mb_%s_counts_%d" % (tag, iters)], rtol=0, atol=0)
        np.testing.assert_allclose(C, G["mb_%s_centroids_%d" % (tag, iters)], rtol=tol, atol=tol)


@pytest.mark.parametrize("tag", ["f64", "f32"])
def test_oracle_lloyd_k_means_matches_reference(tag):
    G = golden("ref_kmeans.npz")
    X = G["X_" + tag].astype(np.float64)
    for iters in (1, 4):
        C = G["ll_%s_init" % tag].astype(np.float64)
        for _ in range(iters):
            labels, counts, sums, _ = orc.kmeans_assign(X, C)
            C = orc.kmeans_lloyd_update(counts, sums, C)
        tol = 1e-12 if tag == "f64" else 2e-5
        np.testing.assert_allclose(C, G["ll_%s_centroids_%d" % (tag, iters)], rtol=tol, atol=tol)
        np.testing.assert_array_equal(labels, G["ll_%s_resp_%d" % (tag, iters)].argmax(axis=1))


# ---- util.py host loops (CPU double and GPU) against the reference -----------------------------
@pytest.mark.parametrize("tag", ["f64", "f32"])
def test_minibatch_k_means_matches_reference(backend, tag):
    from density_deconvolution_b200.util import minibatch_k_means
    dev, ops = backend
    G = golden("ref_kmeans.npz")
    X = torch.tensor(G["X_" + tag], dtype=torch.float32, device=dev)
    K = G["mb_%s_counts_1" % tag].shape[0]
    batches = [(X[lo:lo + 60],) for lo in range(0, X.shape[0], 60)]
    for iters in (1, 3):
        counts, cents = minibatch_k_means(batches, K, max_iters=iters, tol=0.0, device=dev, ops=ops)
        np.testing.assert_array_equal(counts.cpu().numpy(), G["mb_%s_counts_%d" % (tag, iters)])
        np.testing.assert_allclose(cents.cpu().numpy(), G["mb_%s_centroids_%d" % (tag, iters)], rtol=1e-4, atol=1e-4)


@pytest.mark.parametrize("tag", ["f64", "f32"])
def test_lloyd_passes_match_reference(backend, tag):
    dev, ops = backend
    G = golden("ref_kmeans.npz")
    X = torch.tensor(G["X_" + tag], dtype=torch.float32, device=dev)
    for iters in (1, 4):
        C = torch.tensor(G["ll_%s_init" % tag], dtype=torch.float32, device=dev).contiguous()
        for _ in range(iters):
            labels, stats = ops.kmeans_assign(X, C, want_labels=True)
            ops.kmeans_update("lloyd", stats, C)
        np.testing.assert_allclose(C.cpu().numpy(), G["ll_%s_centroids_%d" % (tag, iters)], rtol=1e-4, atol=1e-4)
        np.testing.assert_array_equal(labels.cpu().numpy(), G["ll_%s_resp_%d" % (tag, iters)].argmax(axis=1))


def test_k_means_driver_returns_one_hot(backend):
    from density_deconvolution_b200.util import k_means
    dev, ops = backend
    g = torch.Generator().manual_seed(3)
    centres = torch.tensor([[0., 0.], [6., 6.], [-6., 5.]])
    X = (centres[torch.randint(0, 3, (300,), generator=g)] + 0.3 * torch.randn(300, 2, generator=g)).to(dev)
    gen = torch.Generator(device=dev).manual_seed(11)
    resp, cents = k_means(X, 3, max_iters=30, tol=1e-9, device=dev, generator=gen, ops=ops)
    assert resp.shape == (300, 3) and torch.all(resp.sum(dim=1) == 1)
    assert torch.isfinite(cents).all()           # an empty cluster keeps its centroid instead of NaN


# ---- kernels against the oracle ---------------------------------------------------------------
def _check_assign(N, D, K, seed, spread=3.0):
    from density_deconvolution_b200 import ops
    dev = torch.device("cuda", 0)
    rng = np.random.default_rng(seed)
    C = (rng.normal(size=(K, D)) * spread).astype(np.float32)
    X = (C[rng.integers(0, K, N)] + rng.normal(size=(N, D))).astype(np.float32) if N else np.zeros((0, D), np.float32)
    tX, tC = torch.tensor(X, device=dev), torch.tensor(C, device=dev)
    labels, stats = ops.kmeans_assign(tX, tC, want_labels=True)
    labels2, stats2 = ops.kmeans_assign(tX, tC, want_labels=True)
    assert torch.equal(stats, stats2) and torch.equal(labels, labels2)         # bit-reproducible
    lab_ref, counts, sums, total = orc.kmeans_assign(X, C)
    lab = labels.cpu().numpy().astype(np.int64)
    d2 = ((X[:, None, :].astype(np.float64) - C[None].astype(np.float64)) ** 2).sum(axis=2)
    diff = np.nonzero(lab != lab_ref)[0]
    # a differing label must be a float32-level tie
    for i in diff:
        assert abs(d2[i, lab[i]] - d2[i, lab_ref[i]]) <= 1e-5 * d2[i, lab_ref[i]], (i, d2[i, lab[i]], d2[i, lab_ref[i]])
    s = stats.cpu().numpy().astype(np.float64)
    assert s.shape == (ops.kmeans_stats_len(D, K),)
    got_counts, got_sums = s[:K], s[K:K + K * D].reshape(K, D)
    assert got_counts.sum() == N
    if len(diff) == 0:
        np.testing.assert_array_equal(got_counts, counts)
        np.testing.assert_allclose(got_sums, sums, rtol=1e-5, atol=1e-5 * max(1.0, np.abs(sums).max()))
    np.testing.assert_array_equal(got_counts, np.bincount(lab, minlength=K))   # counts agree with the labels
    assert abs((s[-2] + s[-1]) - total) <= 1e-5 * max(total, 1.0)
    # update kernels
    for mode in ("lloyd", "minibatch"):
        c_run = torch.full((K,), 2.0, device=dev)
        Cn = tC.clone()
        ops.kmeans_update(mode, stats, Cn, c_run if mode == "minibatch" else None)
        if mode == "lloyd":
            want = orc.kmeans_lloyd_update(got_counts, got_sums, C)
        else:
            want, n_new = orc.kmeans_minibatch_update(got_counts, got_sums, C, np.full(K, 2.0))
            np.testing.assert_array_equal(c_run.cpu().numpy(), n_new)
        np.testing.assert_allclose(Cn.cpu().numpy(), want, rtol=1e-5, atol=1e-5)


@pytest.mark.gpu
@pytest.mark.parametrize("D", [1, 2, 3, 4, 5, 7, 8, 9, 10, 11, 12, 15, 16, 17, 24, 31, 32])
def test_assign_all_row_strides(D):
    _check_assign(3000, D, 13, seed=100 + D)


@pytest.mark.gpu
@pytest.mark.parametrize("N", [0, 1, 255, 256, 257, 511, 513, 100003])
def test_assign_ragged_sizes(N):
    _check_assign(N, 10, 64, seed=N)


@pytest.mark.gpu
@pytest.mark.parametrize("D,K", [(10, 1), (10, 64), (9, 128), (16, 128), (32, 256), (3, 256), (2, 5)])
def test_assign_shapes_of_the_configs(D, K):
    _check_assign(20000, D, K, seed=D * 1000 + K)


@pytest.mark.gpu
def test_assign_full_size_properties():
    """c2-sized pass (N=1e7, D=10, K=64): counts sum to N, sums add up to the column sums of X,
    labels of a permuted copy are the permuted labels."""
    from density_deconvolution_b200 import ops
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev).manual_seed(5)
    N, D, K = 10_000_000, 10, 64
    C = 5.0 * torch.randn(K, D, device=dev, generator=g)
    X = C[torch.randint(0, K, (N,), device=dev, generator=g)] + torch.randn(N, D, device=dev, generator=g)
    labels, stats = ops.kmeans_assign(X, C, want_labels=True)
    counts, sums = stats[:K], stats[K:K + K * D].view(K, D)
    assert counts.sum().item() == N
    col = X.double().sum(dim=0)
    assert torch.allclose(sums.double().sum(dim=0), col, rtol=1e-5, atol=1e-2)
    perm = torch.randperm(N, device=dev, generator=g)[:1_000_000]
    lab_p, _ = ops.kmeans_assign(X[perm].contiguous(), C, want_labels=True)
    assert torch.equal(lab_p, labels[perm])
    assert torch.equal(torch.bincount(labels.long(), minlength=K).float(), counts)


@pytest.mark.gpu
def test_assign_rejects_cpu_tensors():
    from density_deconvolution_b200 import ops
    with pytest.raises(Exception):
        ops.kmeans_assign(torch.zeros(4, 2), torch.zeros(2, 2))


# ---- device-resident minibatch gather (dgmm_gather_points_f32) --------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("D,layout", [(1, "full"), (3, "full"), (9, "full"), (10, "full"), (16, "full"), (32, "full"),
                                      (10, "diag"), (7, "diag"), (10, "none")])
def test_gather_points_is_exact(D, layout):
    from density_deconvolution_b200 import ops
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev).manual_seed(D)
    N, B = 5000, 1237
    X = torch.randn(N, D, device=dev, generator=g)
    S = {"full": torch.randn(N, D, D, device=dev, generator=g), "diag": torch.rand(N, D, device=dev, generator=g),
         "none": None}[layout]
    idx = torch.randperm(N, device=dev, generator=g)[:B].contiguous()
    bx, bs = ops.gather_points(X, S, idx)
    assert torch.equal(bx, X.index_select(0, idx))
    if S is None:
        assert bs is None
    else:
        assert torch.equal(bs, S.index_select(0, idx))
    bx0, _ = ops.gather_points(X, S, idx[:0].contiguous())
    assert bx0.shape == (0, D)


@pytest.mark.gpu
def test_device_batches_cover_the_table_once_per_epoch():
    from density_deconvolution_b200 import ops
    from density_deconvolution_b200.data import DeviceBatches
    dev = torch.device("cuda", 0)
    N, D = 1000, 4
    X = torch.arange(N, device=dev, dtype=torch.float32)[:, None].repeat(1, D)
    S = X[:, :, None] * torch.eye(D, device=dev)
    loader = DeviceBatches((X, S), 96, dev, shuffle=True, ops=ops)
    seen = torch.cat([xb[:, 0] for xb, sb in loader])
    assert len(loader) == 11 and torch.equal(seen.sort()[0], X[:, 0])
    for xb, sb in loader:
        assert torch.equal(sb[:, 0, 0], xb[:, 0])          # the noise rows travel with their points

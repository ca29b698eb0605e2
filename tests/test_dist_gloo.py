"""World-size-2 ``gloo`` tests of the sharded path's HOST logic (CPU, test double for the
kernels): every rank holds half of the points, the packed statistics are all-reduced once per
step, and the fitted parameters must equal the single-process fit on the whole table."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _problem():
    sys.path.insert(0, HERE)
    from helpers import synth
    rng = np.random.default_rng(77)
    X, S, alpha, mu, Sigma = synth(rng, 64, 3, 4)
    f = lambda a: torch.tensor(a, dtype=torch.float32)  # noqa: E731
    return f(X), f(S), f(alpha), f(mu), f(Sigma)


def _fit_all(X, S, alpha, mu, Sigma, batch):
    """Runs every estimator on (X, S); returns a dict of result tensors."""
    from density_deconvolution_b200 import estimators as est
    import fake_ops
    est._ops = fake_ops
    dev = torch.device("cpu")
    out = {}
    m = est.DeconvGMM(4, 3, epochs=3, tol=0.0, w=1e-3, device=dev).set_params(alpha, mu, Sigma)
    m.fit((X, S), init=False)
    out["em_means"], out["em_covars"], out["em_weights"] = m.means, m.covars, m.weights
    out["em_ll"] = torch.tensor(m.log_prob_curve)
    out["em_score"] = m.score((X, S)).reshape(1)
    b = est.BatchDeconvGMM(4, 3, epochs=2, w=1e-3, batch_size=batch, device=dev).set_params(alpha, mu, Sigma)
    b.fit((X, S), init=False)
    out["bt_means"], out["bt_covars"] = b.means, b.covars
    o = est.OnlineDeconvGMM(4, 3, epochs=2, w=1e-3, batch_size=batch, step_size=0.2, device=dev)
    o.set_params(alpha, mu, Sigma)
    o.fit((X, S), init=False)
    out["on_means"], out["on_covars"], out["on_weights"] = o.means, o.covars, o.weights
    out["on_ll"] = torch.tensor(o.train_ll_curve)
    s = est.SGDDeconvGMM(4, 3, epochs=3, lr=1e-2, batch_size=batch, w=1e-3, device=dev)
    with torch.no_grad():
        s.module.soft_weights.copy_(alpha.log())
        s.module.means.copy_(mu)
    s.fit((X, S), init=False)
    out["sgd_means"], out["sgd_l_lower"] = s.module.means.detach(), s.module.l_lower.detach()
    out["sgd_loss"] = torch.tensor(s.train_loss_curve)
    out["sgd_score"] = torch.tensor([s.score_batch((X, S))])
    # k-means initialisation must agree across ranks
    torch.manual_seed(5)
    k = est.DeconvGMM(4, 3, epochs=1, tol=0.0, w=1e-3, device=dev)
    k.fit((X, S))
    out["km_means"] = k.means
    return out


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, HERE)
    torch.set_num_threads(1)
    import torch.distributed as dist
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    from density_deconvolution_b200.dist import shard_bounds
    X, S, alpha, mu, Sigma = _problem()
    lo, hi = shard_bounds(X.shape[0])
    res = _fit_all(X[lo:hi], S[lo:hi], alpha, mu, Sigma, batch=(hi - lo))
    q.put((rank, {k: v.clone().numpy() for k, v in res.items()}))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_two_rank_fit_equals_single_process():
    sys.path.insert(0, ROOT)
    sys.path.insert(0, HERE)
    X, S, alpha, mu, Sigma = _problem()
    single = {k: v.clone().numpy() for k, v in _fit_all(X, S, alpha, mu, Sigma, batch=X.shape[0]).items()}
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=240) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for key, want in single.items():
        if key == "km_means":
            continue   # different shards draw the same centroids, but Lloyd's sums differ in order only
        for r in range(2):
            np.testing.assert_allclose(got[r][key], want, rtol=2e-4, atol=2e-5, err_msg="%s rank %d" % (key, r))
    for key in single:
        np.testing.assert_array_equal(got[0][key], got[1][key], err_msg="ranks disagree on " + key)
    np.testing.assert_allclose(got[0]["km_means"], single["km_means"], rtol=1e-3, atol=1e-4)


def test_shard_bounds_cover_everything():
    from density_deconvolution_b200.dist import shard_bounds
    for n in (0, 1, 7, 64, 1000003):
        for ws in (1, 2, 3, 8):
            edges = [shard_bounds(n, r, ws) for r in range(ws)]
            assert edges[0][0] == 0 and edges[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(edges, edges[1:]))
            sizes = [hi - lo for lo, hi in edges]
            assert max(sizes) - min(sizes) <= 1


# ---- unequal shards: the ranks' own batch counts differ (one rank blocked forever in all_reduce) ----------
def _worker_unequal(rank, world, port, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, HERE)
    torch.set_num_threads(1)
    import torch.distributed as dist
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    from density_deconvolution_b200 import estimators as est
    from density_deconvolution_b200 import util
    from density_deconvolution_b200.data import DeviceBatches
    from density_deconvolution_b200.dist import shard_bounds
    import fake_ops
    est._ops = fake_ops
    util._real_ops = fake_ops
    from helpers import synth
    rng = np.random.default_rng(78)
    X, S, alpha, mu, Sigma = synth(rng, 51, 3, 4)          # 26 + 25 points: 3 vs 2 full batches of 10 (+ a tail each)
    f = lambda a: torch.tensor(a, dtype=torch.float32)  # noqa: E731
    X, S, alpha, mu, Sigma = f(X), f(S), f(alpha), f(mu), f(Sigma)
    lo, hi = shard_bounds(X.shape[0])
    dev = torch.device("cpu")
    Xs, Ss = X[lo:hi], S[lo:hi]
    loader = DeviceBatches((Xs[:21 + 5 * (1 - rank)], Ss[:21 + 5 * (1 - rank)]), 10, dev).sync_ranks()   # 26 vs 21 points: 3 vs 3; then
    n_batches = len(loader)
    loader2 = DeviceBatches((Xs[:25 - 10 * rank], Ss[:25 - 10 * rank]), 10, dev, drop_last=True).sync_ranks()   # 25 vs 15: 2 vs 1
    sizes = est._global_batch_sizes(DeviceBatches((Xs, Ss), 10, dev).sync_ranks())
    s = est.SGDDeconvGMM(4, 3, epochs=2, lr=1e-2, batch_size=10, w=1e-3, k_means_iters=2, device=dev)
    s.fit((Xs, Ss))                                         # would dead-lock without the rank-synchronised batch count
    o = est.OnlineDeconvGMM(4, 3, epochs=2, w=1e-3, batch_size=12, step_size=0.2, k_means_iters=2, device=dev)
    o.fit((Xs, Ss))                                         # drop_last: 2 vs 2 here; init loader 1200-point batches: 1 vs 1
    q.put((rank, dict(n_batches=n_batches, n_drop=len(loader2), sizes=sizes, n_local=hi - lo,
                      loss=np.asarray(s.train_loss_curve), means=s.means.numpy(), on_ll=np.asarray(o.train_ll_curve))))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_unequal_shards_do_not_deadlock_and_scale_the_loss_by_the_global_batch():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker_unequal, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=240) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert got[0]["n_local"] == 26 and got[1]["n_local"] == 25
    assert got[0]["n_batches"] == got[1]["n_batches"] == 3
    assert got[0]["n_drop"] == got[1]["n_drop"] == 1            # minimum over ranks of (2, 1)
    # 26 + 25 points in batches of 10 per rank: global batches 20, 20, 6 + 5
    assert got[0]["sizes"] == got[1]["sizes"] == [20, 20, 11]
    np.testing.assert_array_equal(got[0]["loss"], got[1]["loss"])
    np.testing.assert_array_equal(got[0]["means"], got[1]["means"])
    np.testing.assert_array_equal(got[0]["on_ll"], got[1]["on_ll"])
    assert np.all(np.isfinite(got[0]["loss"]))

"""On-hardware multi-rank parity (needs >= 2 GPUs; skipped otherwise): the NCCL all-reduced statistics of two
shards equal the single-GPU statistics of the concatenated table, and a sharded ``DeconvGMM.fit`` /
``SGDDeconvGMM.fit`` ends with the parameters of the single-process fit (one all-reduce of the
packed buffer per step, identical update on every rank).  Run with ``gpurun --gpus 2`` (log under profiles/)."""
import os
import socket

import numpy as np
import pytest
import torch

from helpers import synth

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, payload, out_path):
    import torch.distributed as dist
    from density_deconvolution_b200 import DeconvGMM, SGDDeconvGMM, ops
    from density_deconvolution_b200.dist import shard_bounds, allreduce_stats
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    X, S, alpha, mu, Sigma = [torch.tensor(a, dtype=torch.float32, device=dev) for a in payload]
    N = X.shape[0]
    lo, hi = shard_bounds(N, rank, world)
    la = alpha.log()
    res = {}
    _, st = ops.estep_stats(X[lo:hi].contiguous(), S[lo:hi].contiguous(), la, mu, Sigma, want_ll=False)
    allreduce_stats(st)
    res["stats_sharded"] = st.cpu().numpy()
    K, D = mu.shape
    m = DeconvGMM(K, D, epochs=3, w=1e-3, tol=0.0, device=dev)
    m.set_params(alpha, mu, Sigma)
    m.fit((X[lo:hi].contiguous(), S[lo:hi].contiguous()), init=False)
    res["em_means"], res["em_covars"], res["em_weights"] = (t.cpu().numpy() for t in (m.means, m.covars, m.weights))
    res["em_ll"] = np.asarray(m.log_prob_curve)
    for tag, peer in (("sgd", True), ("sgd_nccl", False)):
        # peer = True: the ranks' statistics are summed inside the Adam kernels over NVLink peer memory
        # (dgmm_sgd_adam_step_peers_f32); False: the NCCL all-reduce between the two phases of the step
        sg = SGDDeconvGMM(K, D, epochs=2, batch_size=64, lr=1e-2, w=1e-3, device=dev)
        sg.shuffle = False
        sg.peer_reduce = peer
        with torch.no_grad():
            sg.module.soft_weights.copy_(la)
            sg.module.means.copy_(mu)
        sg.fit((X[lo:hi].contiguous(), S[lo:hi].contiguous()), init=False)
        res[tag + "_means"] = sg.means.cpu().numpy()
        res[tag + "_covars"] = sg.covars.cpu().numpy()
        res[tag + "_curve"] = np.asarray(sg.train_loss_curve)
        res[tag + "_used_peers"] = np.asarray(getattr(sg, "_peers", None) is not None)
    if rank == 0:
        np.savez(out_path, **res)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_two_rank_nccl_statistics_and_fits_equal_single_gpu(tmp_path):
    import torch.multiprocessing as mp
    from density_deconvolution_b200 import DeconvGMM, SGDDeconvGMM, ops
    rng = np.random.default_rng(31337)
    N, D, K = 4097, 6, 40          # odd N: the two shards differ by one point
    X, S, alpha, mu, Sigma = synth(rng, N, D, K)
    out = str(tmp_path / "rank0.npz")
    mp.spawn(_worker, args=(2, _free_port(), (X, S, alpha, mu, Sigma), out), nprocs=2, join=True)
    z = np.load(out)
    dev = torch.device("cuda", 0)
    tX, tS, ta, tmu, tSg = [torch.tensor(a, dtype=torch.float32, device=dev) for a in (X, S, alpha, mu, Sigma)]
    _, st = ops.estep_stats(tX, tS, ta.log(), tmu, tSg, want_ll=False)
    one = st.cpu().numpy().astype(np.float64)
    two = z["stats_sharded"].astype(np.float64)
    body = slice(0, K * (1 + D + D * D))
    assert np.max(np.abs(two[body] - one[body])) <= 1e-6 * np.max(np.abs(one[body]))
    assert abs((two[-4] + two[-3]) - (one[-4] + one[-3])) <= 1e-6 * abs(one[-4] + one[-3])
    assert two[-2] == one[-2] == N and two[-1] == 0
    m = DeconvGMM(K, D, epochs=3, w=1e-3, tol=0.0, device=dev)
    m.set_params(alpha, mu, Sigma)
    m.fit((tX, tS), init=False)
    for name, t in (("em_means", m.means), ("em_covars", m.covars), ("em_weights", m.weights)):
        ref = t.cpu().numpy()
        assert np.max(np.abs(z[name] - ref)) <= 1e-5 * np.max(np.abs(ref)), name
    np.testing.assert_allclose(z["em_ll"], m.log_prob_curve, rtol=1e-6)
    # SGD: global minibatch = 2 x 64 taken shard-wise, so compare against one process on the interleaved order
    assert np.all(np.isfinite(z["sgd_curve"])) and z["sgd_curve"][-1] > z["sgd_curve"][0] - 1.0
    lo = (N + 1) // 2
    order = []
    nb = min(-(-lo // 64), -(-(N - lo) // 64))
    for b in range(nb):
        order += list(range(b * 64, min((b + 1) * 64, lo))) + list(range(lo + b * 64, min(lo + (b + 1) * 64, N)))
    order = torch.tensor(order, device=dev)
    sg = SGDDeconvGMM(K, D, epochs=2, batch_size=128, lr=1e-2, w=1e-3, device=dev)
    sg.shuffle = False
    with torch.no_grad():
        sg.module.soft_weights.copy_(ta.log())
        sg.module.means.copy_(tmu)
    # the last global batch of the sharded run has 1 + 0 or fewer points per rank; compare the means loosely
    sg.fit((tX[order].contiguous(), tS[order].contiguous()), init=False)
    ref = sg.means.cpu().numpy()
    assert np.max(np.abs(z["sgd_means"] - ref)) <= 2e-3 * np.max(np.abs(ref))
    # fused peer-memory reduction == NCCL all-reduce (two ranks: a + b in either order is the same float)
    assert bool(z["sgd_used_peers"]) and not bool(z["sgd_nccl_used_peers"])
    np.testing.assert_allclose(z["sgd_curve"], z["sgd_nccl_curve"], rtol=1e-7)
    assert np.max(np.abs(z["sgd_means"] - z["sgd_nccl_means"])) <= 1e-6 * np.max(np.abs(ref))
    assert np.max(np.abs(z["sgd_covars"] - z["sgd_nccl_covars"])) <= 1e-6 * np.max(np.abs(z["sgd_nccl_covars"]))

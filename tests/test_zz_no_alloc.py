"""The allocation-free entry point ``dgmm_kmeans_assign_static_f32`` (scratch in a file-scope ``static __device__`` array
of the library; include/dgmm_b200.h) against the existing ``dgmm_kmeans_assign_f32`` (caller-provided workspace), through
the C ABI, on two calls in a row that share the array.  Collected last on purpose:
a failure here hides nothing after it.  Passes alone in a fresh process and after the other tests.

Tolerance: the two entry points launch the same kernels on the same grid, so the packed statistics are expected to be
bit-identical; the assertion allows the 1e-6 relative the other k-means tests use."""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else ctypes.c_void_p(0)


def _problem(seed, N, D, K, dev):
    g = torch.Generator(device="cpu").manual_seed(seed)
    cents = 4.0 * torch.randn(K, D, generator=g)
    lab = torch.randint(0, K, (N,), generator=g)
    X = cents[lab] + torch.randn(N, D, generator=g)
    return X.to(dev).contiguous(), (cents + 0.3 * torch.randn(K, D, generator=g)).to(dev).contiguous()


def _both(L, X, cents, dev):
    N, D = X.shape
    K = cents.shape[0]
    n_stats = K * (1 + D) + 2
    stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    # existing entry point: caller-provided workspace
    ws_bytes = L.dgmm_kmeans_workspace_bytes(D, K)
    assert ws_bytes > 0
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
    lab_a = torch.empty(N, dtype=torch.int32, device=dev)
    st_a = torch.empty(n_stats, dtype=torch.float32, device=dev)
    rc = L.dgmm_kmeans_assign_f32(_ptr(X), N, D, K, _ptr(cents), _ptr(lab_a), _ptr(st_a), _ptr(ws), ws_bytes, stream)
    assert rc == 0, rc
    # new entry point: no workspace argument; outputs allocated BEFORE the call, device memory must not move during it
    lab_b = torch.full((N,), -1, dtype=torch.int32, device=dev)
    st_b = torch.full((n_stats,), float("nan"), dtype=torch.float32, device=dev)
    torch.cuda.synchronize(dev)
    free_before = torch.cuda.mem_get_info(dev)[0]
    rc = L.dgmm_kmeans_assign_static_f32(_ptr(X), N, D, K, _ptr(cents), _ptr(lab_b), _ptr(st_b), stream)
    assert rc == 0, rc
    torch.cuda.synchronize(dev)
    assert torch.cuda.mem_get_info(dev)[0] == free_before, "the allocation-free entry point changed the free device memory"
    return (lab_a, st_a), (lab_b, st_b)


def test_static_scratch_entry_point_matches_workspace_entry_point_twice_in_a_row():
    from density_deconvolution_b200 import _lib
    L = _lib.lib()
    dev = torch.device("cuda", 0)
    with torch.cuda.device(dev):
        assert L.dgmm_kmeans_static_scratch_bytes() == 8 << 20
        # two calls in a row on the shared array, the second with a SMALLER shape (stale partials of the first call
        # stay in the array beyond what the second one writes) and then the first shape again
        for seed, N, D, K in ((1, 100_003, 10, 64), (2, 777, 3, 5), (3, 50_000, 10, 64), (4, 4099, 16, 80)):
            X, cents = _problem(seed, N, D, K, dev)
            (lab_a, st_a), (lab_b, st_b) = _both(L, X, cents, dev)
            assert torch.equal(lab_a, lab_b), (N, D, K)
            a, b = st_a.double().cpu().numpy(), st_b.double().cpu().numpy()
            assert np.all(np.isfinite(b))
            assert np.max(np.abs(a - b)) <= 1e-6 * np.max(np.abs(a)), (N, D, K)
            assert abs(b[:K].sum() - N) < 0.5               # counts: every point assigned once
        # the existing test-suite oracle on the last problem: labels are the nearest centroids
        d2 = ((X[:, None, :] - cents[None, :, :]) ** 2).sum(dim=2)
        near = d2.min(dim=1).values
        picked = d2.gather(1, lab_b.long()[:, None])[:, 0]
        assert torch.all(picked <= near * (1 + 1e-5) + 1e-6)


def test_static_scratch_entry_point_rejects_what_does_not_fit():
    from density_deconvolution_b200 import _lib
    L = _lib.lib()
    dev = torch.device("cuda", 0)
    with torch.cuda.device(dev):
        D, K, N = 32, 256, 64                                # workspace ~43 MB on a 148-SM part: beyond the 8 MiB array
        assert L.dgmm_kmeans_workspace_bytes(D, K) > L.dgmm_kmeans_static_scratch_bytes()
        X, cents = _problem(9, N, D, K, dev)
        lab = torch.empty(N, dtype=torch.int32, device=dev)
        st = torch.empty(K * (1 + D) + 2, dtype=torch.float32, device=dev)
        rc = L.dgmm_kmeans_assign_static_f32(_ptr(X), N, D, K, _ptr(cents), _ptr(lab), _ptr(st), None)
        assert rc == -8                                      # DGMM_ERR_STATIC_SCRATCH
        with pytest.raises(_lib.DgmmError, match="static array"):
            _lib.check(rc, "dgmm_kmeans_assign_static_f32")
        assert L.dgmm_kmeans_assign_static_f32(_ptr(X), N, 0, K, _ptr(cents), _ptr(lab), _ptr(st), None) == -2
        assert L.dgmm_kmeans_assign_static_f32(_ptr(X), N, D, 0, _ptr(cents), _ptr(lab), _ptr(st), None) == -3

"""Host-side input handling (reference: deconv/gmm/data.py:7-17, sgd_deconv_gmm.py:31-41, and the DataLoader semantics of
sgd_gmm.py:139-145): ``as_tensors`` on every accepted container, ``DeviceBatches`` on every noise layout, and the
estimators' layout detection.  CPU only (the kernels are replaced by the oracle-backed test double)."""
import numpy as np
import pytest
import torch

from density_deconvolution_b200.data import DeconvDataset, DeviceBatches, SGDDeconvDataset, as_tensors
from density_deconvolution_b200 import estimators


def _table(N=23, D=3, seed=0):
    g = torch.Generator().manual_seed(seed)
    X = torch.randn(N, D, generator=g)
    Q = 0.3 * torch.randn(N, D, D, generator=g)
    S = Q.transpose(1, 2) @ Q + 0.01 * torch.eye(D)
    return X, S


def test_as_tensors_accepts_every_container():
    X, S = _table()
    for data in ((X, S), [X, S], DeconvDataset(X, S), SGDDeconvDataset(X, S), torch.utils.data.TensorDataset(X, S)):
        a, b = as_tensors(data)
        assert torch.equal(a, X) and torch.equal(b, S)
    for data in ((X,), [X], X, torch.utils.data.TensorDataset(X)):
        a, b = as_tensors(data)
        assert torch.equal(a, X) and b is None

    class PerSample(torch.utils.data.Dataset):            # a map-style dataset without .X / .tensors
        def __len__(self):
            return X.shape[0]

        def __getitem__(self, i):
            return X[i], S[i]
    a, b = as_tensors(PerSample())
    assert torch.equal(a, X) and torch.equal(b, S)
    with pytest.raises(ValueError):
        DeconvDataset(X, S[:-1])


def test_layout_detection():
    X, S = _table(N=23, D=3)
    ia, ib = np.tril_indices(3)
    assert estimators._noise_layout(X, S) == "full"
    assert estimators._noise_layout(X, S[:, ia, ib]) == "packed"
    assert estimators._noise_layout(X, S.diagonal(dim1=1, dim2=2)) == "diag"
    assert estimators._noise_layout(X, S[0]) == "shared"
    assert estimators._noise_layout(X, None) == "none"
    with pytest.raises(ValueError):
        estimators._noise_layout(X, torch.zeros(23, 5))
    # D = 2: packed and diagonal differ in width (3 vs 2); D = 1: all layouts coincide in size, full wins by rank
    X2, S2 = _table(N=9, D=2)
    assert estimators._noise_layout(X2, torch.zeros(9, 3)) == "packed"
    assert estimators._noise_layout(X2, torch.zeros(9, 2)) == "diag"


@pytest.mark.parametrize("layout", ["full", "packed", "diag", "shared", "none"])
@pytest.mark.parametrize("shuffle", [False, True])
def test_device_batches_every_layout(layout, shuffle):
    X, S = _table(N=23, D=3)
    ia, ib = np.tril_indices(3)
    Sx = {"full": S, "packed": S[:, ia, ib].contiguous(), "diag": S.diagonal(dim1=1, dim2=2).contiguous(),
          "shared": S[0].contiguous(), "none": None}[layout]
    g = torch.Generator().manual_seed(5)
    loader = DeviceBatches((X, Sx), 10, torch.device("cpu"), shuffle=shuffle, generator=g)
    assert loader.layout == layout and len(loader) == 3
    seen = []
    for xb, sb in loader:
        assert xb.shape[1] == 3 and xb.shape[0] in (10, 3)
        if layout == "none":
            assert sb is None
        elif layout == "shared":
            assert torch.equal(sb, Sx)
        else:
            assert sb.shape[0] == xb.shape[0] and tuple(sb.shape[1:]) == tuple(Sx.shape[1:])
            # the noise rows travel with their points
            for r in range(xb.shape[0]):
                i = int((X == xb[r]).all(dim=1).nonzero()[0, 0])
                assert torch.equal(sb[r], Sx[i])
        seen.append(xb)
    allx = torch.cat(seen)
    assert allx.shape[0] == 23
    assert torch.equal(torch.sort(allx[:, 0]).values, torch.sort(X[:, 0]).values)   # every point exactly once
    dl = DeviceBatches((X, Sx), 10, torch.device("cpu"), shuffle=shuffle, drop_last=True)
    assert len(dl) == 2 and sum(xb.shape[0] for xb, _ in dl) == 20
    assert dl.sync_ranks() is dl and len(dl) == 2                                   # single process: a no-op
    assert len(DeviceBatches((X, Sx), None, torch.device("cpu"))) == 1              # batch_size None: one batch


def test_global_batch_sizes_single_process():
    X, S = _table(N=23, D=3)
    loader = DeviceBatches((X, S), 10, torch.device("cpu"))
    assert estimators._global_batch_sizes(loader) == [10, 10, 3]
    assert estimators._global_batch_sizes(DeviceBatches((X, S), 10, torch.device("cpu"), drop_last=True)) == [10, 10]

"""Input types of the hot path (reference: deconv/gmm/data.py:7-17,
deconv/gmm/sgd_deconv_gmm.py:31-41) and a device-resident batch iterator that replaces the
per-sample ``DataLoader`` indexing of the reference's fit loops (deconv/gmm/sgd_gmm.py:118-151,
deconv/gmm/online_deconv_gmm.py:43-56)."""
import torch
import torch.utils.data as data_utils


class DeconvDataset(data_utils.Dataset):
    """Map-style dataset of ``(X[i], noise_covars[i])``; ``.X [N,D]`` and
    ``.noise_covars [N,D,D]`` hold the full tensors (the layout contract of the kernels)."""

    def __init__(self, X, noise_covars):
        if X.shape[0] != noise_covars.shape[0]:
            raise ValueError("X and noise_covars disagree on N")
        self.X = X
        self.noise_covars = noise_covars

    def __len__(self):
        return self.X.shape[0]

    def __getitem__(self, i):
        return (self.X[i, :], self.noise_covars[i, :, :])


SGDDeconvDataset = DeconvDataset


def as_tensors(data):
    """``(X, S)`` from a dataset exposing ``.X/.noise_covars``, a tuple/list, or any map-style
    dataset of ``(x_i, S_i)`` pairs (stacked once)."""
    if hasattr(data, "X") and hasattr(data, "noise_covars"):
        return data.X, data.noise_covars
    if isinstance(data, (tuple, list)) and len(data) == 2 and torch.is_tensor(data[0]):
        return data[0], data[1]
    xs, ss = zip(*[data[i] for i in range(len(data))])
    return torch.stack([torch.as_tensor(x) for x in xs]), torch.stack([torch.as_tensor(s) for s in ss])


class DeviceBatches:
    """Iterates ``(X_b, S_b)`` minibatches that already live on ``device``.

    The whole table is uploaded once (float32); every epoch draws a permutation on the device
    and gathers batches there (one ``dgmm_gather_points_f32`` launch per batch when ``ops`` is given).  Semantics of ``DataLoader(batch_size, shuffle, drop_last)``."""

    def __init__(self, data, batch_size, device, shuffle=False, drop_last=False, generator=None, ops=None):
        X, S = as_tensors(data)
        self.X = X.to(device=device, dtype=torch.float32).contiguous()
        self.S = S.to(device=device, dtype=torch.float32).contiguous()
        self.n = self.X.shape[0]
        self.batch_size = int(batch_size) if batch_size else self.n
        self.shuffle = shuffle
        self.drop_last = drop_last
        self.device = device
        self.generator = generator
        # ops.gather_points: X and S rows of a shuffled batch in ONE launch (dgmm_gather_points_f32)
        per_point = self.S.dim() in (2, 3) and self.S.shape[0] == self.n and self.S.shape[-1] == self.X.shape[1]
        self.ops = ops if (ops is not None and per_point) else None

    def __len__(self):
        if self.drop_last:
            return self.n // self.batch_size
        return -(-self.n // self.batch_size)

    def __iter__(self):
        if self.shuffle:
            perm = torch.randperm(self.n, device=self.device, generator=self.generator)
        for b in range(len(self)):
            lo = b * self.batch_size
            hi = min(lo + self.batch_size, self.n)
            if self.shuffle:
                idx = perm[lo:hi]
                if self.ops is not None:
                    yield self.ops.gather_points(self.X, self.S, idx)
                else:
                    yield self.X.index_select(0, idx), self.S.index_select(0, idx)
            else:
                yield self.X[lo:hi], self.S[lo:hi]

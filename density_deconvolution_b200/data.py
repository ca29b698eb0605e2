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
    if isinstance(data, (tuple, list)) and len(data) == 1 and torch.is_tensor(data[0]):
        return data[0], None                                    # SGDGMM: data[0] only (sgd_gmm.py:49-62)
    if torch.is_tensor(data):
        return data, None
    if hasattr(data, "tensors") and len(data.tensors) in (1, 2):            # torch TensorDataset
        return data.tensors[0], (data.tensors[1] if len(data.tensors) == 2 else None)
    items = [data[i] for i in range(len(data))]
    if not isinstance(items[0], (tuple, list)):
        return torch.stack([torch.as_tensor(x) for x in items]), None
    cols = list(zip(*items))
    X = torch.stack([torch.as_tensor(x) for x in cols[0]])
    S = torch.stack([torch.as_tensor(s) for s in cols[1]]) if len(cols) > 1 else None
    return X, S


class DeviceBatches:
    """Iterates ``(X_b, S_b)`` minibatches that already live on ``device``.

    The whole table is uploaded once (float32); every epoch draws a permutation on the device
    and gathers batches there (one ``dgmm_gather_points_f32`` launch per batch when ``ops`` is given).
    Semantics of ``DataLoader(batch_size, shuffle, drop_last)``.  The noise may be per point
    (``[N,D,D]`` full, ``[N,D]`` diagonal: gathered with the points), ONE matrix for all points
    (``[D,D]``: handed out unchanged with every batch) or absent (``None``: batches are ``(X_b, None)``).

    Under ``torch.distributed`` every minibatch step ends in an all-reduce, so all ranks must make the
    same number of batches: ``sync_ranks()`` truncates every rank to the minimum over ranks (shards
    differ by one point when ``N % world_size != 0``)."""

    def __init__(self, data, batch_size, device, shuffle=False, drop_last=False, generator=None, ops=None):
        X, S = as_tensors(data)
        self.X = X.to(device=device, dtype=torch.float32).contiguous()
        self.n = self.X.shape[0]
        D = self.X.shape[1]
        if S is None:
            self.layout, self.S = "none", None
        else:
            self.S = S.to(device=device, dtype=torch.float32).contiguous()
            if self.S.dim() == 3 and self.S.shape[0] == self.n:
                self.layout = "full"
            elif self.S.dim() == 2 and tuple(self.S.shape) == (D, D) and self.n != D:
                self.layout = "shared"
            elif self.S.dim() == 2 and tuple(self.S.shape) == (self.n, D):
                self.layout = "diag"
            elif self.S.dim() == 2 and tuple(self.S.shape) == (self.n, D * (D + 1) // 2):
                self.layout = "packed"
            else:
                raise ValueError("noise_covars has shape %s; expected [N,D,D], [N,D] or [D,D]" % (tuple(self.S.shape),))
        self.batch_size = int(batch_size) if batch_size else self.n
        self.shuffle = shuffle
        self.drop_last = drop_last
        self.device = device
        self.generator = generator
        self.max_batches = None
        # ops.gather_points: X and S rows of a shuffled batch in ONE launch (dgmm_gather_points_f32)
        self.ops = ops if (ops is not None and self.layout in ("full", "diag", "packed")) else None

    def _own_batches(self):
        if self.drop_last:
            return self.n // self.batch_size
        return -(-self.n // self.batch_size)

    def __len__(self):
        nb = self._own_batches()
        return nb if self.max_batches is None else min(nb, self.max_batches)

    def sync_ranks(self):
        """All ranks iterate the MINIMUM batch count over ranks (no-op for one process)."""
        import torch.distributed as td
        if td.is_available() and td.is_initialized() and td.get_world_size() > 1:
            t = torch.tensor([self._own_batches()], dtype=torch.int64, device=self.device)
            td.all_reduce(t, op=td.ReduceOp.MIN)
            self.max_batches = int(t.item())
        return self

    def __iter__(self):
        if self.shuffle:
            perm = torch.randperm(self.n, device=self.device, generator=self.generator)
        per_point = self.layout in ("full", "diag", "packed")
        for b in range(len(self)):
            lo = b * self.batch_size
            hi = min(lo + self.batch_size, self.n)
            if self.shuffle:
                idx = perm[lo:hi]
                if self.ops is not None:
                    yield self.ops.gather_points(self.X, self.S, idx)
                elif per_point:
                    yield self.X.index_select(0, idx), self.S.index_select(0, idx)
                else:
                    yield self.X.index_select(0, idx), self.S
            else:
                yield self.X[lo:hi], (self.S[lo:hi] if per_point else self.S)

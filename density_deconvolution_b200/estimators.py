"""The reference's ``deconv.gmm`` estimator API on top of the sm_100a kernels.

Same class names, constructor arguments, attributes and ``fit / score / score_batch /
predict / posterior_params`` semantics as bayesiains/density-deconvolution (file:line cited on
every method), but the per-(point, component) mathematics runs in ``libdgmm_b200.so``:

* one fused kernel pass per EM iteration / SGD step produces ``sum_i ll_i`` and the packed
  statistics ``(s0, s1, S2)`` - no ``[N,K,D,D]`` tensor is ever
  materialised on the fit path;
* with ``torch.distributed`` initialised (one process per GPU, every rank holding its own
  shard of the points) the packed buffer is all-reduced once per step and every rank applies
  the identical parameter update (``dist.py``);
* data is uploaded once and minibatches are gathered on the device (``data.DeviceBatches``)
  instead of going through a per-sample ``DataLoader``.

There is no CPU fallback: tensors that reach ``ops`` must be CUDA tensors and a missing
``libdgmm_b200.so`` raises ``DgmmError``.
"""
import math

import torch
import torch.nn as nn

from . import dist as _dist
from . import ops as _ops
from . import stats_algebra as _alg
from .data import DeconvDataset, DeviceBatches, SGDDeconvDataset, as_tensors  # noqa: F401
from .util import k_means, minibatch_k_means


def _default_device(device):
    if device is None or device == "":
        if not torch.cuda.is_available():
            # keep construction possible (parameters on the CPU); the first kernel call raises
            return torch.device("cpu")
        return torch.device("cuda", torch.cuda.current_device())
    return torch.device(device)


def _resident(t, device):
    """float32, contiguous, on ``device`` (uploaded once; a no-op for resident inputs)."""
    return torch.as_tensor(t).to(device=device, dtype=torch.float32).contiguous()


def _noise_layout(X, S):
    """The reference contract is ``S [N,D,D]``; ``[N,D]`` (diagonals) and ``[D,D]`` (one matrix
    for every point) are accepted as cheaper layouts of the same thing."""
    N, D = X.shape
    if S is None:
        return "none"
    if S.dim() == 3:
        return "full"
    if S.dim() == 2 and tuple(S.shape) == (N, D) and not (N == D):
        return "diag"
    if S.dim() == 2 and tuple(S.shape) == (D, D):
        return "shared"
    if S.dim() == 2 and tuple(S.shape) == (N, D):
        return "diag"
    if S.dim() == 2 and tuple(S.shape) == (N, D * (D + 1) // 2):
        return "packed"
    raise ValueError("noise_covars has shape %s; expected [N,D,D], [N,D(D+1)/2], [N,D] or [D,D]" % (tuple(S.shape),))


def _global_count(n, device):
    if not _dist.is_distributed():
        return n
    t = torch.tensor([n], dtype=torch.int64, device=device)
    torch.distributed.all_reduce(t)
    return int(t.item())


def _global_batch_sizes(loader):
    """Global size of every minibatch the (rank-synchronised) loader will yield: the sum over ranks of
    ``min(batch, n_r - b * batch)``.  Computed once per fit from an all-gather of the shard sizes, so the
    per-step loss scale is right even when the ranks' last batches differ (``N % world_size != 0``)."""
    nb, bs = len(loader), loader.batch_size
    if _dist.is_distributed():
        t = torch.zeros(_dist.world_size(), dtype=torch.int64, device=loader.device)
        t[_dist.rank()] = loader.n
        torch.distributed.all_reduce(t)
        shard = t.tolist()
    else:
        shard = [loader.n]
    return [sum(max(0, min(bs, n_r - b * bs)) for n_r in shard) for b in range(nb)]


class _Mixture:
    """Device-resident ``(alpha, log_alpha, mu, Sigma)`` in the layout the C ABI takes."""

    def __init__(self, alpha, mu, Sigma):
        self.alpha = alpha.reshape(-1).contiguous()
        self.log_alpha = torch.log(self.alpha)
        self.mu = mu.contiguous()
        self.Sigma = Sigma.contiguous()


# =========================================================================================
# EM estimators
# =========================================================================================
class DeconvGMM:
    """Full-batch EM deconvolution GMM (reference: deconv/gmm/deconv_gmm.py:11-113 on
    deconv/gmm/base.py:11-70).

    ``fit((X, noise_covars))`` runs ``epochs`` iterations of: fused E-step statistics kernel ->
    (all-reduce) -> closed-form M-step kernel, stopping when the summed log-likelihood moves
    by less than ``tol`` or becomes ``-inf`` (base.py:46-55).  Attributes ``weights [K,1]``,
    ``means [K,D]``, ``covars [K,D,D]`` as in the reference (deconv_gmm.py:109)."""

    def __init__(self, components, dimensions, epochs=1000, w=1e-6, tol=1e-6, device=None):
        self.k = components
        self.d = dimensions
        self.epochs = epochs
        self.tol = tol
        self.device = _default_device(device)
        self.w_scalar = float(w)
        self.w = w * torch.eye(self.d, device=self.device)            # deconv_gmm.py:19
        self.weights = torch.full((self.k, 1), 1.0 / self.k, device=self.device)
        self.means = torch.zeros(self.k, self.d, device=self.device)
        self.covars = torch.eye(self.d, device=self.device).repeat(self.k, 1, 1)
        self.n_iter_ = 0
        self.log_prob_curve = []

    # -- helpers ---------------------------------------------------------------------------
    def _prep(self, data):
        X, S = as_tensors(data)
        X = _resident(X, self.device)
        S = _resident(S, self.device) if S is not None else None
        return X, S, _noise_layout(X, S)

    def _mixture(self):
        return _Mixture(self.weights.to(torch.float32), self.means.to(torch.float32), self.covars.to(torch.float32))

    def set_params(self, weights, means, covars):
        """Inject parameters (parity harnesses never go through k-means)."""
        self.weights = _resident(weights, self.device).reshape(self.k, 1).clone()
        self.means = _resident(means, self.device).reshape(self.k, self.d).clone()
        self.covars = _resident(covars, self.device).reshape(self.k, self.d, self.d).clone()
        return self

    # -- initialisation (deconv_gmm.py:21-34 + the first _m_step of base.py:41-42) -----------
    def _kmeans_init(self, X, max_iters=50, tol=1e-9):
        return k_means(X, self.k, max_iters, tol, self.device, ops=_ops)[0]

    def _init_expectations(self, data):
        X, _, _ = self._prep(data)
        n = X.shape[0]
        resps = self._kmeans_init(X)
        log_resps = torch.log(resps)
        cond_means = X[:, None, :].expand(n, self.k, self.d)
        cond_covars = torch.eye(self.d, device=self.device).expand(n, self.k, self.d, self.d)
        return (log_resps, cond_means, cond_covars)

    def _init_from_labels(self, X, labels, centroids=None):
        """The M-step the reference applies to the k-means expectations (cond_means = X,
        cond_covars = I; deconv_gmm.py:21-34,81-109) from the hard LABELS: per-label sums by
        ``index_add_`` over chunks of points, so neither the ``[N,K,D,D]`` broadcast of the reference nor
        any ``[N,K,.]`` intermediate exists (N = 1e7 at the benchmark shapes).  An empty cluster (NaN
        parameters in the reference) is seeded with one pseudo-point at its centroid."""
        n_total = _global_count(X.shape[0], self.device)
        if labels.dim() == 2:                                          # one-hot responsibilities
            labels = labels.argmax(dim=1)
        labels = labels.long()
        d = self.d
        s0 = torch.zeros(self.k, device=self.device, dtype=torch.float64)
        sx = torch.zeros(self.k, d, device=self.device, dtype=torch.float64)
        sxx = torch.zeros(self.k, d * d, device=self.device, dtype=torch.float64)
        chunk = 1 << 20
        for lo in range(0, X.shape[0], chunk):
            xb = X[lo:lo + chunk].double()
            lb = labels[lo:lo + chunk]
            s0.index_add_(0, lb, torch.ones_like(lb, dtype=torch.float64))
            sx.index_add_(0, lb, xb)
            sxx.index_add_(0, lb, (xb[:, :, None] * xb[:, None, :]).reshape(-1, d * d))
        sxx = sxx.view(self.k, d, d)
        if _dist.is_distributed():
            for t in (s0, sx, sxx):
                torch.distributed.all_reduce(t)
        if centroids is not None:
            empty = s0 <= 0
            c64 = centroids.double()
            s0 = torch.where(empty, torch.ones_like(s0), s0)
            sx = torch.where(empty[:, None], c64, sx)
            sxx = torch.where(empty[:, None, None], c64[:, :, None] * c64[:, None, :], sxx)
        means = sx / s0[:, None]
        eye = torch.eye(d, device=self.device, dtype=torch.float64)
        scatter = sxx - s0[:, None, None] * means[:, :, None] * means[:, None, :]
        covars = (scatter + s0[:, None, None] * eye + 2 * self.w_scalar * eye) / s0[:, None, None]
        self.means, self.covars = means.float(), covars.float()
        self.weights = (s0 / n_total)[:, None].float()

    # -- the hot loop (base.py:37-55) ---------------------------------------------------------
    def fit(self, data, init=True):
        X, S, layout = self._prep(data)
        n_total = _global_count(X.shape[0], self.device)
        if init:
            labels, centroids = k_means(X, self.k, 50, 1e-9, self.device, ops=_ops, return_labels=True)
            self._init_from_labels(X, labels, centroids)
        mix = self._mixture()
        stats = torch.empty(_ops.stats_len(self.d, self.k), dtype=torch.float32, device=self.device)
        prev = -math.inf
        self.log_prob_curve = []
        self.n_iter_ = 0
        for _ in range(self.epochs):
            _ops.estep_stats(X, S, mix.log_alpha, mix.mu, mix.Sigma, noise=layout, want_ll=False, out_stats=stats)
            _dist.allreduce_stats(stats)
            tail = stats[-4:].tolist()                                 # the one host sync of the iteration
            log_prob = -math.inf if tail[3] > 0 else tail[0] + tail[1]
            self.log_prob_curve.append(log_prob)
            if log_prob == -math.inf or math.isnan(log_prob):          # base.py:48-49
                break
            _ops.mstep_update(stats, n_total, self.w_scalar, mix.mu, mix.Sigma, mix.alpha, mix.log_alpha)
            self.n_iter_ += 1
            if abs(log_prob - prev) < self.tol:                        # base.py:52-53
                break
            prev = log_prob
        self.weights, self.means, self.covars = mix.alpha[:, None], mix.mu, mix.Sigma
        return self

    # -- reference-shaped E/M steps (small N: they materialise [N,K,D,D]) ---------------------
    def _e_step(self, data):
        """``(sum_i ll_i, (log_resps [N,K], cond_means [N,K,D], cond_covars [N,K,D,D]))``
        (deconv_gmm.py:39-79); ``(-inf, None)`` when some ``Sigma_j + S_i`` is not positive
        definite (deconv_gmm.py:46-47)."""
        X, S, layout = self._prep(data)
        mix = self._mixture()
        ll, log_r, b, B, status = _ops.estep_full(X, S, mix.log_alpha, mix.mu, mix.Sigma, noise=layout)
        if int(status.item()) != 0:
            return torch.tensor(float("-inf")), None
        return ll.sum(), (log_r, b, B)

    def _m_step(self, data, expectations):
        """deconv_gmm.py:81-109 on materialised expectations (vectorised over components; the
        covariance is centred on the NEW means and regularised with ``2 w I`` before the
        division, exactly as the reference)."""
        log_resps, cond_means, cond_covars = expectations
        n = cond_means.shape[0]
        resps = torch.exp(log_resps)
        s0 = resps.sum(dim=0)
        means = torch.einsum("nk,nkd->kd", resps, cond_means) / s0[:, None]
        dev = means[None] - cond_means
        scatter = torch.einsum("nk,nkab->kab", resps, cond_covars) + torch.einsum("nk,nka,nkb->kab", resps, dev, dev)
        self.covars = (scatter + 2 * self.w) / s0[:, None, None]
        self.means = means
        self.weights = (s0 / n)[:, None]

    def predict(self, data):
        """Responsibilities ``[N,K]`` (deconv_gmm.py:36-37)."""
        X, S, layout = self._prep(data)
        mix = self._mixture()
        _, log_r, _, _, _ = _ops.estep_full(X, S, mix.log_alpha, mix.mu, mix.Sigma, noise=layout,
                                            want_means=False, want_covars=False)
        return torch.exp(log_r)

    def score(self, data):
        """Sum of the per-point log-likelihoods as a 0-d tensor.  The reference's ``score``
        (deconv_gmm.py:111-113) forgets to call ``_e_step``; this returns what it meant to."""
        X, S, layout = self._prep(data)
        mix = self._mixture()
        _, sums = _ops.loglik(X, S, mix.log_alpha, mix.mu, mix.Sigma, noise=layout, want_sum=True)
        if _dist.is_distributed():
            torch.distributed.all_reduce(sums)
        # float64: at N ~ 1e6 the summed log-likelihood exceeds float32's integer range and two epochs would
        # compare equal (a spurious "converged"); the reference accumulates per-batch .item() in a Python double
        return torch.where(sums[1] > 0, torch.full_like(sums[0], -math.inf), sums[0])

    def score_samples(self, data):
        """Per-point log-likelihood ``[N]``."""
        X, S, layout = self._prep(data)
        mix = self._mixture()
        return _ops.loglik(X, S, mix.log_alpha, mix.mu, mix.Sigma, noise=layout)


class OnlineDeconvGMM(DeconvGMM):
    """Minibatch stochastic-approximation EM (reference: deconv/gmm/online_deconv_gmm.py:8-205).

    Per minibatch: one fused statistics pass with the reference's ``10 eps`` responsibility
    floor (online_deconv_gmm.py:150,154) -> all-reduce -> the exponential-moving-average update
    of ``sum_resps / sum_cond_means`` and the moment-shifted covariance blend
    (online_deconv_gmm.py:134-188), all on ``[K,.]`` tensors.  Under ``torch.distributed`` the
    global minibatch is ``world_size x batch_size`` (every rank contributes ``batch_size`` points
    of its shard)."""

    def __init__(self, components, dimensions, epochs=1000, w=1e-6, tol=1e-6, step_size=0.1, batch_size=100,
                 max_no_improvement=20, k_means_factor=100, k_means_iters=10, lr_step=10, lr_gamma=0.1,
                 device=None):
        super().__init__(components, dimensions, epochs=epochs, w=w, tol=tol, device=device)
        self.batch_size = batch_size
        self.step_size = step_size
        self.max_no_improvement = max_no_improvement
        self.k_means_factor = k_means_factor
        self.k_means_iters = k_means_iters
        self.lr_step = lr_step
        self.lr_gamma = lr_gamma
        self.resp_floor = 10 * torch.finfo(torch.float32).eps

    def _global_batch(self):
        return self.batch_size * _dist.world_size()

    def _init_sum_stats(self, loader, n):
        """online_deconv_gmm.py:26-40: minibatch k-means -> weights/means, identity covariances."""
        counts, centroids = minibatch_k_means(loader, self.k, max_iters=self.k_means_iters, device=self.device, ops=_ops)
        self.set_params(counts[:, None] / counts.sum(), centroids,
                        torch.eye(self.d, device=self.device).repeat(self.k, 1, 1))
        self._reset_running_sums()

    def _reset_running_sums(self):
        self.sum_resps = self.weights * self._global_batch()
        self.sum_cond_means = self.means * self.sum_resps

    def _m_step_stats(self, stats, step_size, log_alpha=None, crash_flag=None):
        """online_deconv_gmm.py:146-188 on the fused statistics: ONE kernel
        (``dgmm_online_mstep_f32``) moves ``sum_resps, sum_cond_means, means, covars, weights`` in place."""
        for name in ("weights", "means", "covars", "sum_resps", "sum_cond_means"):
            t = getattr(self, name)
            if t.dtype != torch.float32 or not t.is_contiguous():
                setattr(self, name, t.to(torch.float32).contiguous())
        if log_alpha is None:
            log_alpha = torch.empty(self.k, dtype=torch.float32, device=self.device)
        _ops.online_mstep(stats, self.resp_floor, step_size, self.w_scalar, self._global_batch(), self.means,
                          self.covars, self.sum_resps.view(-1), self.sum_cond_means, self.weights.view(-1),
                          log_alpha, crash_flag)
        return log_alpha

    def _m_step(self, expectations, n, step_size):
        """Reference signature (online_deconv_gmm.py:146) on materialised expectations."""
        log_resps, cond_means, cond_covars = expectations
        f = self.resp_floor
        r = torch.exp(log_resps) + f
        s0 = r.sum(dim=0)
        sum_resps = (s0 + f)[:, None]
        sum_cond_means = torch.einsum("nk,nkd->kd", r, cond_means)
        batch_means = sum_cond_means / sum_resps
        dev = cond_means - batch_means
        batch_covars = (torch.einsum("nk,nka,nkb->kab", r, dev, dev) + torch.einsum("nk,nkab->kab", r, cond_covars))
        batch_covars = batch_covars / sum_resps[:, :, None]
        m_old, sum_old = self.means, self.sum_resps
        self.sum_resps = (1 - step_size) * sum_old + step_size * sum_resps
        self.sum_cond_means = (1 - step_size) * self.sum_cond_means + step_size * sum_cond_means
        self.means = self.sum_cond_means / self.sum_resps
        self.covars = ((1 - step_size) * _alg.moment_shift(self.covars - self.w, sum_old / self.sum_resps, m_old,
                                                           self.means)
                       + step_size * _alg.moment_shift(batch_covars, sum_resps / self.sum_resps, batch_means,
                                                       self.means) + self.w)
        self.weights = self.sum_resps / self._global_batch()

    def fit(self, data, val_data=None, verbose=False, interval=1, init=True):
        X, S, layout = self._prep(data)
        loader = DeviceBatches((X, S), self.batch_size, self.device, shuffle=True, drop_last=True, ops=_ops).sync_ranks()
        if init:
            init_loader = DeviceBatches((X, S), self.k_means_factor * self.batch_size, self.device, shuffle=True,
                                        ops=_ops).sync_ranks()
            self._init_sum_stats(init_loader, X.shape[0])
        elif not hasattr(self, "sum_resps"):
            self._reset_running_sums()
        self.train_ll_curve = []
        if val_data:
            self.val_ll_curve = []
        stats = torch.empty(_ops.stats_len(self.d, self.k), dtype=torch.float32, device=self.device)
        prev_ll, max_val_ll, no_improvements = -math.inf, -math.inf, 0
        # per minibatch: statistics kernel -> (all-reduce) -> M-step kernel; no host sync inside an epoch
        # (a non-PD minibatch freezes the state through the sticky device flag, read once per epoch)
        self.weights, self.means, self.covars = (t.to(torch.float32).contiguous()
                                                 for t in (self.weights, self.means, self.covars))
        log_alpha = torch.log(self.weights.view(-1))
        crash_flag = torch.zeros(1, dtype=torch.int32, device=self.device)
        for i in range(self.epochs):
            seen = 0
            for Xb, Sb in loader:
                _ops.estep_stats(Xb, Sb, log_alpha, self.means, self.covars, resp_floor=self.resp_floor, noise=layout,
                                 want_ll=False, out_stats=stats)
                _dist.allreduce_stats(stats)
                self._m_step_stats(stats, self.step_size, log_alpha, crash_flag)  # online_deconv_gmm.py:79-80 inside
                seen += 1
            crashed = bool(crash_flag.item())
            if crashed or seen == 0:                                             # :84-88
                print("Log prob 0, crashed.")
                break
            train_ll = self.score_batch((X, S))                                  # :90
            self.train_ll_curve.append(train_ll)
            if val_data:
                val_ll = self.score_batch(val_data)
                self.val_ll_curve.append(val_ll)
            if (i + 1) % self.lr_step == 0:                                      # :97-98
                self.step_size *= self.lr_gamma
            if verbose and i % interval == 0:
                if val_data:
                    print("Epoch {}, Train LL: {}, Val LL: {}".format(i, train_ll, val_ll))
                else:
                    print("Epoch {}, Train LL: {}".format(i, train_ll))
            if abs(train_ll - prev_ll) < self.tol:                               # :112-116
                print("Train LL converged within tolerance at {}".format(train_ll))
                break
            if val_data:
                if val_ll > max_val_ll:
                    no_improvements, max_val_ll = 0, val_ll
                else:
                    no_improvements += 1
                if no_improvements > self.max_no_improvement:
                    print("No improvement in val LL for {} epochs. Early Stopping at {}".format(
                        self.max_no_improvement, val_ll))
                    break
            prev_ll = train_ll
        return self

    def score_batch(self, dataset):
        """Sum of log-likelihoods over a dataset as a Python float (online_deconv_gmm.py:190-205):
        one kernel launch over the resident table instead of a loader loop."""
        return float(self.score(dataset).item())


class BatchDeconvGMM(OnlineDeconvGMM):
    """Exact EM with accumulated sufficient statistics (reference:
    deconv/gmm/batch_deconv_gmm.py:7-120).  The reference needs two passes over the loader per
    epoch (means first, then covariances centred on the new means); with the fused statistics the
    covariance can be re-centred after the fact, so an epoch is ONE pass +
    one all-reduce + the closed-form update.  As in the reference all ``epochs`` are run
    (batch_deconv_gmm.py:37,89: ``prev_ll`` is never updated)."""

    def fit(self, data, val_data=None, verbose=False, interval=1, init=True):
        X, S, layout = self._prep(data)
        n_total = _global_count(X.shape[0], self.device)
        if init:
            init_loader = DeviceBatches((X, S), self.k_means_factor * self.batch_size, self.device, shuffle=True,
                                        ops=_ops).sync_ranks()
            self._init_sum_stats(init_loader, X.shape[0])
        self.train_ll_curve = []
        if val_data:
            self.val_ll_curve = []
        mix = self._mixture()
        stats = torch.empty(_ops.stats_len(self.d, self.k), dtype=torch.float32, device=self.device)
        prev_ll = -math.inf
        for i in range(self.epochs):
            _ops.estep_stats(X, S, mix.log_alpha, mix.mu, mix.Sigma, noise=layout, want_ll=False, out_stats=stats)
            _dist.allreduce_stats(stats)
            tail = stats[-4:].tolist()
            train_ll = -math.inf if tail[3] > 0 else tail[0] + tail[1]
            self.train_ll_curve.append(train_ll)
            if val_data:
                self.weights, self.means, self.covars = mix.alpha[:, None], mix.mu, mix.Sigma
                val_ll = self.score_batch(val_data)
                self.val_ll_curve.append(val_ll)
            if train_ll == -math.inf:
                break
            _ops.mstep_update(stats, n_total, self.w_scalar, mix.mu, mix.Sigma, mix.alpha, mix.log_alpha)
            if verbose and i % interval == 0:
                if val_data:
                    print("Epoch {}, Train LL: {}, Val LL: {}".format(i, train_ll, val_ll))
                else:
                    print("Epoch {}, Train LL: {}".format(i, train_ll))
            if abs(train_ll - prev_ll) < self.tol:
                print("Train LL converged within tolerance at {}".format(train_ll))
                break
        self.weights, self.means, self.covars = mix.alpha[:, None], mix.mu, mix.Sigma
        return self

    def _step_1(self, data):
        """batch_deconv_gmm.py:95-104: ``(ll_sum, (sum_r [K,1], sum_r_b [K,D]))``."""
        X, S, layout = self._prep(data)
        mix = self._mixture()
        _, stats = _ops.estep_stats(X, S, mix.log_alpha, mix.mu, mix.Sigma, noise=layout, want_ll=False)
        s0, s1, _, tail = _ops.unpack_stats(stats, self.d, self.k)
        return tail[0] + tail[1], (s0[:, None], _alg.sum_resp_means(s0, s1, mix.mu, mix.Sigma))

    def _step_2(self, data, new_means):
        """batch_deconv_gmm.py:106-120: ``sum_i r (B + (b - m)(b - m)^T)`` for ``m = new_means``."""
        X, S, layout = self._prep(data)
        mix = self._mixture()
        _, stats = _ops.estep_stats(X, S, mix.log_alpha, mix.mu, mix.Sigma, noise=layout, want_ll=False)
        s0, s1, S2, _ = _ops.unpack_stats(stats, self.d, self.k)
        return _alg.centred_scatter(s0, s1, S2, mix.mu, mix.Sigma, _resident(new_means, self.device))


# =========================================================================================
# SGD estimator
# =========================================================================================
class _DeconvLogLik(torch.autograd.Function):
    """Differentiable per-point ``ll_i`` (deconv/gmm/sgd_deconv_gmm.py:14-28).  Forward = the
    log-likelihood kernel; backward = ONE fused statistics pass with ``g = grad_output``:
    ``d/d log_alpha = s0``, ``d/d mu = s1``, ``d/d Sigma = S2 / 2``.
    No gradient flows to the data ``(x, noise_covars)``."""

    @staticmethod
    def forward(ctx, X, S, log_alpha, mu, Sigma, layout):
        ll, sums = _ops.loglik(X, S, log_alpha, mu, Sigma, noise=layout, want_sum=True)
        ctx.save_for_backward(X, S if S is not None else X.new_empty(0), log_alpha, mu, Sigma)
        ctx.layout = layout
        ctx.nonpd = sums
        return ll

    @staticmethod
    def backward(ctx, grad_ll):
        X, S, log_alpha, mu, Sigma = ctx.saved_tensors
        K, D = mu.shape
        _, stats = _ops.estep_stats(X, S if ctx.layout != "none" else None, log_alpha, mu, Sigma,
                                    g=grad_ll.contiguous(), noise=ctx.layout, want_ll=False)
        s0, s1, S2, _ = _ops.unpack_stats(stats, D, K)
        return None, None, s0.reshape(log_alpha.shape), s1, 0.5 * S2, None


class _DeconvLogLikSum(torch.autograd.Function):
    """``sum_i ll_i`` over the (global) minibatch with its parameter gradients from the SAME
    kernel pass: the fit loop's loss is ``-mean(ll)``, whose upstream gradient is a constant, so
    forward and backward of deconv/gmm/sgd_gmm.py:173-181 collapse into one fused statistics
    launch (+ one all-reduce under ``torch.distributed``)."""

    @staticmethod
    def forward(ctx, X, S, log_alpha, mu, Sigma, layout):
        K, D = mu.shape
        _, stats = _ops.estep_stats(X, S, log_alpha, mu, Sigma, noise=layout, want_ll=False)
        _dist.allreduce_stats(stats)
        s0, s1, S2, tail = _ops.unpack_stats(stats, D, K)
        ctx.save_for_backward(s0, s1, S2)
        ctx.la_shape = log_alpha.shape
        total = tail[0] + tail[1]
        return torch.where(tail[3] > 0, torch.full_like(total, float("nan")), total)

    @staticmethod
    def backward(ctx, grad_sum):
        s0, s1, S2 = ctx.saved_tensors
        return None, None, (grad_sum * s0).reshape(ctx.la_shape), grad_sum * s1, (0.5 * grad_sum) * S2, None


class SGDGMMModule(nn.Module):
    """Parametrisation of deconv/gmm/sgd_gmm.py:14-65: softmax weights and
    ``Sigma_j = L_j L_j^T`` with ``L_j = strict_lower(l_lower_j) + diag(exp(l_diag_j))``.
    ``forward`` without noise is the ``S = 0`` case of the same kernel (sgd_gmm.py:49-62)."""

    def __init__(self, components, dimensions, w, device=None):
        super().__init__()
        self.k = components
        self.d = dimensions
        self.device = _default_device(device)
        self.soft_weights = nn.Parameter(torch.zeros(self.k))
        self.soft_max = nn.Softmax(dim=0)
        self.means = nn.Parameter(torch.rand(self.k, self.d))
        self.l_diag = nn.Parameter(torch.zeros(self.k, self.d))
        self.l_lower = nn.Parameter(torch.zeros(self.k, self.d * (self.d - 1) // 2))
        self.register_buffer("l_idx", torch.tril_indices(self.d, self.d, -1), persistent=False)
        self.register_buffer("w", w * torch.eye(self.d), persistent=False)
        self.to(self.device)

    @property
    def L(self):
        L = torch.diag_embed(torch.exp(self.l_diag))
        if self.l_idx.shape[1]:
            flat = self.l_idx[0] * self.d + self.l_idx[1]
            low = torch.zeros(self.k, self.d * self.d, device=self.l_lower.device, dtype=self.l_lower.dtype)
            L = L + low.index_copy(1, flat, self.l_lower).view(self.k, self.d, self.d)
        return L

    @property
    def covars(self):
        L = self.L
        return L @ L.transpose(-2, -1)

    def _mixture_tensors(self):
        return torch.log_softmax(self.soft_weights, dim=0), self.means, self.covars

    def forward(self, data):
        x = _resident(data[0], self.means.device)
        log_alpha, mu, Sigma = self._mixture_tensors()
        return _DeconvLogLik.apply(x, None, log_alpha, mu, Sigma, "none")

    def log_prob(self, data):
        return self.forward([data])

    def _data(self, data):
        """``(x, S, layout)`` of one minibatch: this module reads ``data[0]`` only (sgd_gmm.py:49-62)."""
        return _resident(data[0], self.means.device), None, "none"

    def log_prob_sum(self, data):
        """``sum_i ll_i`` (all-reduced over ranks) with gradients from a single kernel pass."""
        x, S, layout = self._data(data)
        log_alpha, mu, Sigma = self._mixture_tensors()
        return _DeconvLogLikSum.apply(x, S, log_alpha, mu, Sigma, layout)


class SGDDeconvGMMModule(SGDGMMModule):
    """deconv/gmm/sgd_deconv_gmm.py:12-28: ``forward((x, noise_covars)) -> ll [B]``, differentiable
    w.r.t. ``soft_weights, means, l_diag, l_lower``."""

    def _data(self, data):
        x = _resident(data[0], self.means.device)
        S = _resident(data[1], self.means.device) if data[1] is not None else None
        return x, S, _noise_layout(x, S)

    def forward(self, data):
        x, S, layout = self._data(data)
        log_alpha, mu, Sigma = self._mixture_tensors()
        return _DeconvLogLik.apply(x, S, log_alpha, mu, Sigma, layout)


class BaseSGDGMM:
    """Adam / MultiStepLR training loop of deconv/gmm/sgd_gmm.py:68-261."""

    def __init__(self, components, dimensions, epochs=10000, lr=1e-3, batch_size=64, tol=1e-6,
                 max_no_improvement=20, k_means_factor=100, w=1e-6, k_means_iters=10, lr_step=5, lr_gamma=0.1,
                 device=None):
        self.k = components
        self.d = dimensions
        self.epochs = epochs
        self.batch_size = batch_size
        self.tol = 1e-6                                    # sgd_gmm.py:78 (the ctor argument is ignored)
        self.lr = lr
        self.w = w
        self.k_means_factor = k_means_factor
        self.k_means_iters = k_means_iters
        self.max_no_improvement = max_no_improvement
        self.device = _default_device(device)
        self.module.to(self.device)
        self.optimiser = torch.optim.Adam(params=self.module.parameters(), lr=self.lr)
        self.scheduler = torch.optim.lr_scheduler.MultiStepLR(self.optimiser, milestones=[lr_step, lr_step + 5],
                                                              gamma=lr_gamma)

    @property
    def means(self):
        return self.module.means.detach()

    @property
    def covars(self):
        return self.module.covars.detach()

    @property
    def weights(self):
        return torch.softmax(self.module.soft_weights.detach(), dim=0)

    def reg_loss(self, n, n_total):
        """sgd_gmm.py:110-112."""
        l = (n / n_total) * self.w / torch.diagonal(self.module.covars, dim1=-1, dim2=-2)
        return l.sum()

    def init_params(self, loader):
        """sgd_gmm.py:256-261."""
        counts, centroids = minibatch_k_means(loader, self.k, max_iters=self.k_means_iters, device=self.device, ops=_ops)
        with torch.no_grad():
            self.module.soft_weights.copy_(torch.log(counts / counts.sum()))
            self.module.means.copy_(centroids)
            self.module.l_diag.zero_()
            self.module.l_lower.zero_()

    def _batches(self, data, shuffle, batch_size=None):
        X, S = as_tensors(data)
        if not isinstance(self.module, SGDDeconvGMMModule):
            S = None                                       # SGDGMM reads data[0] only (sgd_gmm.py:49-62)
        return DeviceBatches((X, S), batch_size or self.batch_size, self.device, shuffle=shuffle and self.shuffle,
                             ops=_ops).sync_ranks()

    # ---- fused optimiser step: no autograd graph, five launches per minibatch -------------------
    fused_step = True   # False: the autograd route (module.log_prob_sum -> loss.backward -> optimiser.step)
    shuffle = True      # DataLoader(shuffle=True) of sgd_gmm.py:139-145; tests fix the batch order

    def _fused_step_ok(self):
        g = self.optimiser.param_groups[0]
        return (self.fused_step and isinstance(self.module, SGDGMMModule)
                and isinstance(self.optimiser, torch.optim.Adam) and len(self.optimiser.param_groups) == 1
                and not g.get("weight_decay") and not g.get("amsgrad") and not g.get("maximize"))

    def _adam_state(self):
        """The parameters and the torch optimiser's OWN state tensors (created here the way
        ``torch.optim.Adam`` creates them on its first step), so ``optimiser.state`` stays truthful."""
        params = [self.module.soft_weights, self.module.means, self.module.l_diag, self.module.l_lower]
        states = []
        for prm in params:
            st = self.optimiser.state[prm]
            if len(st) == 0:
                st["step"] = torch.tensor(0.0)
                st["exp_avg"] = torch.zeros_like(prm, memory_format=torch.preserve_format)
                st["exp_avg_sq"] = torch.zeros_like(prm, memory_format=torch.preserve_format)
            states.append(st)
        return params, states

    peer_reduce = True   # distributed fused step: sum the ranks' statistics inside the Adam kernels (NVLink peer memory)

    def _peer_reduce(self):
        """The symmetric buffer of the fused cross-GPU reduction (created once per estimator; collective).  ``None``
        when it is switched off (``peer_reduce = False`` / ``DGMM_PEER_REDUCE=0``), the process group is not NCCL, or the
        ranks do not all get peer access: the step then all-reduces through NCCL (every rank agrees on the choice)."""
        import os
        if hasattr(self, "_peers"):
            return self._peers
        self._peers = None
        import torch.distributed as dist
        want = (self.peer_reduce and os.environ.get("DGMM_PEER_REDUCE", "1") != "0" and hasattr(_ops, "PeerReduce")
                and dist.get_backend() == "nccl" and _dist.world_size() <= 8)
        ok = torch.tensor([1 if want else 0], dtype=torch.int32, device=self.device)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        if int(ok.item()) == 0:
            return None
        peers = None
        try:
            peers = _ops.PeerReduce(self.d, self.k, self.device)
        except Exception:  # noqa: BLE001  (no symmetric memory on this box)
            peers = None
        ok = torch.tensor([1 if peers is not None else 0], dtype=torch.int32, device=self.device)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        self._peers = peers if int(ok.item()) == 1 else None
        return self._peers

    def _fused_epoch(self, loader, n_total, sizes):
        """One epoch of sgd_gmm.py:167-182 on the fused kernels: per minibatch ONE statistics pass
        (forward and backward of the loss), one all-reduce, one Adam step on the ``[K,.]`` parameters
        (``dgmm_sgd_adam_step_f32``).  Returns the epoch's mean log-likelihood (:183)."""
        params, states = self._adam_state()
        group = self.optimiser.param_groups[0]
        data = [prm.data for prm in params]
        exp_avg = [st["exp_avg"] for st in states]
        exp_avg_sq = [st["exp_avg_sq"] for st in states]
        step = int(states[0]["step"].item())
        log_alpha, Sigma = _ops.sgd_params(data[0], data[2], data[3])
        epoch_acc = torch.zeros(2, dtype=torch.float64, device=self.device)
        if hasattr(_ops, "SgdStepper") and loader.X.is_cuda:
            # one C call per minibatch: gather + statistics (+ all-reduce) + Adam (dgmm_sgd_fused_step_f32)
            stepper = _ops.SgdStepper(loader.X, loader.S, loader.layout, loader.batch_size, data, exp_avg, exp_avg_sq,
                                      log_alpha, Sigma, epoch_acc, self.w, group["betas"], group["eps"])
            perm = torch.randperm(loader.n, device=self.device, generator=loader.generator) if loader.shuffle else None
            distributed = _dist.is_distributed()
            peers = self._peer_reduce() if distributed else None
            bs = loader.batch_size
            for b in range(len(loader)):
                lo = b * bs
                nb = min(bs, loader.n - lo)
                step += 1
                if distributed and peers is not None:
                    # no collective call: the Adam kernels add the ranks' statistics over NVLink peer memory
                    stepper.step_peers(perm, lo, nb, sizes[b], n_total, group["lr"], step, peers)
                elif distributed:
                    stepper.step(perm, lo, nb, 1, sizes[b], n_total, group["lr"], step)
                    _dist.allreduce_stats(stepper.stats)
                    stepper.step(perm, lo, nb, 2, sizes[b], n_total, group["lr"], step)
                else:
                    stepper.step(perm, lo, nb, 3, sizes[b], n_total, group["lr"], step)
        else:
            stats = None
            for b, (X_b, S_b) in enumerate(loader):
                n = sizes[b]                                                         # global minibatch
                _, stats = _ops.estep_stats(X_b, S_b, log_alpha, data[1], Sigma, noise=loader.layout,
                                            want_ll=False, out_stats=stats)
                _dist.allreduce_stats(stats)
                step += 1
                _ops.sgd_adam_step(stats, n, n_total, self.w, group["lr"], group["betas"], group["eps"], step, data,
                                   exp_avg, exp_avg_sq, log_alpha, Sigma, epoch_acc)
        for st in states:
            st["step"].fill_(float(step))
        self.optimiser._opt_called = True   # what optimiser.step() records for the lr scheduler's order check
        acc = epoch_acc.cpu()
        peers = getattr(self, "_peers", None)
        if peers is not None and _dist.is_distributed():
            # every rank's last Adam kernel has read its peers' buffers before anybody moves on (and could free one);
            # a peer that never published its statistics within the kernels' bounded wait is an error of its own
            timed_out = bool(peers.timeout.item())
            torch.distributed.barrier()
            if timed_out:
                raise _ops.DgmmError("fused cross-GPU reduction: a peer rank did not publish its statistics within the "
                                     "time limit (is every rank running the same number of minibatch steps?)")
        if acc[1].item() > 0:
            return math.nan
        return acc[0].item() / n_total

    # the `_tim` fork of the loop (deconv/gmm/sgd_gmm_tim.py:112-227) differs in bookkeeping only: see BaseSGDGMMTim
    _tim = False

    def fit(self, data, val_data=None, verbose=False, interval=1, init=True):
        return self._fit(data, val_data, verbose, interval, init)

    def _fit(self, data, val_data, verbose, interval, init, logger=None, es_tl=None):
        tim = self._tim
        sign = -1.0 if tim else 1.0        # the fork tracks the LOSS (-mean ll, lower is better): sgd_gmm_tim.py:176,192,210
        loader = self._batches(data, shuffle=True)
        n_local = loader.n
        n_total = _global_count(n_local, self.device)
        sizes = _global_batch_sizes(loader)
        if init:
            # sgd_gmm.py:153 initialises from the training loader; the fork from ONE batch holding the whole
            # table (sgd_gmm_tim.py:152-153)
            self.init_params(self._batches(data, shuffle=True, batch_size=n_local) if tim else loader)
        self.train_loss_curve = []
        if val_data:
            self.val_loss_curve = []
            best_val_loss, no_improvement_epochs = -math.inf, 0
        prev_loss = -math.inf
        fused = self._fused_step_ok()
        for i in range(self.epochs):
            if fused:
                train_loss = self._fused_epoch(loader, n_total, sizes)
            else:
                epoch_ll = torch.zeros((), dtype=torch.float64, device=self.device)
                for b, d in enumerate(loader):
                    self.optimiser.zero_grad(set_to_none=True)
                    n = sizes[b]                                                 # global minibatch
                    ll_sum = self.module.log_prob_sum(d)                         # sgd_gmm.py:173 (+ all-reduce)
                    loss = -ll_sum / n + self.reg_loss(n, n_total)               # :174-179
                    epoch_ll += ll_sum.detach()                                  # :176 without the per-batch sync
                    loss.backward()                                              # :181
                    self.optimiser.step()                                        # :182
                train_loss = epoch_ll.item() / n_total
            if math.isnan(train_loss):
                raise ValueError("Sigma_j + S_i is not positive definite for some (point, component) pair")
            self.train_loss_curve.append(sign * train_loss)
            if tim and es_tl is not None and sign * train_loss < es_tl:          # sgd_gmm_tim.py:187-189
                break
            if val_data:
                n_val = as_tensors(val_data)[0].shape[0]                         # len(val_data), also for (X, S) tuples
                val_loss = self.score_batch(val_data) / _global_count(n_val, self.device)
                self.val_loss_curve.append(sign * val_loss)
            if not tim:
                self.scheduler.step()                                            # disabled in the fork (:195)
            if verbose and i % interval == 0:
                if val_data:
                    msg = "Epoch {}, Train Loss: {}, Val Loss :{}".format(i, sign * train_loss, sign * val_loss)
                else:
                    msg = "Epoch {}, Loss: {}".format(i, sign * train_loss)
                if tim and logger is not None:
                    logger.info(msg)                                             # :204,207
                else:
                    print(msg)
            if val_data:
                if val_loss > best_val_loss:          # higher mean ll == lower loss of the fork (:210)
                    no_improvement_epochs, best_val_loss = 0, val_loss
                else:
                    no_improvement_epochs += 1
                if no_improvement_epochs > self.max_no_improvement:
                    print("No improvement in val loss for {} epochs. Early Stopping at {}".format(
                        self.max_no_improvement, sign * val_loss))
                    if tim:
                        return sign * best_val_loss                              # :221
                    break
            if not tim and abs(train_loss - prev_loss) < self.tol:               # commented out in the fork (:223-227)
                print("Training loss converged within tolerance at {}".format(train_loss))
                break
            prev_loss = train_loss
        return None if tim else self

    def score(self, data):
        """Per-point log-likelihood ``[B]`` (sgd_gmm.py:225-226)."""
        return self.module(data)

    def score_batch(self, dataset):
        """Sum of log-likelihoods as a float (sgd_gmm.py:228-242): one launch over the table."""
        X, S = as_tensors(dataset)
        with torch.no_grad():
            X = _resident(X, self.device)
            if S is None or not isinstance(self.module, SGDDeconvGMMModule):
                S = None
            else:
                S = _resident(S, self.device)
            log_alpha, mu, Sigma = self.module._mixture_tensors()
            _, sums = _ops.loglik(X, S, log_alpha.contiguous(), mu.contiguous(), Sigma.contiguous(),
                                  noise=_noise_layout(X, S), want_sum=True)
            if _dist.is_distributed():
                torch.distributed.all_reduce(sums)
            s = sums.tolist()
        return -math.inf if s[1] > 0 else s[0]

    def _sample(self, num_samples):
        """sgd_gmm.py:244-254: draws from the fitted (noise-free) mixture."""
        with torch.no_grad():
            idx = torch.multinomial(self.weights, num_samples, replacement=True)
            eps = torch.randn(num_samples, self.d, device=self.device)
            return self.module.means[idx] + torch.einsum("nab,nb->na", self.module.L[idx], eps)


class SGDGMM(BaseSGDGMM):
    """deconv/gmm/sgd_gmm.py:264-272."""

    def __init__(self, components, dimensions, epochs=10000, lr=1e-3, batch_size=64, tol=1e-6, w=1e-3, device=None):
        self.module = SGDGMMModule(components, dimensions, w, device)
        super().__init__(components, dimensions, epochs=epochs, lr=lr, batch_size=batch_size, w=w, tol=tol,
                         device=device)


class SGDDeconvGMM(BaseSGDGMM):
    """deconv/gmm/sgd_deconv_gmm.py:44-130."""

    def __init__(self, components, dimensions, epochs=10000, lr=1e-3, batch_size=64, tol=1e-6, w=1e-3,
                 k_means_factor=100, k_means_iters=10, lr_step=5, lr_gamma=0.1, device=None):
        self.module = SGDDeconvGMMModule(components, dimensions, w, device)
        super().__init__(components, dimensions, epochs=epochs, lr=lr, batch_size=batch_size, w=w, tol=tol,
                         k_means_factor=k_means_factor, k_means_iters=k_means_iters, lr_step=lr_step,
                         lr_gamma=lr_gamma, device=device)

    def posterior_params(self, x):
        """``(log r [N,K], b [N,K,D], B [N,K,D,D])`` (sgd_deconv_gmm.py:71-94); materialising,
        meant for small N."""
        with torch.no_grad():
            X = _resident(x[0], self.device)
            S = _resident(x[1], self.device)
            log_alpha, mu, Sigma = self.module._mixture_tensors()
            _, log_r, b, B, status = _ops.estep_full(X, S, log_alpha.contiguous(), mu.contiguous(),
                                                     Sigma.contiguous(), noise=_noise_layout(X, S))
        if int(status.item()) != 0:
            raise ValueError("Sigma_j + S_i is not positive definite for some (point, component) pair")
        return log_r, b, B

    def posterior_log_prob(self, x, context):
        """sgd_deconv_gmm.py:96-105: log density of ``x`` (``[N,D]`` -> ``[N]``... one query per context
        point gives ``[N,N]`` in the reference's broadcasting; ``[N,M,D]`` -> ``[N,M]``) under each context
        point's posterior mixture, in one kernel (``dgmm_posterior_logprob_f32``): no ``[N,K,D,D]`` tensor."""
        with torch.no_grad():
            X = _resident(context[0], self.device)
            S = _resident(context[1], self.device)
            n = X.shape[0]
            xq = _resident(x, self.device)
            if xq.dim() == 2:
                # reference: log_prob(x[:, None, None, :]) against [N,K] posteriors -> [len(x), N]
                V = xq[None, :, :].expand(n, xq.shape[0], self.d).contiguous()
            else:
                V = xq.contiguous()
            log_alpha, mu, Sigma = self.module._mixture_tensors()
            out, status = _ops.posterior_logprob(X, S, log_alpha.contiguous(), mu.contiguous(), Sigma.contiguous(), V,
                                                 noise=_noise_layout(X, S))
        if int(status.item()) != 0:
            raise ValueError("a posterior covariance B_ij (or some Sigma_j + S_i) is not positive definite")
        return out.transpose(0, 1) if xq.dim() == 2 else out

    def _sample_prior(self, num_samples, context=None):
        return self._sample(num_samples)

    def sample_prior(self, num_samples, device=torch.device("cpu")):
        """sgd_deconv_gmm.py:61-69 (chunked by ``batch_size`` like deconv/utils/sampling.py:5-25)."""
        out = torch.empty(num_samples, self.d, device=device)
        step = max(int(self.batch_size or num_samples), 1)
        for lo in range(0, num_samples, step):
            out[lo:lo + step] = self._sample(min(step, num_samples - lo)).to(device)
        return out

    def _sample_posterior(self, x, num_samples, context=None, generator=None):
        """sgd_deconv_gmm.py:107-119 -> ``[N, num_samples, D]``: a component per draw from ``r_i`` and a
        draw of ``N(b_ij, B_ij)`` for the picked component only, in one kernel
        (``dgmm_posterior_sample_f32``); no ``[N,K,D,D]`` tensor is formed."""
        with torch.no_grad():
            X = _resident(x[0], self.device)
            S = _resident(x[1], self.device)
            log_alpha, mu, Sigma = self.module._mixture_tensors()
            n = X.shape[0]
            u = torch.rand(n, num_samples, device=self.device, generator=generator)
            z = torch.randn(n, num_samples, self.d, device=self.device, generator=generator)
            out, _, _, status = _ops.posterior_sample(X, S, log_alpha.contiguous(), mu.contiguous(),
                                                      Sigma.contiguous(), u, z, noise=_noise_layout(X, S))
        if int(status.item()) != 0:
            raise ValueError("a posterior covariance B_ij (or some Sigma_j + S_i) is not positive definite")
        return out

    def sample_posterior(self, x, num_samples, device=torch.device("cpu")):
        """sgd_deconv_gmm.py:121-130."""
        with torch.no_grad():
            return self._sample_posterior(x, num_samples).to(device)


# =========================================================================================
# The `_tim` forks (deconv/gmm/sgd_gmm_tim.py, deconv/gmm/sgd_deconv_gmm_tim.py)
# =========================================================================================
class BaseSGDGMMTim(BaseSGDGMM):
    """deconv/gmm/sgd_gmm_tim.py:65-259: the same module, loss and Adam step as ``BaseSGDGMM`` with the
    fork's loop bookkeeping - ``fit(data, logger, val_data=None, verbose=False, interval=1, es_tl=None)``
    (:112), curves hold the LOSS ``-mean ll`` (:176,192), k-means initialisation from one batch holding the
    whole table (:152-153), no LR scheduler step (:195), no tolerance stop (:223-227), ``es_tl`` early stop on
    the training loss (:187-189), messages to ``logger.info`` (:204,207), ``max_no_improvement=30`` (:68) and
    ``best_val_loss`` returned on a validation early stop (:221)."""
    _tim = True

    def __init__(self, components, dimensions, epochs=10000, lr=1e-3, batch_size=64, tol=1e-6,
                 max_no_improvement=30, k_means_factor=100, w=1e-6, k_means_iters=10, lr_step=5, lr_gamma=0.1,
                 device=None):
        BaseSGDGMM.__init__(self, components, dimensions, epochs=epochs, lr=lr, batch_size=batch_size, tol=tol,
                            max_no_improvement=max_no_improvement, k_means_factor=k_means_factor, w=w,
                            k_means_iters=k_means_iters, lr_step=lr_step, lr_gamma=lr_gamma, device=device)

    def fit(self, data, logger=None, val_data=None, verbose=False, interval=1, es_tl=None, init=True):
        return self._fit(data, val_data, verbose, interval, init, logger=logger, es_tl=es_tl)

    def score(self, data):
        with torch.no_grad():                                                    # sgd_gmm_tim.py:231-233
            return self.module(data)


class SGDGMMTim(BaseSGDGMMTim):
    """deconv/gmm/sgd_gmm_tim.py:262-270."""

    def __init__(self, components, dimensions, epochs=10000, lr=1e-3, batch_size=64, tol=1e-6, w=1e-3, device=None):
        self.module = SGDGMMModule(components, dimensions, w, device)
        BaseSGDGMMTim.__init__(self, components, dimensions, epochs=epochs, lr=lr, batch_size=batch_size, w=w,
                               tol=tol, device=device)


class SGDDeconvGMMTim(BaseSGDGMMTim, SGDDeconvGMM):
    """deconv/gmm/sgd_deconv_gmm_tim.py:44-143 (``SGDDeconvGMM`` there): the posterior / sampling methods of
    ``SGDDeconvGMM`` on the fork's fit loop, plus ``log_prob_prior`` (:81-97)."""

    def __init__(self, components, dimensions, epochs=10000, lr=1e-3, batch_size=64, tol=1e-6, w=1e-3,
                 k_means_factor=100, k_means_iters=10, lr_step=5, lr_gamma=0.1, device=None):
        self.module = SGDDeconvGMMModule(components, dimensions, w, device)
        BaseSGDGMMTim.__init__(self, components, dimensions, epochs=epochs, lr=lr, batch_size=batch_size, w=w,
                               tol=tol, k_means_factor=k_means_factor, k_means_iters=k_means_iters,
                               lr_step=lr_step, lr_gamma=lr_gamma, device=device)

    def log_prob_prior(self, x, device=torch.device("cpu")):
        """Log density of ``x [N,D]`` under the fitted NOISE-FREE mixture (sgd_deconv_gmm_tim.py:81-97; the
        reference loops over components on the CPU): the ``S = 0`` layout of the log-likelihood kernel."""
        with torch.no_grad():
            X = _resident(x, self.device)
            log_alpha, mu, Sigma = self.module._mixture_tensors()
            ll = _ops.loglik(X, None, log_alpha.contiguous(), mu.contiguous(), Sigma.contiguous(), noise="none")
        return ll.to(device)

"""Tensor-level wrappers around the C ABI (device pointers of torch CUDA tensors in,
launches on torch's current stream).  torch is plumbing here: memory, streams."""
import ctypes

import torch

from . import _lib
from ._lib import NOISE_LAYOUTS, DgmmError, check

_workspaces = {}


def stats_len(D, K):
    return K * (1 + D + D * D) + 4


def _f32c(t, name, device=None):
    if not isinstance(t, torch.Tensor):
        raise TypeError("%s must be a torch.Tensor" % name)
    if not t.is_cuda:
        raise DgmmError("%s must live on a CUDA device (the B200 path has no CPU fallback)" % name)
    if t.dtype != torch.float32:
        raise TypeError("%s must be float32, got %s" % (name, t.dtype))
    if device is not None and t.device != device:
        raise DgmmError("%s is on %s, expected %s" % (name, t.device, device))
    return t if t.is_contiguous() else t.contiguous()


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else ctypes.c_void_p(0)


def _stream(device):
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def workspace(D, K, device):
    key = (D, K, device.index if device.index is not None else torch.cuda.current_device())
    ws = _workspaces.get(key)
    if ws is None:
        with torch.cuda.device(device):
            nbytes = _lib.lib().dgmm_workspace_bytes(D, K)
        if nbytes == 0:
            raise DgmmError("dgmm_workspace_bytes(D=%d, K=%d) failed (unsupported shape or no sm_100 device)" % (D, K))
        ws = torch.empty(nbytes, dtype=torch.uint8, device=device)
        _workspaces[key] = ws
    return ws


def clear_workspaces():
    """Drops the cached per-(D, K, device) workspaces and scratch vectors (they are otherwise kept for the life of
    the process so that a fit allocates nothing per iteration)."""
    _workspaces.clear()
    _ll_scratches.clear()
    _km_workspaces.clear()


_ll_scratches = {}


def _ll_scratch(N, device):
    """[N] float32 scratch for the D > 16 statistics pass (grown on demand, one per device)."""
    key = device.index if device.index is not None else torch.cuda.current_device()
    t = _ll_scratches.get(key)
    if t is None or t.numel() < N:
        t = torch.empty(max(N, 1), dtype=torch.float32, device=device)
        _ll_scratches[key] = t
    return t[:N]


def _noise(S, layout, N, D, device):
    code = NOISE_LAYOUTS[layout] if isinstance(layout, str) else int(layout)
    if code == _lib.NOISE_NONE:
        return None, code
    S = _f32c(S, "noise_covars", device)
    want = {_lib.NOISE_FULL: (N, D, D), _lib.NOISE_DIAG: (N, D), _lib.NOISE_SHARED: (D, D),
            _lib.NOISE_PACKED: (N, D * (D + 1) // 2)}[code]
    if tuple(S.shape) != want:
        raise ValueError("noise_covars has shape %s, expected %s for layout %r" % (tuple(S.shape), want, layout))
    return S, code


def _params(log_alpha, mu, Sigma, device):
    mu = _f32c(mu, "means", device)
    K, D = mu.shape
    Sigma = _f32c(Sigma, "covars", device)
    log_alpha = _f32c(log_alpha, "log_alpha", device).reshape(-1)
    if tuple(Sigma.shape) != (K, D, D) or log_alpha.numel() != K:
        raise ValueError("parameter shapes disagree: means %s covars %s log_alpha %s"
                         % (tuple(mu.shape), tuple(Sigma.shape), tuple(log_alpha.shape)))
    return log_alpha, mu, Sigma, K, D


def loglik(X, S, log_alpha, mu, Sigma, noise="full", want_sum=False):
    """``ll_i = log sum_j alpha_j N(x_i | mu_j, Sigma_j + S_i)`` -> ``[N]`` float32.
    With ``want_sum`` also returns a float64 device tensor ``[sum_i ll_i, n_nonpd]``."""
    X = _f32c(X, "X")
    dev = X.device
    N, D = X.shape
    log_alpha, mu, Sigma, K, D2 = _params(log_alpha, mu, Sigma, dev)
    if D2 != D:
        raise ValueError("X has D=%d but means have D=%d" % (D, D2))
    S, code = _noise(S, noise, N, D, dev)
    ll = torch.empty(N, dtype=torch.float32, device=dev)
    sums = torch.empty(2, dtype=torch.float64, device=dev)
    ws = workspace(D, K, dev)
    with torch.cuda.device(dev):
        rc = _lib.lib().dgmm_loglik_f32(_ptr(X), _ptr(S), code, N, D, K, _ptr(log_alpha), _ptr(mu), _ptr(Sigma),
                                        _ptr(ll), _ptr(sums), None, _ptr(ws), ws.numel(), _stream(dev))
    check(rc, "dgmm_loglik_f32")
    return (ll, sums) if want_sum else ll


def estep_stats(X, S, log_alpha, mu, Sigma, g=None, resp_floor=0.0, noise="full", want_ll=True, out_stats=None):
    """Fused E-step + sufficient statistics.  Returns ``(ll [N] or None, stats)`` with
    ``stats`` the packed float32 buffer of ``include/dgmm_b200.h``."""
    X = _f32c(X, "X")
    dev = X.device
    N, D = X.shape
    log_alpha, mu, Sigma, K, D2 = _params(log_alpha, mu, Sigma, dev)
    if D2 != D:
        raise ValueError("X has D=%d but means have D=%d" % (D, D2))
    S, code = _noise(S, noise, N, D, dev)
    if g is not None:
        g = _f32c(g, "g", dev).reshape(-1)
        if g.numel() != N:
            raise ValueError("g must have N elements")
    ll = torch.empty(N, dtype=torch.float32, device=dev) if want_ll else None
    if ll is None and D > 16 and _lib.lib().dgmm_stats_needs_ll(D, K, code):
        ll = _ll_scratch(N, dev)      # the two-pass cooperative path carries ll_i between its passes
    if out_stats is None:
        out_stats = torch.empty(stats_len(D, K), dtype=torch.float32, device=dev)
    ws = workspace(D, K, dev)
    with torch.cuda.device(dev):
        rc = _lib.lib().dgmm_estep_stats_f32(_ptr(X), _ptr(S), code, N, D, K, _ptr(log_alpha), _ptr(mu), _ptr(Sigma),
                                             _ptr(g), float(resp_floor), _ptr(ll), _ptr(out_stats), _ptr(ws),
                                             ws.numel(), _stream(dev))
    check(rc, "dgmm_estep_stats_f32")
    return (ll if want_ll else None), out_stats


def estep_full(X, S, log_alpha, mu, Sigma, noise="full", want_means=True, want_covars=True):
    """Materialising E-step: ``(ll [N], log_r [N,K], b [N,K,D], B [N,K,D,D], status)``."""
    X = _f32c(X, "X")
    dev = X.device
    N, D = X.shape
    log_alpha, mu, Sigma, K, D2 = _params(log_alpha, mu, Sigma, dev)
    if D2 != D:
        raise ValueError("X has D=%d but means have D=%d" % (D, D2))
    S, code = _noise(S, noise, N, D, dev)
    ll = torch.empty(N, dtype=torch.float32, device=dev)
    log_r = torch.empty(N, K, dtype=torch.float32, device=dev)
    b = torch.empty(N, K, D, dtype=torch.float32, device=dev) if want_means else None
    B = torch.empty(N, K, D, D, dtype=torch.float32, device=dev) if want_covars else None
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    ws = workspace(D, K, dev)
    with torch.cuda.device(dev):
        rc = _lib.lib().dgmm_estep_full_f32(_ptr(X), _ptr(S), code, N, D, K, _ptr(log_alpha), _ptr(mu), _ptr(Sigma),
                                            _ptr(log_r), _ptr(b), _ptr(B), _ptr(ll), _ptr(status), _ptr(ws),
                                            ws.numel(), _stream(dev))
    check(rc, "dgmm_estep_full_f32")
    return ll, log_r, b, B, status


def mstep_update(stats, n_total, w, mu, Sigma, alpha, log_alpha):
    """In-place closed-form EM update of ``mu, Sigma, alpha, log_alpha`` from packed stats."""
    dev = stats.device
    K, D = mu.shape
    for name, t in (("stats", stats), ("means", mu), ("covars", Sigma), ("alpha", alpha), ("log_alpha", log_alpha)):
        if not (t.is_cuda and t.dtype == torch.float32 and t.is_contiguous()):
            raise DgmmError("%s must be a contiguous float32 CUDA tensor (updated in place)" % name)
    with torch.cuda.device(dev):
        rc = _lib.lib().dgmm_mstep_update_f32(_ptr(stats), float(n_total), float(w), D, K, _ptr(mu), _ptr(Sigma),
                                              _ptr(alpha), _ptr(log_alpha), _stream(dev))
    check(rc, "dgmm_mstep_update_f32")


def unpack_stats(stats, D, K):
    """Views into the packed buffer: ``s0 [K], s1 [K,D], S2 [K,D,D], tail [4]``
    (tail = ll_hi, ll_lo, sum_g, nonpd)."""
    o = 0
    s0 = stats[o:o + K]; o += K
    s1 = stats[o:o + K * D].view(K, D); o += K * D
    S2 = stats[o:o + K * D * D].view(K, D, D); o += K * D * D
    return s0, s1, S2, stats[o:o + 4]


def set_tuning(sigp=-1, accp=-1, nslot=-1, pt=-1, nst=-1, blocks_per_sm=-1, rc=-1, no_bulk=0):
    check(_lib.lib().dgmm_debug_set_tuning(sigp, accp, nslot, pt, nst, blocks_per_sm, rc, no_bulk), "set_tuning")


def get_plan(mode, D, K, N, noise="full", has_g=False):
    out = (ctypes.c_int * 11)()
    code = NOISE_LAYOUTS[noise] if isinstance(noise, str) else int(noise)
    check(_lib.lib().dgmm_debug_get_plan(mode, D, K, N, code, int(has_g), out), "get_plan")
    keys = ("sigp", "accp", "rc", "GS", "NSLOT", "TPB", "PT", "NST", "grid", "smem", "use_bulk")
    return dict(zip(keys, list(out)))


def pair_timing(enable=True):
    """Start (or stop) recording the device time of every pair-kernel launch."""
    check(_lib.lib().dgmm_debug_pair_timing(int(bool(enable))), "pair_timing")


def pair_times(capacity=512):
    """Device times (ms) of the pair-kernel launches recorded since ``pair_timing(True)``."""
    buf = (ctypes.c_float * capacity)()
    n = ctypes.c_int(capacity)
    check(_lib.lib().dgmm_debug_pair_times(buf, ctypes.byref(n)), "pair_times")
    return [buf[i] for i in range(n.value)]


def fp32_probe(device, iters=4096, reps=5):
    """Measured FP32-FMA throughput (FLOP/s) of a register-only FFMA chain on ``device``."""
    sink = torch.zeros(1, dtype=torch.float32, device=device)
    flops = ctypes.c_double(0.0)
    best = 0.0
    with torch.cuda.device(device):
        st = _stream(device)
        check(_lib.lib().dgmm_fp32_probe(_ptr(sink), iters, ctypes.byref(flops), st), "fp32_probe")
        torch.cuda.synchronize(device)
        for _ in range(reps):
            a = torch.cuda.Event(enable_timing=True)
            b = torch.cuda.Event(enable_timing=True)
            a.record()
            check(_lib.lib().dgmm_fp32_probe(_ptr(sink), iters, ctypes.byref(flops), st), "fp32_probe")
            b.record()
            torch.cuda.synchronize(device)
            best = max(best, flops.value / (a.elapsed_time(b) * 1e-3))
    return best


class HostContext:
    """Host-buffer entry points of the C ABI (``dgmm_ctx_*``, ``dgmm_em_iteration_host``,
    ``dgmm_estep_stats_host``, ``dgmm_score_host``): data and parameters are CPU tensors
    (pinned memory makes the chunked upload overlap with compute), every call streams the table
    to the GPU and returns with the result on the host."""

    def __init__(self, D, K, chunk_points=1 << 18, device=0):
        self.D, self.K = int(D), int(K)
        self.device = int(device)
        self._ctx = ctypes.c_void_p(0)
        check(_lib.lib().dgmm_ctx_create(ctypes.byref(self._ctx), self.device, self.D, self.K, int(chunk_points)),
              "dgmm_ctx_create")

    def close(self):
        if self._ctx:
            _lib.lib().dgmm_ctx_destroy(self._ctx)
            self._ctx = ctypes.c_void_p(0)

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass

    @staticmethod
    def _host(t, name, shape=None):
        if not isinstance(t, torch.Tensor) or t.is_cuda or t.dtype != torch.float32 or not t.is_contiguous():
            raise DgmmError("%s must be a contiguous float32 CPU tensor" % name)
        if shape is not None and tuple(t.shape) != tuple(shape):
            raise ValueError("%s has shape %s, expected %s" % (name, tuple(t.shape), tuple(shape)))
        return ctypes.c_void_p(t.data_ptr())

    def _data(self, X, S, noise):
        code = NOISE_LAYOUTS[noise] if isinstance(noise, str) else int(noise)
        N = X.shape[0]
        want = {_lib.NOISE_FULL: (N, self.D, self.D), _lib.NOISE_DIAG: (N, self.D), _lib.NOISE_SHARED: (self.D, self.D),
                _lib.NOISE_PACKED: (N, self.D * (self.D + 1) // 2)}
        pX = self._host(X, "X", (N, self.D))
        pS = self._host(S, "noise_covars", want[code]) if code != _lib.NOISE_NONE else ctypes.c_void_p(0)
        return pX, pS, code, N

    def em_iteration(self, X, S, w, alpha, mu, Sigma, noise="full"):
        """One EM iteration; ``alpha [K]``, ``mu``, ``Sigma`` (CPU) are updated in place.
        Returns ``(ll_sum, n_nonpd)``."""
        pX, pS, code, N = self._data(X, S, noise)
        ll, bad = ctypes.c_double(0.0), ctypes.c_int64(0)
        check(_lib.lib().dgmm_em_iteration_host(
            self._ctx, pX, pS, code, N, float(w), self._host(alpha, "alpha", (self.K,)),
            self._host(mu, "means", (self.K, self.D)), self._host(Sigma, "covars", (self.K, self.D, self.D)),
            ctypes.byref(ll), ctypes.byref(bad)), "dgmm_em_iteration_host")
        return ll.value, bad.value

    def estep_stats(self, X, S, alpha, mu, Sigma, resp_floor=0.0, noise="full", out=None):
        pX, pS, code, N = self._data(X, S, noise)
        if out is None:
            out = torch.empty(stats_len(self.D, self.K), dtype=torch.float32)
        check(_lib.lib().dgmm_estep_stats_host(
            self._ctx, pX, pS, code, N, self._host(alpha, "alpha", (self.K,)),
            self._host(mu, "means", (self.K, self.D)), self._host(Sigma, "covars", (self.K, self.D, self.D)),
            float(resp_floor), self._host(out, "out_stats", (stats_len(self.D, self.K),))), "dgmm_estep_stats_host")
        return out

    def score(self, X, S, alpha, mu, Sigma, noise="full", want_ll=False):
        pX, pS, code, N = self._data(X, S, noise)
        ll = torch.empty(N, dtype=torch.float32) if want_ll else None
        tot, bad = ctypes.c_double(0.0), ctypes.c_int64(0)
        check(_lib.lib().dgmm_score_host(
            self._ctx, pX, pS, code, N, self._host(alpha, "alpha", (self.K,)), self._host(mu, "means", (self.K, self.D)),
            self._host(Sigma, "covars", (self.K, self.D, self.D)),
            self._host(ll, "out_ll") if want_ll else ctypes.c_void_p(0), ctypes.byref(tot), ctypes.byref(bad)),
            "dgmm_score_host")
        return (tot.value, bad.value, ll) if want_ll else (tot.value, bad.value)


    # ---- resident table: upload once, iterate many times ------------------------------------
    def upload(self, X, S, noise="full", pack=False):
        """Copies the table into device memory owned by the context (``dgmm_ctx_upload``); with ``pack`` a
        full ``[N,D,D]`` noise tensor is packed to lower triangles on the device as it streams in."""
        pX, pS, code, N = self._data(X, S, noise)
        check(_lib.lib().dgmm_ctx_upload(self._ctx, pX, pS, code, N, 1 if pack else 0), "dgmm_ctx_upload")
        self.n_resident = N
        return self

    def _params_host(self, alpha, mu, Sigma):
        return (self._host(alpha, "alpha", (self.K,)), self._host(mu, "means", (self.K, self.D)),
                self._host(Sigma, "covars", (self.K, self.D, self.D)))

    def em_iteration_resident(self, w, alpha, mu, Sigma):
        ll, bad = ctypes.c_double(0.0), ctypes.c_int64(0)
        check(_lib.lib().dgmm_em_iteration_resident(self._ctx, float(w), *self._params_host(alpha, mu, Sigma),
                                                    ctypes.byref(ll), ctypes.byref(bad)), "dgmm_em_iteration_resident")
        return ll.value, bad.value

    def em_fit_resident(self, max_iters, w, tol, alpha, mu, Sigma):
        """``base.py:46-55`` on the resident table: returns ``(iterations done, ll trace, n_nonpd)``."""
        done, bad = ctypes.c_int(0), ctypes.c_int64(0)
        trace = (ctypes.c_double * max(int(max_iters), 1))()
        check(_lib.lib().dgmm_em_fit_resident(self._ctx, int(max_iters), float(w), float(tol),
                                              *self._params_host(alpha, mu, Sigma), ctypes.byref(done), trace,
                                              ctypes.byref(bad)), "dgmm_em_fit_resident")
        n = min(int(max_iters), done.value + 1)
        return done.value, [trace[i] for i in range(n)], bad.value

    def score_resident(self, alpha, mu, Sigma, want_ll=False):
        ll = torch.empty(self.n_resident, dtype=torch.float32) if want_ll else None
        tot, bad = ctypes.c_double(0.0), ctypes.c_int64(0)
        check(_lib.lib().dgmm_score_resident(self._ctx, *self._params_host(alpha, mu, Sigma),
                                             self._host(ll, "out_ll") if want_ll else ctypes.c_void_p(0),
                                             ctypes.byref(tot), ctypes.byref(bad)), "dgmm_score_resident")
        return (tot.value, bad.value, ll) if want_ll else (tot.value, bad.value)


def pack_noise(S):
    """``S [N,D,D]`` -> packed lower triangles ``[N, D(D+1)/2]`` (row-major: (0,0),(1,0),(1,1),(2,0),...), the
    ``"packed"`` noise layout: CUDA tensors through ``dgmm_pack_noise_f32``, CPU tensors through the threaded
    ``dgmm_pack_noise_host``."""
    if S.dim() != 3 or S.shape[1] != S.shape[2]:
        raise ValueError("expected [N,D,D], got %s" % (tuple(S.shape),))
    N, D = S.shape[0], S.shape[1]
    if S.dtype != torch.float32:
        raise TypeError("noise_covars must be float32")
    S = S.contiguous()
    out = torch.empty(N, D * (D + 1) // 2, dtype=torch.float32, device=S.device)
    if S.is_cuda:
        with torch.cuda.device(S.device):
            rc = _lib.lib().dgmm_pack_noise_f32(_ptr(S), N, D, _ptr(out), _stream(S.device))
        check(rc, "dgmm_pack_noise_f32")
    else:
        check(_lib.lib().dgmm_pack_noise_host(ctypes.c_void_p(S.data_ptr()), N, D, ctypes.c_void_p(out.data_ptr())),
              "dgmm_pack_noise_host")
    return out


# -----------------------------------------------------------------------------------------
# initialisation: fused k-means assignment
# -----------------------------------------------------------------------------------------
_km_workspaces = {}


def kmeans_stats_len(D, K):
    return K * (1 + D) + 2


def _km_workspace(D, K, device):
    key = (D, K, device.index if device.index is not None else torch.cuda.current_device())
    ws = _km_workspaces.get(key)
    if ws is None:
        with torch.cuda.device(device):
            nbytes = _lib.lib().dgmm_kmeans_workspace_bytes(D, K)
        if nbytes == 0:
            raise DgmmError("dgmm_kmeans_workspace_bytes(D=%d, K=%d) failed" % (D, K))
        ws = torch.empty(nbytes, dtype=torch.uint8, device=device)
        _km_workspaces[key] = ws
    return ws


def kmeans_assign(X, centroids, want_labels=False, out_stats=None):
    """Nearest-centroid labels and, in the same pass, the packed ``[counts K | sums K*D | dist_hi |
    dist_lo]`` buffer (deconv/gmm/util.py:19-24,51-64).  Returns ``(labels int32 [N] or None, stats)``."""
    X = _f32c(X, "X")
    dev = X.device
    centroids = _f32c(centroids, "centroids", dev)
    N, D = X.shape
    K = centroids.shape[0]
    if centroids.shape[1] != D:
        raise ValueError("centroids %s do not match X %s" % (tuple(centroids.shape), tuple(X.shape)))
    labels = torch.empty(N, dtype=torch.int32, device=dev) if want_labels else None
    stats = out_stats if out_stats is not None else torch.empty(kmeans_stats_len(D, K), dtype=torch.float32, device=dev)
    ws = _km_workspace(D, K, dev)
    with torch.cuda.device(dev):
        rc = _lib.lib().dgmm_kmeans_assign_f32(_ptr(X), N, D, K, _ptr(centroids), _ptr(labels), _ptr(stats), _ptr(ws),
                                               ws.numel(), _stream(dev))
    check(rc, "dgmm_kmeans_assign_f32")
    return labels, stats


def kmeans_update(mode, stats, centroids, counts_run=None):
    """In-place centroid update from the (all-reduced) packed buffer: ``mode='lloyd'`` or
    ``'minibatch'`` (the latter also advances ``counts_run [K]``)."""
    dev = stats.device
    K, D = centroids.shape
    for name, t in (("stats", stats), ("centroids", centroids), ("counts_run", counts_run)):
        if t is not None and not (t.is_cuda and t.dtype == torch.float32 and t.is_contiguous()):
            raise DgmmError("%s must be a contiguous float32 CUDA tensor (updated in place)" % name)
    code = {"lloyd": 0, "minibatch": 1}[mode]
    with torch.cuda.device(dev):
        rc = _lib.lib().dgmm_kmeans_update_f32(code, _ptr(stats), D, K, _ptr(centroids), _ptr(counts_run), _stream(dev))
    check(rc, "dgmm_kmeans_update_f32")


def kmeans_total_distance(stats):
    """``sum_i min_j |x_i - c_j|`` from the packed buffer (a float pair carrying a double)."""
    return float(stats[-2].item()) + float(stats[-1].item())


# -----------------------------------------------------------------------------------------
# minibatch SGD: parameter side of a step (deconv/gmm/sgd_gmm.py:167-182)
# -----------------------------------------------------------------------------------------
def _inplace(name, t):
    if t is not None and not (t.is_cuda and t.dtype == torch.float32 and t.is_contiguous()):
        raise DgmmError("%s must be a contiguous float32 CUDA tensor (updated in place)" % name)
    return t


def sgd_params(soft_weights, l_diag, l_lower, out_log_alpha=None, out_Sigma=None):
    """``log_softmax(soft_weights)`` and ``Sigma_j = L_j L_j^T`` (sgd_gmm.py:38-47)."""
    dev = soft_weights.device
    K, D = l_diag.shape
    for n, t in (("soft_weights", soft_weights), ("l_diag", l_diag), ("l_lower", l_lower)):
        _inplace(n, t)
    la = out_log_alpha if out_log_alpha is not None else torch.empty(K, dtype=torch.float32, device=dev)
    Sg = out_Sigma if out_Sigma is not None else torch.empty(K, D, D, dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        rc = _lib.lib().dgmm_sgd_params_f32(_ptr(soft_weights), _ptr(l_diag), _ptr(l_lower if D > 1 else None), D, K,
                                            _ptr(la), _ptr(Sg), _stream(dev))
    check(rc, "dgmm_sgd_params_f32")
    return la, Sg


def sgd_adam_step(stats, n_batch, n_total, w, lr, betas, eps, step, params, exp_avg, exp_avg_sq, log_alpha, Sigma,
                  epoch_acc=None):
    """Chain rule + Adam + re-parametrisation from the packed statistics of one minibatch, in place.
    ``params = (soft_weights, means, l_diag, l_lower)``; ``exp_avg`` / ``exp_avg_sq`` the matching Adam
    moments; ``epoch_acc`` an optional float64 ``[2]`` accumulator of (sum ll, non-PD pairs)."""
    dev = stats.device
    sw, mu, ld, ll = params
    K, D = mu.shape
    for n, t in (("stats", stats), ("soft_weights", sw), ("means", mu), ("l_diag", ld), ("l_lower", ll),
                 ("log_alpha", log_alpha), ("Sigma", Sigma)) + tuple(("exp_avg", t) for t in exp_avg) + \
            tuple(("exp_avg_sq", t) for t in exp_avg_sq):
        _inplace(n, t)
    if epoch_acc is not None and not (epoch_acc.is_cuda and epoch_acc.dtype == torch.float64 and epoch_acc.numel() == 2):
        raise DgmmError("epoch_acc must be a float64 CUDA tensor of 2 elements")
    arr = ctypes.c_void_p * 4
    m = arr(*[t.data_ptr() if t.numel() else None for t in exp_avg])
    v = arr(*[t.data_ptr() if t.numel() else None for t in exp_avg_sq])
    with torch.cuda.device(dev):
        rc = _lib.lib().dgmm_sgd_adam_step_f32(_ptr(stats), float(n_batch), float(n_total), float(w), float(lr),
                                               float(betas[0]), float(betas[1]), float(eps), int(step), D, K,
                                               _ptr(sw), _ptr(mu), _ptr(ld), _ptr(ll if D > 1 else None), m, v,
                                               _ptr(log_alpha), _ptr(Sigma), _ptr(epoch_acc), _stream(dev))
    check(rc, "dgmm_sgd_adam_step_f32")


class SgdStepper:
    """One minibatch of ``BaseSGDGMM.fit`` per ctypes call (``dgmm_sgd_fused_step_f32``): gather of the shuffled
    batch from the resident table + fused statistics pass + Adam step.  Everything that does not change between
    steps (pointers, shapes, workspace) is validated and marshalled ONCE here; ``step`` only passes the batch
    offset, the step count and the learning rate.  At D = 10, K = 64 the host side of a step drops from three
    wrapper calls with their argument checks (~0.1 ms) to one call."""

    def __init__(self, X, S, noise, batch_size, params, exp_avg, exp_avg_sq, log_alpha, Sigma, epoch_acc, w, betas, eps):
        X = _f32c(X, "X")
        dev = X.device
        N, D = X.shape
        sw, mu, ld, ll = params
        K = mu.shape[0]
        S, code = _noise(S, noise, N, D, dev)
        for n, t in (("soft_weights", sw), ("means", mu), ("l_diag", ld), ("l_lower", ll), ("log_alpha", log_alpha),
                     ("Sigma", Sigma)) + tuple(("exp_avg", t) for t in exp_avg) + \
                tuple(("exp_avg_sq", t) for t in exp_avg_sq):
            _inplace(n, t)
        if not (epoch_acc.is_cuda and epoch_acc.dtype == torch.float64 and epoch_acc.numel() == 2):
            raise DgmmError("epoch_acc must be a float64 CUDA tensor of 2 elements")
        self.dev, self.N, self.D, self.K, self.code = dev, N, D, K, code
        B = int(batch_size)
        per_point = code in (_lib.NOISE_FULL, _lib.NOISE_DIAG, _lib.NOISE_PACKED)
        self.bufX = torch.empty(B, D, dtype=torch.float32, device=dev)
        self.bufS = torch.empty((B,) + tuple(S.shape[1:]), dtype=torch.float32, device=dev) if per_point else None
        self.ll = torch.empty(B, dtype=torch.float32, device=dev)
        self.stats = torch.empty(stats_len(D, K), dtype=torch.float32, device=dev)
        self.ws = workspace(D, K, dev)
        arr = ctypes.c_void_p * 4
        self._m = arr(*[t.data_ptr() if t.numel() else None for t in exp_avg])
        self._v = arr(*[t.data_ptr() if t.numel() else None for t in exp_avg_sq])
        self._keep = (X, S, params, exp_avg, exp_avg_sq, log_alpha, Sigma, epoch_acc)
        self._head = (_ptr(X), _ptr(S), code, D, K)
        self._bufs = (_ptr(self.bufX), _ptr(self.bufS), _ptr(self.ll))
        self._tail = (_ptr(sw), _ptr(mu), _ptr(ld), _ptr(ll if D > 1 else None), self._m, self._v, _ptr(log_alpha),
                      _ptr(Sigma), _ptr(epoch_acc), _ptr(self.ws), self.ws.numel())
        self._stats_p = _ptr(self.stats)
        self._hyper = (float(w), float(betas[0]), float(betas[1]), float(eps))
        self._fn = _lib.lib().dgmm_sgd_fused_step_f32
        self._stream = _stream(dev)

    def step(self, perm, batch_lo, B, phase, n_batch, n_total, lr, step, stats_ptr=None):
        """``perm``: int64 CUDA permutation of the table rows (None: contiguous batch).  ``stats_ptr``: device address
        the packed statistics go to (default: ``self.stats``)."""
        w, b1, b2, eps = self._hyper
        sp = self._stats_p if stats_ptr is None else ctypes.c_void_p(int(stats_ptr))
        with torch.cuda.device(self.dev):
            rc = self._fn(*self._head, ctypes.c_void_p(perm.data_ptr()) if perm is not None else None, int(batch_lo),
                          int(B), *self._bufs, int(phase), sp, float(n_batch), float(n_total), w, float(lr),
                          b1, b2, eps, int(step), *self._tail, _stream(self.dev))
        check(rc, "dgmm_sgd_fused_step_f32")

    def step_peers(self, perm, batch_lo, B, n_batch, n_total, lr, step, peers):
        """The distributed step WITHOUT a collective call: statistics into this rank's symmetric buffer, then the Adam
        kernels add the ``world`` ranks' statistics themselves over NVLink peer memory (``PeerReduce``)."""
        stamp = peers.next_stamp()          # the buffer's own step counter (every rank advances it in lockstep)
        off = peers.stats_offset(stamp)
        self.step(perm, batch_lo, B, 1, n_batch, n_total, lr, step, stats_ptr=peers.local_ptr + 4 * off)
        w, b1, b2, eps = self._hyper
        t = self._tail
        with torch.cuda.device(self.dev):
            rc = _lib.lib().dgmm_sgd_adam_step_peers_f32(
                peers.ptr_array, peers.world, peers.rank, off, peers.flags_off, stamp & 0xffffffff,
                _ptr(peers.timeout), float(n_batch), float(n_total), w, float(lr), b1, b2, eps, int(step), self.D, self.K,
                t[0], t[1], t[2], t[3], t[4], t[5], t[6], t[7], t[8], _stream(self.dev))
        check(rc, "dgmm_sgd_adam_step_peers_f32")


class PeerReduce:
    """Symmetric (peer-mapped) buffer for the fused cross-GPU reduction of the SGD step
    (``dgmm_sgd_adam_step_peers_f32``): two packed-statistics regions (alternating with the step parity) and the
    hand-shake flags, allocated and exchanged through ``torch.distributed._symmetric_memory`` (plumbing: the
    allocation and the handle exchange; every access to it is made by this library's kernels).  Collective: every
    rank of ``group`` must construct it at the same point.  Raises when the box has no peer access."""

    def __init__(self, D, K, device, group=None):
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm_mem
        group = group if group is not None else dist.group.WORLD
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        if self.world > 8:
            raise DgmmError("peer reduction is for one NVLink box (world <= 8)")
        L = stats_len(D, K)
        self.region = (L + 63) // 64 * 64
        self.flags_off = 2 * self.region
        self.buf = symm_mem.empty(2 * self.region + 64, dtype=torch.float32, device=device)
        self.buf.zero_()
        torch.cuda.synchronize(device)
        self.hdl = symm_mem.rendezvous(self.buf, group)
        ptrs = [int(p) for p in self.hdl.buffer_ptrs]
        if len(ptrs) != self.world or any(p == 0 for p in ptrs):
            raise DgmmError("symmetric memory rendezvous returned no peer pointers")
        self.ptr_array = (ctypes.c_void_p * 8)(*(ptrs + [None] * (8 - self.world)))
        self.local_ptr = ptrs[self.rank]
        self.timeout = torch.zeros(1, dtype=torch.int32, device=device)
        self.stamp = 0
        dist.barrier(group)

    def next_stamp(self):
        self.stamp += 1
        return self.stamp

    def stats_offset(self, stamp):
        return (int(stamp) & 1) * self.region


# -----------------------------------------------------------------------------------------
# posterior sampling (deconv/gmm/sgd_deconv_gmm.py:107-119) without [N,K,D,D] tensors
# -----------------------------------------------------------------------------------------
def posterior_sample(X, S, log_alpha, mu, Sigma, u, z, noise="full", want_comp=False):
    """``u [N,M]`` uniforms pick the component (inverse CDF of ``r_i``), ``z [N,M,D]`` standard normals
    give ``b_ij + chol(B_ij) z``.  Returns ``(samples [N,M,D], comp int32 [N,M] or None, log_r [N,K],
    status)``."""
    X = _f32c(X, "X")
    dev = X.device
    N, D = X.shape
    log_alpha, mu, Sigma, K, D2 = _params(log_alpha, mu, Sigma, dev)
    if D2 != D:
        raise ValueError("X has D=%d, parameters have D=%d" % (D, D2))
    S, code = _noise(S, noise, N, D, dev)
    u = _f32c(u, "u", dev)
    z = _f32c(z, "z", dev)
    if u.dim() != 2 or u.shape[0] != N or tuple(z.shape) != (N, u.shape[1], D):
        raise ValueError("u must be [N,M] and z [N,M,D]; got %s and %s" % (tuple(u.shape), tuple(z.shape)))
    M = u.shape[1]
    out = torch.empty(N, M, D, dtype=torch.float32, device=dev)
    comp = torch.empty(N, M, dtype=torch.int32, device=dev) if want_comp else None
    log_r = torch.empty(N, K, dtype=torch.float32, device=dev)
    ll = torch.empty(N, dtype=torch.float32, device=dev)
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    ws = workspace(D, K, dev)
    with torch.cuda.device(dev):
        rc = _lib.lib().dgmm_posterior_sample_f32(_ptr(X), _ptr(S), code, N, D, K, _ptr(log_alpha), _ptr(mu),
                                                  _ptr(Sigma), M, _ptr(u), _ptr(z), _ptr(out), _ptr(comp),
                                                  _ptr(log_r), _ptr(ll), _ptr(status), _ptr(ws), ws.numel(),
                                                  _stream(dev))
    check(rc, "dgmm_posterior_sample_f32")
    return out, comp, log_r, status


def online_mstep(stats, resp_floor, step_size, w, global_batch, mu, Sigma, sum_resps, sum_cond_means, alpha, log_alpha,
                 crash_flag=None):
    """In-place Online-EM M-step of one minibatch (online_deconv_gmm.py:134-188) from packed statistics.
    ``sum_resps [K]`` (or ``[K,1]``), ``sum_cond_means [K,D]``, ``alpha [K]`` / ``log_alpha [K]`` are updated
    with ``mu`` and ``Sigma``; ``crash_flag`` is an optional int32 ``[1]`` device tensor (sticky)."""
    dev = stats.device
    K, D = mu.shape
    for n, t in (("stats", stats), ("means", mu), ("covars", Sigma), ("sum_resps", sum_resps),
                 ("sum_cond_means", sum_cond_means), ("alpha", alpha), ("log_alpha", log_alpha)):
        _inplace(n, t)
    if crash_flag is not None and not (crash_flag.is_cuda and crash_flag.dtype == torch.int32):
        raise DgmmError("crash_flag must be an int32 CUDA tensor")
    with torch.cuda.device(dev):
        rc = _lib.lib().dgmm_online_mstep_f32(_ptr(stats), float(resp_floor), float(step_size), float(w),
                                              float(global_batch), D, K, _ptr(mu), _ptr(Sigma), _ptr(sum_resps),
                                              _ptr(sum_cond_means), _ptr(alpha), _ptr(log_alpha), _ptr(crash_flag),
                                              _stream(dev))
    check(rc, "dgmm_online_mstep_f32")


def posterior_logprob(X, S, log_alpha, mu, Sigma, V, noise="full"):
    """``out[i,m] = logsumexp_j(log r_ij + log N(V[i,m] | b_ij, B_ij))`` (sgd_deconv_gmm.py:96-105) for
    ``V [N,M,D]``.  Returns ``(logprob [N,M], status)``."""
    X = _f32c(X, "X")
    dev = X.device
    N, D = X.shape
    log_alpha, mu, Sigma, K, D2 = _params(log_alpha, mu, Sigma, dev)
    if D2 != D:
        raise ValueError("X has D=%d, parameters have D=%d" % (D, D2))
    S, code = _noise(S, noise, N, D, dev)
    V = _f32c(V, "V", dev)
    if V.dim() != 3 or V.shape[0] != N or V.shape[2] != D:
        raise ValueError("V must be [N,M,D]; got %s" % (tuple(V.shape),))
    M = V.shape[1]
    out = torch.empty(N, M, dtype=torch.float32, device=dev)
    log_r = torch.empty(N, K, dtype=torch.float32, device=dev)
    ll = torch.empty(N, dtype=torch.float32, device=dev)
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    ws = workspace(D, K, dev)
    with torch.cuda.device(dev):
        rc = _lib.lib().dgmm_posterior_logprob_f32(_ptr(X), _ptr(S), code, N, D, K, _ptr(log_alpha), _ptr(mu),
                                                   _ptr(Sigma), M, _ptr(V), _ptr(out), _ptr(log_r), _ptr(ll),
                                                   _ptr(status), _ptr(ws), ws.numel(), _stream(dev))
    check(rc, "dgmm_posterior_logprob_f32")
    return out, status


def gather_points(X, S, idx, outX=None, outS=None):
    """Rows ``idx`` (int64, on the device) of ``X [N,D]`` and of the noise table ``S`` (``[N,D,D]``, ``[N,D]``
    or None) as contiguous batch tensors, in one launch (the device-resident DataLoader)."""
    X = _f32c(X, "X")
    dev = X.device
    N, D = X.shape
    if idx.dtype != torch.int64 or not idx.is_cuda or not idx.is_contiguous():
        raise DgmmError("idx must be a contiguous int64 CUDA tensor")
    B = idx.numel()
    if S is None:
        code = _lib.NOISE_NONE
    else:
        S = _f32c(S, "noise_covars", dev)
        code = _lib.NOISE_FULL if S.dim() == 3 else (_lib.NOISE_DIAG if S.shape[1] == D else _lib.NOISE_PACKED)
        if S.shape[0] != N or tuple(S.shape[1:]) not in ((D, D), (D,), (D * (D + 1) // 2,)):
            raise ValueError("noise_covars has shape %s for X %s" % (tuple(S.shape), tuple(X.shape)))
    if outX is None:
        outX = torch.empty(B, D, dtype=torch.float32, device=dev)
    if outS is None and S is not None:
        outS = torch.empty((B,) + tuple(S.shape[1:]), dtype=torch.float32, device=dev)
    with torch.cuda.device(dev):
        rc = _lib.lib().dgmm_gather_points_f32(_ptr(X), _ptr(S), code, _ptr(idx), B, D, _ptr(outX), _ptr(outS),
                                               _stream(dev))
    check(rc, "dgmm_gather_points_f32")
    return outX, outS

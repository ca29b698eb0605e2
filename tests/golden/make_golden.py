"""Generate golden vectors by importing the UNMODIFIED reference.

Run in the build container only (needs /root/reference; the GPU box has none):

    python tests/golden/make_golden.py

Writes ``tests/golden/ref_*.npz``.  Everything the reference computes here is
driven with injected parameters (never through its k-means RNG).
The files are small and committed; tests read them and never the reference.
"""
import os
import sys
import warnings

import numpy as np
import torch

REF = os.environ.get("DGMM_REFERENCE", "/root/reference")
sys.path.insert(0, REF)
warnings.filterwarnings("ignore")

from deconv.gmm.deconv_gmm import DeconvGMM  # noqa: E402
from deconv.gmm.online_deconv_gmm import OnlineDeconvGMM  # noqa: E402
from deconv.gmm.batch_deconv_gmm import BatchDeconvGMM  # noqa: E402
from deconv.gmm.sgd_deconv_gmm import SGDDeconvGMM  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def synth(rng, N, D, K, noise_scale=0.5, spread=3.0):
    """Synthetic mixture + per-point full noise in the style of
    experiments/gmm/checks/data.py:5-10 (smaller spread so that
    responsibilities are not all one-hot)."""
    mu = rng.uniform(-spread, spread, size=(K, D))
    q = rng.standard_normal((K, D, D))
    Sigma = np.einsum("kba,kbc->kac", q, q) + 0.1 * np.eye(D)
    alpha = rng.dirichlet(np.ones(K) * 3.0)
    qn = noise_scale * rng.standard_normal((N, D, D))
    S = np.einsum("nba,nbc->nac", qn, qn) + 1e-3 * np.eye(D)
    comp = rng.integers(0, K, size=N)
    X = np.empty((N, D))
    for i in range(N):
        X[i] = rng.multivariate_normal(mu[comp[i]], Sigma[comp[i]] + S[i])
    return X, S, alpha, mu, Sigma


def inject(model, alpha, mu, Sigma, dtype):
    model.weights = torch.tensor(alpha, dtype=dtype)[:, None]
    model.means = torch.tensor(mu, dtype=dtype)
    model.covars = torch.tensor(Sigma, dtype=dtype)
    model.w = model.w.to(dtype)


def em_case(name, rng, N, D, K, dtype, w=1e-3, iters=(1, 3)):
    X, S, alpha, mu, Sigma = synth(rng, N, D, K)
    torch.set_default_dtype(dtype)
    m = DeconvGMM(K, D, w=w)
    inject(m, alpha, mu, Sigma, dtype)
    data = (torch.tensor(X, dtype=dtype), torch.tensor(S, dtype=dtype))
    ll, (log_r, b, B) = m._e_step(data)
    out = dict(X=X, S=S, alpha=alpha, mu=mu, Sigma=Sigma, w=w,
               ll_sum=ll.item(), log_resps=log_r.numpy(), cond_means=b.numpy(),
               cond_covars=B.numpy())
    m._m_step(data, (log_r, b, B))
    out.update(m1_weights=m.weights.numpy().copy(), m1_means=m.means.numpy().copy(),
               m1_covars=m.covars.numpy().copy())
    # further full EM iterations (E then M), recording params after each count
    done = 1
    for target in iters:
        while done < target:
            _, e = m._e_step(data)
            m._m_step(data, e)
            done += 1
        out["it%d_weights" % target] = m.weights.numpy().copy()
        out["it%d_means" % target] = m.means.numpy().copy()
        out["it%d_covars" % target] = m.covars.numpy().copy()
        out["it%d_ll" % target] = m._e_step(data)[0].item()
    torch.set_default_dtype(torch.float32)
    np.savez_compressed(os.path.join(HERE, name), **out)
    print(name, "ll_sum", out["ll_sum"])


def sgd_case(name, rng, N, D, K, dtype):
    X, S, _, mu, _ = synth(rng, N, D, K)
    torch.set_default_dtype(dtype)
    m = SGDDeconvGMM(K, D, w=1e-3, batch_size=16)
    sw = rng.standard_normal(K)
    l_diag = 0.3 * rng.standard_normal((K, D))
    l_lower = 0.5 * rng.standard_normal((K, D * (D - 1) // 2))
    m.module.soft_weights.data = torch.tensor(sw, dtype=dtype)
    m.module.means.data = torch.tensor(mu, dtype=dtype)
    m.module.l_diag.data = torch.tensor(l_diag, dtype=dtype)
    m.module.l_lower.data = torch.tensor(l_lower, dtype=dtype)
    data = (torch.tensor(X, dtype=dtype), torch.tensor(S, dtype=dtype))
    g = rng.standard_normal(N)
    ll = m.module(data)
    (ll * torch.tensor(g, dtype=dtype)).sum().backward()
    out = dict(X=X, S=S, soft_weights=sw, means=mu, l_diag=l_diag, l_lower=l_lower, g=g,
               ll=ll.detach().numpy(),
               d_soft_weights=m.module.soft_weights.grad.numpy().copy(),
               d_means=m.module.means.grad.numpy().copy(),
               d_l_diag=m.module.l_diag.grad.numpy().copy(),
               d_l_lower=m.module.l_lower.grad.numpy().copy(),
               covars=m.covars.numpy().copy(),
               reg_loss=m.reg_loss(16, 64).item())
    with torch.no_grad():
        pw, pm, pc = m.posterior_params(data)
    out.update(post_log_w=pw.numpy(), post_means=pm.numpy(), post_covars=pc.numpy())
    # the training objective of sgd_gmm.py:173-181 on this one batch
    for p in m.module.parameters():
        p.grad = None
    lp = m.module(data)
    loss = -1 * torch.mean(lp) + m.reg_loss(N, 4 * N)
    loss.backward()
    out.update(loss=loss.item(),
               loss_d_soft_weights=m.module.soft_weights.grad.numpy().copy(),
               loss_d_means=m.module.means.grad.numpy().copy(),
               loss_d_l_diag=m.module.l_diag.grad.numpy().copy(),
               loss_d_l_lower=m.module.l_lower.grad.numpy().copy())
    torch.set_default_dtype(torch.float32)
    np.savez_compressed(os.path.join(HERE, name), **out)
    print(name, "ll mean", out["ll"].mean())


def online_batch_case(name, rng, N, D, K, dtype):
    X, S, alpha, mu, Sigma = synth(rng, N, D, K)
    torch.set_default_dtype(dtype)
    w = 1e-3
    on = OnlineDeconvGMM(K, D, w=w, batch_size=N, step_size=0.1)
    inject(on, alpha, mu, Sigma, dtype)
    on.sum_resps = on.weights * on.batch_size
    on.sum_cond_means = on.means * on.sum_resps
    data = (torch.tensor(X, dtype=dtype), torch.tensor(S, dtype=dtype))
    ll, e = on._e_step(data)
    on._m_step(e, N, 0.1)
    out = dict(X=X, S=S, alpha=alpha, mu=mu, Sigma=Sigma, w=w, step_size=0.1,
               eps=float(torch.finfo(dtype).eps),
               on_weights=on.weights.numpy().copy(), on_means=on.means.numpy().copy(),
               on_covars=on.covars.numpy().copy(), on_sum_resps=on.sum_resps.numpy().copy(),
               on_sum_cond_means=on.sum_cond_means.numpy().copy(), ll_sum=ll.item())
    # second online step on the same batch from the updated state
    ll2, e2 = on._e_step(data)
    on._m_step(e2, N, 0.1)
    out.update(on2_weights=on.weights.numpy().copy(), on2_means=on.means.numpy().copy(),
               on2_covars=on.covars.numpy().copy(), ll2_sum=ll2.item())

    bt = BatchDeconvGMM(K, D, w=w, batch_size=N)
    inject(bt, alpha, mu, Sigma, dtype)
    half = N // 2
    parts = [(data[0][:half], data[1][:half]), (data[0][half:], data[1][half:])]
    sr = torch.zeros(K, 1)
    sm = torch.zeros(K, D)
    tl = 0.0
    for p in parts:
        lp, (a, b) = bt._step_1(p)
        sr += a
        sm += b
        tl += lp.item()
    new_means = sm / sr
    sc = torch.zeros(K, D, D)
    for p in parts:
        sc += bt._step_2(p, new_means)
    out.update(bt_ll=tl, bt_sum_resps=sr.numpy().copy(), bt_sum_cond_means=sm.numpy().copy(),
               bt_sum_cond_covars=sc.numpy().copy(),
               bt_weights=(sr / N).numpy().copy(), bt_means=new_means.numpy().copy(),
               bt_covars=((sc + 2 * bt.w) / sr[:, :, None]).numpy().copy())
    torch.set_default_dtype(torch.float32)
    np.savez_compressed(os.path.join(HERE, name), **out)
    print(name, "online ll", out["ll_sum"], "batch ll", tl)


def data_gen_case(name):
    """deconv/utils/data_gen.py fixture (seeds :42,:63): mean ll under the
    generating mixture, plus a 512-point slice with per-point values."""
    from deconv.utils.data_gen import generate_mixture_data
    from deconv.gmm.sgd_deconv_gmm import SGDDeconvDataset
    torch.set_default_dtype(torch.float32)
    gmm, S, (z_tr, x_tr), (z_va, x_va), (z_te, x_te) = generate_mixture_data()
    x_tr, x_va, x_te = x_tr[0], x_va[0], x_te[0]   # minibatch_sample adds a leading 1
    out = {}
    for tag, x in (("train", x_tr), ("val", x_va), ("test", x_te)):
        ds = SGDDeconvDataset(x, S.repeat(x.shape[0], 1, 1))
        out["mean_ll_" + tag] = gmm.score_batch(ds) / x.shape[0]
        out["n_" + tag] = x.shape[0]
    sl = slice(0, 512)
    Sx = S.repeat(512, 1, 1)
    with torch.no_grad():
        ll = gmm.module((x_te[sl], Sx))
        pw, pm, pc = gmm.posterior_params((x_te[sl], Sx))
    out.update(x_test=x_te[sl].numpy(), S=S.numpy(), ll_test=ll.numpy(),
               soft_weights=gmm.module.soft_weights.detach().numpy(),
               means=gmm.module.means.detach().numpy(),
               l_diag=gmm.module.l_diag.detach().numpy(),
               l_lower=gmm.module.l_lower.detach().numpy(),
               post_log_w=pw.numpy(), post_means=pm.numpy(), post_covars=pc.numpy())
    np.savez_compressed(os.path.join(HERE, name), **out)
    print(name, {k: v for k, v in out.items() if k.startswith("mean_ll")})


def robustness_case(name):
    """huge / tiny noise stays finite in the fp32 reference."""
    torch.set_default_dtype(torch.float32)
    K, D = 3, 2
    alpha = np.full(K, 1.0 / 3)
    mu = np.array([[0.0, 0.0], [2.0, 3.0], [2.0, -3.0]])
    Sigma = np.array([np.diag([0.1, 1.5]), np.diag([1.0, 0.1]), np.diag([1.0, 0.1])])
    X = np.array([[0.5, 0.25], [2.0, 2.5], [-1.0, 4.0], [0.1, -0.2]])
    S = np.array([np.diag([1e12, 1e-2]), 1e-4 * np.eye(2), np.diag([1e-2, 1e12]),
                  np.array([[2.0, -0.5], [-0.5, 1.0]])])
    m = DeconvGMM(K, D)
    inject(m, alpha, mu, Sigma, torch.float32)
    ll, (lr, b, B) = m._e_step((torch.tensor(X, dtype=torch.float32), torch.tensor(S, dtype=torch.float32)))
    np.savez_compressed(os.path.join(HERE, name), X=X, S=S, alpha=alpha, mu=mu, Sigma=Sigma,
                        ll_sum=ll.item(), log_resps=lr.numpy(), cond_means=b.numpy(),
                        cond_covars=B.numpy())
    print(name, ll.item(), np.isfinite(B.numpy()).all())


def boston_case(name):
    """BASELINE.json configs[0]: boston (506x11), K=10, synthetic per-point
    diagonal noise, injected init, a few full EM iterations (float64)."""
    dtype = torch.float64
    torch.set_default_dtype(dtype)
    rng = np.random.default_rng(20260101)
    X0 = np.load(os.path.join(REF, "data_small", "boston_no_discrete.npy"))
    N, D = X0.shape
    K = 10
    var = 0.1 * (0.5 + rng.uniform(size=(N, D)))
    X = X0 + np.sqrt(var) * rng.standard_normal((N, D))
    S = np.zeros((N, D, D))
    S[:, np.arange(D), np.arange(D)] = var
    rows = rng.choice(N, size=K, replace=False)
    alpha = np.full(K, 1.0 / K)
    mu = X[rows].copy()
    Sigma = np.tile(np.eye(D), (K, 1, 1))
    m = DeconvGMM(K, D, w=1e-3)
    inject(m, alpha, mu, Sigma, dtype)
    data = (torch.tensor(X), torch.tensor(S))
    out = dict(noise_var=var, X=X, rows=rows, w=1e-3)
    lls = []
    for it in range(1, 11):
        ll, e = m._e_step(data)
        lls.append(ll.item())
        m._m_step(data, e)
        if it in (1, 5, 10):
            out["it%d_weights" % it] = m.weights.numpy().copy()
            out["it%d_means" % it] = m.means.numpy().copy()
            out["it%d_covars" % it] = m.covars.numpy().copy()
    out["ll_trace"] = np.asarray(lls)
    torch.set_default_dtype(torch.float32)
    np.savez_compressed(os.path.join(HERE, name), **out)
    print(name, lls[0], lls[-1])


def kmeans_case(name):
    """deconv/gmm/util.py: ``minibatch_k_means`` over a list-of-batches loader and ``k_means`` (its
    ``torch.rand`` initial centroids are re-drawn here from the same seed), float32 and float64."""
    import contextlib
    import io
    from deconv.gmm.util import k_means, minibatch_k_means
    rng = np.random.default_rng(777)
    out = {}
    for tag, dtype, N, D, K in (("f64", torch.float64, 240, 3, 4), ("f32", torch.float32, 300, 5, 6)):
        centres = rng.normal(size=(K, D)) * 4.0
        X = torch.tensor(centres[rng.integers(0, K, N)] + rng.normal(size=(N, D)), dtype=dtype)
        old = torch.get_default_dtype()
        torch.set_default_dtype(dtype)

        class Loader:     # fresh batch tensors per pass, like a DataLoader (the reference's centroids
            def __iter__(self):  # are a view of the first batch and are updated in place)
                return iter([(X[lo:lo + 60].clone(),) for lo in range(0, N, 60)])
        batches = Loader()
        for iters in (1, 3):
            with contextlib.redirect_stdout(io.StringIO()):
                counts, cents = minibatch_k_means(batches, K, max_iters=iters, tol=0.0)
            out["mb_%s_counts_%d" % (tag, iters)] = counts.numpy().copy()
            out["mb_%s_centroids_%d" % (tag, iters)] = cents.numpy().copy()
        for seed in range(1000):          # the reference turns an empty cluster into NaN: find a seed without one
            torch.manual_seed(seed)
            with contextlib.redirect_stdout(io.StringIO()):
                if not torch.isnan(k_means(X, K, max_iters=4, tol=0.0)[1]).any():
                    break
        else:
            raise RuntimeError("no seed without an empty cluster")
        torch.manual_seed(seed)
        init = torch.rand(K, D) * (X.max(dim=0)[0] - X.min(dim=0)[0]) + X.min(dim=0)[0]
        for iters in (1, 4):
            torch.manual_seed(seed)
            resp, cents = k_means(X, K, max_iters=iters, tol=0.0)
            assert not torch.isnan(cents).any()
            out["ll_%s_resp_%d" % (tag, iters)] = resp.numpy().copy()
            out["ll_%s_centroids_%d" % (tag, iters)] = cents.numpy().copy()
        out["ll_%s_init" % tag] = init.numpy().copy()
        out["X_%s" % tag] = X.numpy().copy()
        torch.set_default_dtype(old)
    np.savez_compressed(os.path.join(HERE, name), **out)
    print(name, sorted(out))


def sgd_fit_case(name, N, D, K, dtype, epochs=7, batch=16):
    """The inner loop of BaseSGDGMM.fit (deconv/gmm/sgd_gmm.py:163-186) on the reference's own module,
    reg_loss, Adam optimiser and MultiStepLR scheduler, with injected parameters and the batches taken
    in order (the reference's DataLoader shuffles; everything else is its code path)."""
    rng = np.random.default_rng(2024)
    X, S, _, mu, _ = synth(rng, N, D, K)
    torch.set_default_dtype(dtype)
    m = SGDDeconvGMM(K, D, w=1e-3, batch_size=batch, lr=2e-2, lr_step=2)
    sw = rng.standard_normal(K)
    l_diag = 0.3 * rng.standard_normal((K, D))
    l_lower = 0.5 * rng.standard_normal((K, D * (D - 1) // 2))
    m.module.soft_weights.data = torch.tensor(sw, dtype=dtype)
    m.module.means.data = torch.tensor(mu, dtype=dtype)
    m.module.l_diag.data = torch.tensor(l_diag, dtype=dtype)
    m.module.l_lower.data = torch.tensor(l_lower, dtype=dtype)
    tX, tS = torch.tensor(X, dtype=dtype), torch.tensor(S, dtype=dtype)
    out = dict(X=X, S=S, soft_weights=sw, means=mu, l_diag=l_diag, l_lower=l_lower)
    curve = []
    for ep in range(epochs):
        train_loss = 0.0
        for lo in range(0, N, batch):
            d = [tX[lo:lo + batch], tS[lo:lo + batch]]
            m.optimiser.zero_grad()
            log_prob = m.module(d)
            loss = -1 * torch.mean(log_prob)
            train_loss += torch.sum(log_prob).item()
            loss += m.reg_loss(d[0].shape[0], N)
            loss.backward()
            m.optimiser.step()
        curve.append(train_loss / N)
        m.scheduler.step()
        if ep + 1 in (1, 3, epochs):
            for pn in ("soft_weights", "means", "l_diag", "l_lower"):
                out["%s_after_%d" % (pn, ep + 1)] = getattr(m.module, pn).detach().numpy().copy()
    out["train_loss_curve"] = np.asarray(curve)
    out["lr_final"] = m.optimiser.param_groups[0]["lr"]
    torch.set_default_dtype(torch.float32)
    np.savez_compressed(os.path.join(HERE, name), **out)
    print(name, curve)


def posterior_sample_case(name, N=4, D=3, K=4, M=60000):
    """Monte-Carlo moments of the reference's SGDDeconvGMM._sample_posterior (sgd_deconv_gmm.py:107-119):
    its torch RNG streams cannot be reproduced elsewhere, its distribution can."""
    rng = np.random.default_rng(99)
    X, S, _, mu, _ = synth(rng, N, D, K, spread=1.5)
    torch.set_default_dtype(torch.float64)
    m = SGDDeconvGMM(K, D, w=1e-3, batch_size=16)
    sw = 0.5 * rng.standard_normal(K)
    l_diag = 0.3 * rng.standard_normal((K, D))
    l_lower = 0.5 * rng.standard_normal((K, D * (D - 1) // 2))
    m.module.soft_weights.data = torch.tensor(sw)
    m.module.means.data = torch.tensor(mu)
    m.module.l_diag.data = torch.tensor(l_diag)
    m.module.l_lower.data = torch.tensor(l_lower)
    m.device = torch.device("cpu")
    torch.manual_seed(7)
    with torch.no_grad():
        smp = m._sample_posterior((torch.tensor(X), torch.tensor(S)), M).numpy()   # [N,M,D]
        pw, pm, pc = m.posterior_params((torch.tensor(X), torch.tensor(S)))
    mean = smp.mean(axis=1)
    cov = np.stack([np.cov(smp[i].T) for i in range(N)])
    # posterior_log_prob (sgd_deconv_gmm.py:96-105) for [N,M,D] and [Q,D] queries
    Vq = X[:, None, :] + rng.standard_normal((N, 3, D))
    x2 = X[:3] + 0.5 * rng.standard_normal((3, D))
    with torch.no_grad():
        lp3 = m.posterior_log_prob(torch.tensor(Vq), (torch.tensor(X), torch.tensor(S))).numpy()
        lp2 = m.posterior_log_prob(torch.tensor(x2), (torch.tensor(X), torch.tensor(S))).numpy()
    torch.set_default_dtype(torch.float32)
    np.savez_compressed(os.path.join(HERE, name), X=X, S=S, soft_weights=sw, means=mu, l_diag=l_diag, l_lower=l_lower,
                        M=M, sample_mean=mean, sample_cov=cov, post_log_w=pw.numpy(), post_means=pm.numpy(),
                        post_covars=pc.numpy(), query3=Vq, query2=x2, logprob3=lp3, logprob2=lp2)
    print(name, mean, lp3.shape, lp2.shape)


def fit_api_case(name):
    """END-TO-END ``fit`` of the reference's ``SGDGMM`` (deconv/gmm/sgd_gmm.py:264-272: data[0] only, no noise)
    and of the ``_tim`` fork's ``SGDDeconvGMM`` (deconv/gmm/sgd_deconv_gmm_tim.py + sgd_gmm_tim.py:112-227) - the
    reference's own loops INCLUDING their minibatch-k-means initialisation, scheduler / logger / early-stop
    bookkeeping.  The only intervention: ``DataLoader`` is asked not to shuffle, so that the batch order (and with
    it the k-means seeds = the first K rows) is reproducible."""
    import logging
    import torch.utils.data as data_utils
    import deconv.gmm.sgd_gmm as ref_sgd
    import deconv.gmm.sgd_gmm_tim as ref_tim
    import deconv.gmm.sgd_deconv_gmm_tim as ref_dtim

    real_loader = data_utils.DataLoader

    class OrderedLoader(real_loader):
        def __init__(self, *a, **kw):
            kw["shuffle"] = False
            super().__init__(*a, **kw)

    rng = np.random.default_rng(777)
    N, D, K = 96, 3, 3
    centres = np.array([[-6.0, 0.0, 3.0], [5.0, 4.0, -2.0], [0.0, -7.0, 0.0]])
    comp = np.arange(N) % K                       # the first K rows belong to K different clusters
    qn = 0.3 * rng.standard_normal((N, D, D))
    S = np.einsum("nba,nbc->nac", qn, qn) + 1e-2 * np.eye(D)
    X = centres[comp] + rng.standard_normal((N, D)) + np.einsum(
        "nab,nb->na", np.linalg.cholesky(S), rng.standard_normal((N, D)))
    Xv = centres[comp[:48]] + rng.standard_normal((48, D))
    Sv = S[:48]
    out = dict(X=X.astype(np.float32), S=S.astype(np.float32), Xv=Xv.astype(np.float32), Sv=Sv.astype(np.float32))
    tX, tS = torch.tensor(out["X"]), torch.tensor(out["S"])
    tXv, tSv = torch.tensor(out["Xv"]), torch.tensor(out["Sv"])
    data_utils.DataLoader = OrderedLoader
    try:
        # (1) SGDGMM.fit: plain GMM on X, validation set, scheduler milestones at [2, 7]
        torch.manual_seed(0)
        m = ref_sgd.SGDGMM(K, D, epochs=5, lr=2e-2, batch_size=16, w=1e-3)
        m.k_means_iters = 3
        m.fit(data_utils.TensorDataset(tX), val_data=data_utils.TensorDataset(tXv), verbose=False)
        out["plain_train_curve"] = np.asarray(m.train_loss_curve)
        out["plain_val_curve"] = np.asarray(m.val_loss_curve)
        for pn in ("soft_weights", "means", "l_diag", "l_lower"):
            out["plain_" + pn] = getattr(m.module, pn).detach().numpy().copy()
        out["plain_score_batch"] = m.score_batch(data_utils.TensorDataset(tX))
        # (2) the _tim fork: SGDDeconvGMM.fit(data, logger, val_data, verbose, interval, es_tl)
        records = []

        class Sink(logging.Handler):
            def emit(self, rec):
                records.append(rec.getMessage())
        logger = logging.getLogger("dgmm_golden")
        logger.setLevel(logging.INFO)
        logger.addHandler(Sink())
        t = ref_dtim.SGDDeconvGMM(K, D, epochs=4, lr=2e-2, batch_size=16, w=1e-3, k_means_iters=3)
        ds = ref_dtim.SGDDeconvDataset(tX, tS)
        ret = t.fit(ds, logger, val_data=ref_dtim.SGDDeconvDataset(tXv, tSv), verbose=True, interval=2)
        out["tim_train_curve"] = np.asarray(t.train_loss_curve)
        out["tim_val_curve"] = np.asarray(t.val_loss_curve)
        out["tim_returned_none"] = ret is None
        out["tim_n_messages"] = len(records)
        out["tim_lr_final"] = t.optimiser.param_groups[0]["lr"]
        for pn in ("soft_weights", "means", "l_diag", "l_lower"):
            out["tim_" + pn] = getattr(t.module, pn).detach().numpy().copy()
        # es_tl: stops after the first epoch whose training loss is below the threshold
        t2 = ref_dtim.SGDDeconvGMM(K, D, epochs=50, lr=2e-2, batch_size=16, w=1e-3, k_means_iters=3)
        thr = float(out["tim_train_curve"][1]) + 1e-6
        t2.fit(ds, logger, es_tl=thr)
        out["tim_es_tl"] = thr
        out["tim_es_epochs"] = len(t2.train_loss_curve)
        assert ref_tim.BaseSGDGMM is not ref_sgd.BaseSGDGMM
    finally:
        data_utils.DataLoader = real_loader
    np.savez_compressed(os.path.join(HERE, name), **out)
    print(name, out["plain_train_curve"], out["tim_train_curve"], out["tim_es_epochs"], records)


def real_noise(seed, N, D):
    """Per-point full noise for the real tables (regenerated from the seed by the tests, not stored):
    S_i = Q_i^T Q_i + 1e-3 I with Q_i ~ (0.5 / sqrt(D)) N(0,1)^{DxD} (experiments/gmm/checks/data.py:9-10
    scaled so that trace(S_i)/D ~ 0.25 on standardised columns)."""
    rng = np.random.default_rng(seed)
    Q = (0.5 / np.sqrt(D)) * rng.standard_normal((N, D, D))
    S = np.einsum("nba,nbc->nac", Q, Q) + 1e-3 * np.eye(D)
    L = np.linalg.cholesky(S)
    noise = np.einsum("nab,nb->na", L, rng.standard_normal((N, D)))
    return S, noise


def tril_pack(C):
    D = C.shape[-1]
    ia, ib = np.tril_indices(D)
    return C[..., ia, ib]


def real_em_case(name, X0, K, iters, seed, w=1e-3, sgd=None):
    """EM from the reference's own style of injected init (alpha = 1/K, mu = K distinct rows, Sigma = I:
    deconv/gmm/util.py:42, online_deconv_gmm.py:32-36) on a real table with synthetic per-point full
    noise; float64 reference, parameters stored as float32 lower triangles after the listed iteration counts."""
    dtype = torch.float64
    torch.set_default_dtype(dtype)
    N, D = X0.shape
    S, noise = real_noise(seed, N, D)
    X = X0 + noise
    rng = np.random.default_rng(seed + 1)
    rows = np.sort(rng.choice(N, size=K, replace=False))
    m = DeconvGMM(K, D, w=w)
    inject(m, np.full(K, 1.0 / K), X[rows].copy(), np.tile(np.eye(D), (K, 1, 1)), dtype)
    data = (torch.tensor(X), torch.tensor(S))
    out = dict(X=X.astype(np.float32), X64=X, rows=rows, w=w, seed=seed, K=K)
    lls = []
    for it in range(1, max(iters) + 1):
        ll, e = m._e_step(data)
        lls.append(ll.item())
        m._m_step(data, e)
        del e
        if it in iters:
            out["it%d_weights" % it] = m.weights.numpy().reshape(-1).astype(np.float32)
            out["it%d_means" % it] = m.means.numpy().astype(np.float32)
            out["it%d_covars_tril" % it] = tril_pack(m.covars.numpy()).astype(np.float32)
        print(name, "iteration", it, "ll", lls[-1], flush=True)
    out["ll_trace"] = np.asarray(lls)
    if sgd is not None:
        # a short SGD trajectory of the reference's module / reg_loss / Adam from the same init, batches in order
        epochs, batch, lr = sgd
        sg = SGDDeconvGMM(K, D, w=w, batch_size=batch, lr=lr)
        sg.module.soft_weights.data = torch.zeros(K, dtype=dtype)
        sg.module.means.data = torch.tensor(X[rows].copy())
        sg.module.l_diag.data = torch.zeros(K, D, dtype=dtype)
        sg.module.l_lower.data = torch.zeros(K, D * (D - 1) // 2, dtype=dtype)
        tX, tS = data
        curve = []
        for ep in range(epochs):
            tot = 0.0
            for lo in range(0, N, batch):
                d = [tX[lo:lo + batch], tS[lo:lo + batch]]
                sg.optimiser.zero_grad()
                lp = sg.module(d)
                loss = -1 * torch.mean(lp)
                tot += torch.sum(lp).item()
                loss += sg.reg_loss(d[0].shape[0], N)
                loss.backward()
                sg.optimiser.step()
            curve.append(tot / N)
            sg.scheduler.step()
            print(name, "sgd epoch", ep, curve[-1], flush=True)
        out["sgd_epochs"], out["sgd_batch"], out["sgd_lr"] = epochs, batch, lr
        out["sgd_curve"] = np.asarray(curve)
        for pn in ("soft_weights", "means", "l_diag", "l_lower"):
            out["sgd_" + pn] = getattr(sg.module, pn).detach().numpy().astype(np.float32)
    torch.set_default_dtype(torch.float32)
    np.savez_compressed(os.path.join(HERE, name), **out)
    print(name, lls)


def ionosphere_case(name):
    """BASELINE.json configs[3] / SURVEY.md 8d item 4: ionosphere (351x32), K = 256, full noise, injected
    init, 5 EM iterations, w = 1e-3."""
    X0 = np.load(os.path.join(REF, "data_small", "ionosphere_no_discrete_no_corr_0.98.npy"))
    real_em_case(name, X0, K=256, iters=(1, 5), seed=4004)


def wine_case(name):
    """BASELINE.json configs[2]: white + red wine (6497x9), K = 128, full noise, injected init, parameters
    after 1 / 5 / 20 EM iterations and a 2-epoch SGD trajectory."""
    X0 = np.concatenate([np.load(os.path.join(REF, "data_small", "white_no_discrete_no_corr_0.98.npy")),
                         np.load(os.path.join(REF, "data_small", "red_no_discrete_no_corr_0.98.npy"))])
    real_em_case(name, X0, K=128, iters=(1, 5, 20), seed=3003, sgd=(2, 500, 1e-3))


def main():
    if len(sys.argv) > 1 and sys.argv[1] == "posterior_sample":
        posterior_sample_case("ref_posterior_sample.npz")
        return
    if len(sys.argv) > 1 and sys.argv[1] == "sgd_fit":
        sgd_fit_case("ref_sgd_fit_d5_k3_f32.npz", N=64, D=5, K=3, dtype=torch.float32)
        sgd_fit_case("ref_sgd_fit_d3_k4_f64.npz", N=48, D=3, K=4, dtype=torch.float64)
        return
    if len(sys.argv) > 1 and sys.argv[1] == "fit_api":
        fit_api_case("ref_fit_api.npz")
        return
    if len(sys.argv) > 1 and sys.argv[1] == "ionosphere":
        ionosphere_case("ref_ionosphere_k256_f64.npz")
        return
    if len(sys.argv) > 1 and sys.argv[1] == "wine":
        wine_case("ref_wine_k128_f64.npz")
        return
    if len(sys.argv) > 1 and sys.argv[1] == "kmeans":
        kmeans_case("ref_kmeans.npz")
        return
    rng = np.random.default_rng(12345)
    em_case("ref_em_d3_k4_f64.npz", rng, N=40, D=3, K=4, dtype=torch.float64)
    em_case("ref_em_d10_k6_f64.npz", rng, N=24, D=10, K=6, dtype=torch.float64)
    em_case("ref_em_d10_k6_f32.npz", rng, N=24, D=10, K=6, dtype=torch.float32)
    em_case("ref_em_d16_k5_f64.npz", rng, N=12, D=16, K=5, dtype=torch.float64, iters=(1,))
    sgd_case("ref_sgd_d4_k5_f64.npz", rng, N=30, D=4, K=5, dtype=torch.float64)
    sgd_case("ref_sgd_d7_k3_f32.npz", rng, N=20, D=7, K=3, dtype=torch.float32)
    online_batch_case("ref_online_batch_d3_k4_f64.npz", rng, N=32, D=3, K=4, dtype=torch.float64)
    online_batch_case("ref_online_batch_d5_k3_f32.npz", rng, N=32, D=5, K=3, dtype=torch.float32)
    data_gen_case("ref_data_gen.npz")
    robustness_case("ref_robust_f32.npz")
    boston_case("ref_boston_k10_f64.npz")
    kmeans_case("ref_kmeans.npz")
    sgd_fit_case("ref_sgd_fit_d5_k3_f32.npz", N=64, D=5, K=3, dtype=torch.float32)
    sgd_fit_case("ref_sgd_fit_d3_k4_f64.npz", N=48, D=3, K=4, dtype=torch.float64)
    posterior_sample_case("ref_posterior_sample.npz")
    fit_api_case("ref_fit_api.npz")
    ionosphere_case("ref_ionosphere_k256_f64.npz")
    wine_case("ref_wine_k128_f64.npz")


if __name__ == "__main__":
    main()

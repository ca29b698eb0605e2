#!/usr/bin/env python
"""Benchmark of the deconvolution-GMM hot path (BASELINE.json metric:
"(point,component) log-lik evals/sec and EM iters/sec").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c2|c3|c5|small] [--impl reference]

A STEP is one full EM iteration over the rank's resident shard: the fused statistics kernel
(log-likelihood + E-step + M-step reductions for every (point, component) pair), the all-reduce
of the packed statistics (N > 1) and the closed-form parameter update.  ``value`` is
(point, component) evaluations per second over the whole job inside those steps;
``em_iters_per_sec`` and the log-likelihood-only pass are reported beside it.

Workloads (per GPU; weak scaling, every rank generates its own shard on the device):
  c2    synthetic N=1e7,   D=10, K=64,  full per-point noise   (BASELINE.json configs[1], default)
  c3    wine-shaped N=1.25e7, D=9,  K=128 (configs[2]: 1e8 points over 8 GPUs)
  c5    synthetic N=1.25e7, D=16, K=128 (configs[4]: 1e8 points over 8 GPUs, headline shape)
  small N=1e6, D=10, K=64 (quick check)

The ``--impl reference`` arm and the ``cpu_baseline`` object time a torch-CPU port of the
reference's own E-step/M-step call sequence (oracle/torch_port.py) on the host cores.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WORKLOADS = {
    "c2": dict(N=10_000_000, D=10, K=64, noise="full", name="synthetic N=1e7/GPU D=10 K=64 full per-point noise"),
    "c3": dict(N=12_500_000, D=9, K=128, noise="full", name="wine-shaped N=1.25e7/GPU D=9 K=128 full noise"),
    "c5": dict(N=12_500_000, D=16, K=128, noise="full", name="synthetic N=1.25e7/GPU D=16 K=128 full noise"),
    "small": dict(N=1_000_000, D=10, K=64, noise="full", name="synthetic N=1e6/GPU D=10 K=64 full noise"),
}
METRIC = "pair_loglik_evals_per_sec"
UNIT = "(point,component) evals/s"


def F_ll(D):
    """Algorithmic flops per pair of a log-likelihood pass."""
    return D ** 3 / 3 + 2 * D ** 2 + 17 / 3 * D + 8


def F_em(D):
    """Algorithmic flops per pair of an EM iteration / SGD forward+backward."""
    return D ** 3 + 6.5 * D ** 2 + 10.5 * D + 10


def ncu_traffic(workload, kernel):
    """DRAM bytes per launch of ``kernel`` measured by ncu on this workload (None if not captured)."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            return json.load(f).get(workload, {}).get(kernel)
    except (OSError, ValueError):
        return None


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--points", type=int, default=0, help="override points per GPU")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-sgd", action="store_true")
    ap.add_argument("--sgd-batch", type=int, default=65536, help="SGD minibatch per GPU")
    return ap.parse_args()


# -----------------------------------------------------------------------------------------
# CPU arm: torch port of the reference on the host cores
# -----------------------------------------------------------------------------------------
def cpu_problem(wl, n, seed=99):
    import torch
    from density_deconvolution_b200.synthetic import make_mixture, make_points
    alpha, mu, Sigma, q = make_mixture(wl["D"], wl["K"], seed=0, device="cpu")
    X, S = make_points(n, alpha, mu, q, seed=seed, noise=wl["noise"])
    g = torch.Generator().manual_seed(7)
    mu0 = mu + 0.3 * torch.randn(mu.shape, generator=g)
    return X, S, alpha, mu0, Sigma * 1.2


def cpu_em_rate(wl, n, reps=1):
    """(pairs/s, seconds per iteration) of the reference port's EM iteration on ``n`` points."""
    import torch
    from oracle import torch_port as tp
    X, S, alpha, mu, Sigma = cpu_problem(wl, n)
    chunk = max(256, min(4096, (1 << 28) // (wl["K"] * wl["D"] * wl["D"] * 4 * 8)))
    best = float("inf")
    for _ in range(reps):
        t0 = time.perf_counter()
        tp.em_iteration_single_pass(X, S, alpha, mu, Sigma, 1e-3, chunk=chunk)
        best = min(best, time.perf_counter() - t0)
    return n * wl["K"] / best, best, chunk


def cpu_baseline(wl, budget_s=15.0):
    import torch
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    rate, _, chunk = cpu_em_rate(wl, 2048)                       # warm-up + speed estimate
    n = int(min(131072, max(4096, rate * budget_s / wl["K"])))
    n = (n // 1024) * 1024
    rate, secs, chunk = cpu_em_rate(wl, n)
    return dict(value=rate, unit=UNIT, cores=cores, kind="port",
                sample="one EM iteration (E-step + M-step, torch-CPU port of deconv/gmm/deconv_gmm.py:39-109, "
                       "float32, %d torch threads, chunks of %d points) on %d points of the same workload: %.1f s"
                       % (torch.get_num_threads(), chunk, n, secs))


def run_reference(args, wl):
    """``--impl reference``: the reference's CPU path (torch port) on the host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    rate, _, _ = cpu_em_rate(wl, 1024)
    # each step: a bounded sample sized for ~3 s so that K+W steps end within a few minutes
    n = int(min(65536, max(1024, rate * 3.0 / wl["K"])))
    n = (n // 512) * 512
    from oracle import torch_port as tp
    X, S, alpha, mu, Sigma = cpu_problem(wl, n)
    chunk = max(256, min(4096, (1 << 28) // (wl["K"] * wl["D"] * wl["D"] * 4 * 8)))
    w = alpha[:, None]
    for _ in range(args.warmup):
        _, w, mu, Sigma = tp.em_iteration_single_pass(X, S, w.reshape(-1), mu, Sigma, 1e-3, chunk=chunk)
        w = w
    t0 = time.perf_counter()
    for _ in range(args.steps):
        _, w, mu, Sigma = tp.em_iteration_single_pass(X, S, w.reshape(-1), mu, Sigma, 1e-3, chunk=chunk)
    dt = time.perf_counter() - t0
    value = n * wl["K"] * args.steps / dt
    sample = ("each step = one EM iteration of the torch-CPU port of the reference (oracle/torch_port.py) on a "
              "%d-point sample of the workload, float32, %d torch threads" % (n, torch.get_num_threads()))
    line = dict(impl="reference", metric=METRIC, value=value, unit=UNIT, n_gpus=args.gpus, steps=args.steps,
                warmup=args.warmup, ms_per_step=dt / args.steps * 1e3, higher_is_better=True, scaling="weak",
                vs_baseline=None, dtype="f32", data="synthetic",
                config=dict(workload=wl["name"], D=wl["D"], K=wl["K"], sample_points=n),
                em_iters_per_sec_full_workload=value / (wl["N"] * wl["K"]),
                cpu_baseline=dict(value=value, unit=UNIT, cores=cores, kind="port", sample=sample),
                e2e=dict(value=value, unit=UNIT, h2d_bytes_per_step=0, d2h_bytes_per_step=0),
                gpu_launches=0)
    emit(line)


# -----------------------------------------------------------------------------------------
# clocks during the timed region
# -----------------------------------------------------------------------------------------
class ClockSampler:
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_id):
        self.gpu_id = gpu_id
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.gpu_id), "--query-gpu=" + self.FIELDS, "--format=csv,noheader,nounits",
                 "-lms", "50"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), [c.strip() for c in line.split(",")]))

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except subprocess.TimeoutExpired:
                self.proc.kill()

    def summary(self, t0, t1):
        rows = [r for (t, r) in self.rows if t0 <= t <= t1 and len(r) >= 7] or [r for (_, r) in self.rows if len(r) >= 7]
        if not rows:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=["nvidia-smi unavailable"])
        sm = sorted(float(r[0]) for r in rows)
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        reasons = [n for i, n in enumerate(names) if any(r[3 + i].lower().startswith("active") for r in rows)]
        power = max(float(r[2]) for r in rows if r[2].replace(".", "", 1).isdigit())
        return dict(sm_mhz=sm[len(sm) // 2], sm_max_mhz=float(rows[0][1]), reasons=reasons, samples=len(rows),
                    power_w_max=power)


# -----------------------------------------------------------------------------------------
# GPU arm
# -----------------------------------------------------------------------------------------
def run_b200(args, wl):
    import torch
    import torch.distributed as dist
    from density_deconvolution_b200 import ops
    from density_deconvolution_b200.synthetic import make_mixture, make_points

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the B200 path has no CPU fallback; use --impl reference)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    if world != args.gpus and rank == 0:
        print("bench.py: WORLD_SIZE=%d but --gpus %d; using WORLD_SIZE" % (world, args.gpus), file=sys.stderr)

    N, D, K = (args.points or wl["N"]), wl["D"], wl["K"]
    alpha, mu_true, Sigma_true, q = make_mixture(D, K, seed=0, device=dev)
    X, S = make_points(N, alpha, mu_true, q, seed=1234 + rank, noise=wl["noise"])
    del q
    g = torch.Generator(device="cpu").manual_seed(7)
    mu0 = (mu_true.cpu() + 0.3 * torch.randn(K, D, generator=g)).to(dev)
    Sigma0 = (1.2 * Sigma_true).contiguous()

    def fresh():
        a = alpha.clone()
        return a, a.log(), mu0.clone(), Sigma0.clone()

    stats = torch.empty(ops.stats_len(D, K), dtype=torch.float32, device=dev)
    n_total = N * world
    w_reg = 1e-3

    def em_step(p):
        a, la, m, Sg = p
        ops.estep_stats(X, S, la, m, Sg, noise=wl["noise"], want_ll=False, out_stats=stats)
        if world > 1:
            dist.all_reduce(stats)
        ops.mstep_update(stats, n_total, w_reg, m, Sg, a, la)

    def ll_step(p):
        a, la, m, Sg = p
        ops.loglik(X, S, la, m, Sg, noise=wl["noise"])

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, params, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        for _ in range(steps):
            fn(params)
        e1.record()
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        barrier()
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return ms.item() * 1e-3, (t0, t1)

    sampler = None
    if rank == 0:
        try:
            gpu_id = "GPU-" + str(torch.cuda.get_device_properties(dev).uuid)
        except Exception:  # noqa: BLE001
            gpu_id = str(local_rank)
        sampler = ClockSampler(gpu_id)
        sampler.start()

    fp32_peak = ops.fp32_probe(dev)

    # ---- the timed region: K EM iterations ------------------------------------------------
    params = fresh()
    for _ in range(max(args.warmup, 3)):
        em_step(params)
    ops.pair_timing(True)
    t_em, window = timed(em_step, params, args.steps)
    pair_ms = ops.pair_times()
    ops.pair_timing(False)
    ll_after = (stats[-4] + stats[-3]).item() / n_total
    nonpd = stats[-1].item()

    # ---- log-likelihood-only pass (score_batch), same protocol ------------------------------
    for _ in range(3):
        ll_step(params)
    ops.pair_timing(True)
    t_ll, _ = timed(ll_step, params, args.steps)
    pair_ms_ll = ops.pair_times()
    ops.pair_timing(False)

    # ---- minibatch SGD steps through the estimator API (configs[4]: Adam, lr 1e-3, w 1e-3) --------
    sgd = None
    if not args.no_sgd:
        sgd = run_sgd(args, wl, dev, world, X, S, fresh, barrier)

    if sampler:
        time.sleep(0.06)
        sampler.stop()
    clocks = sampler.summary(*window) if sampler else None

    # ---- end to end through the host-buffer C ABI (pinned host memory -> GPU -> host) --------
    e2e = None
    if not args.no_e2e:
        e2e = run_e2e(args, wl, ops, dist, dev, world, X, S, fresh, n_total, w_reg, barrier)

    pairs_per_step = float(n_total) * K
    value = pairs_per_step * args.steps / t_em
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    pair_avg_s = (sum(pair_ms) / len(pair_ms)) * 1e-3 if pair_ms else None
    pair_ll_avg_s = (sum(pair_ms_ll) / len(pair_ms_ll)) * 1e-3 if pair_ms_ll else None
    peaks = {}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peaks = json.load(f)
    except OSError:
        pass
    hbm_peak = peaks.get("hbm_gbs", 6650.0)
    bytes_per_point = 4 * (D + (D * D if wl["noise"] == "full" else D))
    plan = ops.get_plan(1, D, K, N, noise=wl["noise"])
    roofline = dict(
        bound="fp32", kernel="pair_kernel<MODE_STATS>",
        achieved=N * K * F_em(D) / pair_avg_s / 1e12, peak=fp32_peak / 1e12, unit="TFLOP/s",
        frac=N * K * F_em(D) / pair_avg_s / fp32_peak,
        peak_source="in-bench register-only FFMA probe on this GPU (MEASURED_PEAKS.json has no FP32 entry; "
                    "nominal 148 SM x 128 lanes x 2 x 1.965 GHz = 74.4)",
        flops_per_pair=F_em(D), pairs_per_launch=N * K, launch_ms=pair_avg_s * 1e3,
        share_of_step=pair_avg_s * args.steps / t_em if world == 1 else None,
        traffic=ncu_traffic(args.workload, "pair_kernel<MODE_STATS>") if not args.points else None,
        traffic_note="DRAM bytes per launch from the committed ncu --set full capture (profiles/ncu_traffic.json); "
                     "algorithmic bytes per launch = %d" % (N * bytes_per_point),
        hbm=dict(achieved=N * bytes_per_point / pair_avg_s / 1e9, peak=hbm_peak, unit="GB/s",
                 frac=N * bytes_per_point / pair_avg_s / 1e9 / hbm_peak,
                 peak_source="MEASURED_PEAKS.json" if peaks else "fallback of B200_PROFILING.md"),
        loglik_kernel=dict(kernel="pair_kernel<MODE_LOGLIK>", launch_ms=pair_ll_avg_s * 1e3,
                           achieved=N * K * F_ll(D) / pair_ll_avg_s / 1e12,
                           frac=N * K * F_ll(D) / pair_ll_avg_s / fp32_peak, flops_per_pair=F_ll(D)))
    line = dict(
        metric=METRIC, value=value, unit=UNIT, n_gpus=world, steps=args.steps, warmup=max(args.warmup, 3),
        ms_per_step=t_em / args.steps * 1e3, higher_is_better=True, scaling="weak", vs_baseline=None,
        dtype="f32", data="synthetic",
        config=dict(workload=wl["name"], points_per_gpu=N, points_total=n_total, D=D, K=K, noise=wl["noise"],
                    step="one EM iteration: fused log-lik + E-step + M-step statistics kernel, "
                         "stats all-reduce (N>1), closed-form parameter update",
                    l2="inputs (%.1f GB per GPU) are far larger than the 126 MB L2; no flush needed"
                       % (N * bytes_per_point / 1e9),
                    parallelism="points sharded by rank, %d rank(s)" % world, launch_plan=plan),
        em_iters_per_sec=args.steps / t_em,
        loglik_only=dict(value=pairs_per_step * args.steps / t_ll, unit=UNIT, ms_per_pass=t_ll / args.steps * 1e3),
        mean_loglik_after=ll_after, nonpd_pairs=nonpd,
        gpu_launches=4 * args.steps,
        gpu_launches_note="per step: prepare_kernel, pair_kernel<STATS>, finalize_stats_kernel, mstep_kernel",
        sgd=sgd, clocks=clocks, roofline=roofline, e2e=e2e)
    if world == 1 and not args.no_cpu:
        line["cpu_baseline"] = cpu_baseline(wl)
    emit(line)
    if world > 1:
        dist.destroy_process_group()


def run_sgd(args, wl, dev, world, X, S, fresh, barrier, n_steps=20):
    """Minibatch-SGD steps (forward + backward in ONE fused statistics launch, stats all-reduce,
    Adam on the [K,.] parameters) through ``SGDDeconvGMM.fit`` on a resident slice of the shard:
    ``n_steps`` minibatches of ``--sgd-batch`` points per GPU, one epoch, device-side shuffling."""
    import torch
    from density_deconvolution_b200.estimators import SGDDeconvGMM
    N, D = X.shape
    K = wl["K"]
    B = min(args.sgd_batch, N)
    n_steps = max(1, min(n_steps, N // B))
    n_sub = n_steps * B
    a, _, m, Sg = fresh()

    def make():
        est = SGDDeconvGMM(K, D, epochs=1, lr=1e-3, batch_size=B, w=1e-3, device=dev)
        with torch.no_grad():
            est.module.soft_weights.copy_(a.log())
            est.module.means.copy_(m)
            L = torch.linalg.cholesky(Sg.double()).float()
            est.module.l_diag.copy_(torch.diagonal(L, dim1=1, dim2=2).log())
            idx = est.module.l_idx
            est.module.l_lower.copy_(L[:, idx[0], idx[1]])
        return est

    make().fit((X[:min(3 * B, n_sub)], S[:min(3 * B, n_sub)]), init=False)       # warm-up (3 steps)
    est = make()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    est.fit((X[:n_sub], S[:n_sub]), init=False)
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    barrier()
    if world > 1:
        import torch.distributed as dist
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    t = ms.item() * 1e-3
    return dict(api="SGDDeconvGMM.fit (1 epoch, device-resident batches, shuffle on device)",
                batch_per_gpu=B, global_batch=B * world, steps=n_steps, ms_per_step=t / n_steps * 1e3,
                steps_per_sec=n_steps / t, points_per_sec=float(n_sub) * world / t,
                pair_evals_per_sec=float(n_sub) * world * K / t, mean_loglik=est.train_loss_curve[-1])


def run_e2e(args, wl, ops, dist, dev, world, X, S, fresh, n_total, w_reg, barrier):
    """The same EM iteration through the host-buffer C ABI: every step streams the rank's shard
    from pinned host memory (H2D inside the timed region) and reads the updated parameters and
    the log-likelihood back (D2H)."""
    import torch
    N, D = X.shape
    K = wl["K"]
    hX = torch.empty(X.shape, dtype=torch.float32, pin_memory=True)
    hS = torch.empty(S.shape, dtype=torch.float32, pin_memory=True)
    hX.copy_(X)
    hS.copy_(S)
    a, _, m, Sg = fresh()
    ha, hm, hSg = a.cpu(), m.cpu(), Sg.cpu()
    ctx = ops.HostContext(D, K, chunk_points=1 << 18, device=dev.index)
    h_stats = torch.empty(ops.stats_len(D, K), dtype=torch.float32, pin_memory=True)
    d_stats = torch.empty(ops.stats_len(D, K), dtype=torch.float32, device=dev)
    d_par = [t.to(dev) for t in (ha, ha.log(), hm, hSg)]

    def step():
        if world == 1:
            return ctx.em_iteration(hX, hS, w_reg, ha, hm, hSg, noise=wl["noise"])[0]
        ctx.estep_stats(hX, hS, ha, hm, hSg, noise=wl["noise"], out=h_stats)
        d_stats.copy_(h_stats, non_blocking=True)
        dist.all_reduce(d_stats)
        d_par[2].copy_(hm)
        d_par[3].copy_(hSg)
        ops.mstep_update(d_stats, n_total, w_reg, d_par[2], d_par[3], d_par[0], d_par[1])
        ha.copy_(d_par[0])
        hm.copy_(d_par[2])
        hSg.copy_(d_par[3])
        return float(d_stats[-4].item() + d_stats[-3].item())

    steps = max(2, min(args.steps, 5))
    step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        ll = step()
    torch.cuda.synchronize()
    dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    barrier()
    if world > 1:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    ctx.close()
    bytes_in = 4 * (hX.numel() + hS.numel() + K * (1 + D + D * D))
    bytes_out = 4 * (K * (1 + D + D * D)) + 16
    return dict(value=float(n_total) * K * steps / dt.item(), unit=UNIT, steps=steps,
                ms_per_step=dt.item() / steps * 1e3, em_iters_per_sec=steps / dt.item(),
                h2d_bytes_per_step=bytes_in, d2h_bytes_per_step=bytes_out, mean_loglik=ll / n_total,
                api="dgmm_em_iteration_host (N=1) / dgmm_estep_stats_host + all-reduce + dgmm_mstep_update_f32 (N>1), "
                    "pinned host buffers, 262144-point chunks, copy/compute overlap on two streams")


_JSON_OUT = None


def emit(line):
    """The one JSON line of the contract, on the process's ORIGINAL stdout."""
    out = _JSON_OUT or sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def main():
    global _JSON_OUT
    args = parse_args()
    # stdout carries exactly one JSON line: everything else written to fd 1 by this process or the libraries it
    # loads (NCCL's version banner, extension build chatter) is sent to stderr instead
    sys.stdout.flush()
    _JSON_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    wl = dict(WORKLOADS[args.workload])
    if args.points:
        wl["N"] = args.points
    if args.impl == "reference":
        run_reference(args, wl)
    else:
        run_b200(args, wl)


if __name__ == "__main__":
    main()

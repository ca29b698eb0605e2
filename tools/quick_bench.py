"""Quick timing sweep of the pair kernel over launch-plan variants (GPU box only)."""
import argparse
import json
import sys
import os

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from density_deconvolution_b200 import ops
from density_deconvolution_b200.synthetic import make_mixture, make_points


def F_ll(D): return D ** 3 / 3 + 2 * D ** 2 + 17 / 3 * D + 8
def F_em(D): return D ** 3 + 6.5 * D ** 2 + 10.5 * D + 10


def timeit(fn, reps=3):
    fn(); torch.cuda.synchronize()
    best = 1e9
    for _ in range(reps):
        a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b) * 1e-3)
    return best


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--N", type=int, default=2_000_000)
    ap.add_argument("--D", type=int, default=10)
    ap.add_argument("--K", type=int, default=64)
    ap.add_argument("--variants", type=str, default="auto")
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    N, D, K = args.N, args.D, args.K
    alpha, mu, Sigma, q = make_mixture(D, K, seed=0, device=dev)
    X, S = make_points(N, alpha, mu, q, seed=1)
    la = alpha.log()
    peak = ops.fp32_probe(dev)
    print(json.dumps({"fp32_probe_tflops": peak / 1e12}), flush=True)
    variants = [dict()]
    if args.variants != "auto":
        variants = json.loads(args.variants)
    for v in variants:
        try:
            ops.set_tuning(**v)
            plan_s = ops.get_plan(1, D, K, N)
            plan_l = ops.get_plan(0, D, K, N)
            t_ll = timeit(lambda: ops.loglik(X, S, la, mu, Sigma))
            t_em = timeit(lambda: ops.estep_stats(X, S, la, mu, Sigma, want_ll=False))
            out = dict(variant=v, N=N, D=D, K=K, t_loglik_ms=t_ll * 1e3, t_stats_ms=t_em * 1e3,
                       loglik_pairs_per_s=N * K / t_ll, stats_pairs_per_s=N * K / t_em,
                       loglik_tflops=N * K * F_ll(D) / t_ll / 1e12, stats_tflops=N * K * F_em(D) / t_em / 1e12,
                       frac_ll_nominal=N * K * F_ll(D) / t_ll / 74.4e12, frac_em_nominal=N * K * F_em(D) / t_em / 74.4e12,
                       plan_stats=plan_s, plan_loglik=plan_l)
            print(json.dumps(out), flush=True)
        except Exception as e:  # noqa: BLE001
            print(json.dumps(dict(variant=v, error=str(e))), flush=True)
    ops.set_tuning()


if __name__ == "__main__":
    main()

"""BASELINE config 3: exact-GP refit sweep n = 1k/4k/8k, SE-ARD and Matern-5/2, P = 3 outputs, G = 16 candidates.
Prints one JSON line per (n, kernel): fits/s ((candidate, output) pairs per second), FP64 TFLOP/s on the n^3/3 + 2n^2 P
flop count of SURVEY 8(d), and -- with --cpu -- the sequential scipy loop (oracle.exact_fit) timed on a bounded sample."""
import argparse, json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def measure(n, reps=3, kernels=("se_ard", "matern52"), cpu=False):
    """One JSON-able dict per kernel family for training-set size n."""
    import torch
    from gp_mpc_rocket_landing_b200.gp import engine as E, refit
    D, P = 11, 3
    rng = np.random.default_rng(1)
    X = rng.normal(0, 0.4, (n, D))
    Y = np.stack([np.sin(2 * X[:, 0]) + 0.05 * rng.normal(size=n), X[:, 1] * X[:, 2], np.cos(X[:, 3])], axis=1)
    ls, sv, nz = refit.default_grid(D)
    G = ls.shape[0]
    lines = []
    for kind, name in ((E.SE, "se_ard"), (E.M52, "matern52")):
        if name not in kernels:
            continue
        sweep = E.ExactSweep(kind, X, Y)
        sweep.fit(ls, sv, nz)
        torch.cuda.synchronize()
        best = 1e9
        for _ in range(reps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); r = sweep.fit(ls, sv, nz); e1.record(); torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        npad = sweep.n_pad
        flops = G * (npad ** 3 / 3.0 + 2.0 * npad ** 2 * P)
        line = {"config": "BASELINE configs[2]: exact GP refit sweep", "n": n, "kernel": name, "candidates": G, "outputs": P,
                "ms": best, "fits_per_s": G * P / (best * 1e-3), "factorisations_per_s": G / (best * 1e-3),
                "tflops": flops / (best * 1e-3) / 1e12, "bound": "tensor (FP64 DMMA)", "flops_per_sweep": flops,
                "n_jittered": int((r["jitter"] > 0).sum())}
        if cpu and n <= 4000:
            from oracle import gp as ogp
            t0 = time.perf_counter()
            k = 2 if n >= 4000 else 4
            for g in range(k):
                ogp.exact_fit(kind, X, Y[:, 0], ls[g], sv[g], nz[g])
            dt = (time.perf_counter() - t0) / k
            line["cpu_fit_ms"] = dt * 1e3
            line["cpu_fits_per_s"] = 1.0 / dt
            line["cpu_threads"] = os.cpu_count()
        lines.append(line)
        del sweep, r
        torch.cuda.empty_cache()
    return lines


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sizes", default="1000,4000,8000")
    ap.add_argument("--cpu", action="store_true")
    ap.add_argument("--reps", type=int, default=3)
    a = ap.parse_args()
    for n in [int(s) for s in a.sizes.split(",")]:
        for line in measure(n, a.reps, cpu=a.cpu):
            print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()

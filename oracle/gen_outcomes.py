"""TEST INFRASTRUCTURE ONLY.  Generates tests/golden/mc_outcomes_64.npz: the CPU oracle's closed loop (oracle.mpc.run_single
= monte_carlo.py:401-583, controller = GPMPCOracle on the 1000-point exact GP at noise 1e-4, N = 20, S = 1, ADMM max_iter
1000 -- bench.py's problem) for the first 64 scenarios of the measured workload (sample_initial_condition(seed = 42 + i),
drag plant 0.25), for both target modes of the loop, run to termination.  The GPU test compares the device loop's outcome
histogram and per-scenario outcomes on the same 64 seeds.  About 10 minutes on 8 cores:

    python oracle/gen_outcomes.py
"""
import multiprocessing as mp
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
N_SEEDS = 64
_GP = None


def _run(job):
    from threadpoolctl import threadpool_limits
    threadpool_limits(1)
    from gp_mpc_rocket_landing_b200 import workloads
    from oracle import mpc as ompc
    i, mode = job
    ctl = ompc.GPMPCOracle(_GP.predict, N=20, dt=0.1, settings=ompc.OSQPSettings(max_iter=1000), max_sqp_iter=1)
    its = []

    def solve(x, xt):
        r = ctl.solve(x, xt)
        its.append(sum(r.qp_iters))
        return r.U_opt[0], r.success

    X0, _ = workloads.sample_initial_conditions(i, 1)
    code, xf, steps, nfail = ompc.run_single(solve, X0[0], dt=0.1, max_time=30.0, target_mode=mode,
                                             on_solve_failure="continue", drag_cda=0.25)
    return i, mode, code, xf, steps, nfail, int(np.sum(its))


def main():
    global _GP
    from gp_mpc_rocket_landing_b200 import workloads
    from oracle import gp as ogp
    X, U, D = workloads.residual_dataset(1000, seed=0)
    _GP = ogp.Simple3DoFGPOracle(noise_variance=1e-4)
    _GP.add_data(X, U, D)
    _GP.fit()
    jobs = [(i, m) for m in ("per_step", "fixed") for i in range(N_SEEDS)]
    out = {}
    with mp.get_context("fork").Pool(os.cpu_count()) as pool:
        for k, r in enumerate(pool.imap_unordered(_run, jobs)):
            out[(r[0], r[1])] = r
            print(k + 1, "/", len(jobs), r[:3], r[4:], flush=True)
    save = {}
    for m in ("per_step", "fixed"):
        rs = [out[(i, m)] for i in range(N_SEEDS)]
        save[m + "_outcome"] = np.array([r[2] for r in rs], dtype=np.int32)
        save[m + "_x_final"] = np.stack([r[3] for r in rs])
        save[m + "_steps"] = np.array([r[4] for r in rs], dtype=np.int32)
        save[m + "_failed_solves"] = np.array([r[5] for r in rs], dtype=np.int32)
        save[m + "_admm_iters"] = np.array([r[6] for r in rs], dtype=np.int64)
    np.savez_compressed(os.path.join(ROOT, "tests", "golden", "mc_outcomes_64.npz"), **save)


if __name__ == "__main__":
    main()

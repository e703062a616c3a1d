"""GPU-box host profile (not part of the product) of bench.py's end-to-end step: where the host time of the numpy API goes."""
import cProfile, os, pstats, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import bench
from gp_mpc_rocket_landing_b200 import workloads


class A:
    batch, sqp_iters, qp_max_iter = 4096, 1, 1000


dyn, gp, ctl, mc = bench.build_problem(A, 0)
X0, XT = workloads.sample_initial_conditions(0, 4096)
x = X0.copy()


def step(x):
    s = ctl.solve_batch(x, XT, per_step_target=True)
    xn = dyn.step_batch(x, s.u0, 0.1)
    done = (xn[:, 1] < 0.0) | (xn[:, 0] <= 1.01) | ((xn[:, 1] < 1.0) & (np.abs(xn[:, 4]) < 5.0)) | ~np.all(np.abs(xn) <= 1e6, axis=1)
    if done.any():
        xn[done] = X0[done]
        ctl.reset_scenarios(done)
    return xn


for _ in range(10):
    x = step(x)
t0 = time.perf_counter()
for _ in range(40):
    x = step(x)
print("ms per e2e step", (time.perf_counter() - t0) / 40 * 1e3)
pr = cProfile.Profile(); pr.enable()
for _ in range(40):
    x = step(x)
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(28)

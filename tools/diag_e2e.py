"""Diagnostic: where the host-buffer (e2e) step spends its time next to the device-resident step."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import bench
from gp_mpc_rocket_landing_b200 import workloads

class A: pass
a = A(); a.batch = 4096; a.sqp_iters = 1; a.qp_max_iter = 1000
dyn, gp, ctl, mc = bench.build_problem(a, 0)
X0, XT = workloads.sample_initial_conditions(0, a.batch)
x = X0.copy()
for _ in range(10):
    s = ctl.solve_batch(x, XT); x = dyn.step_batch(x, s.u0, bench.DT)
t = {"solve_batch": 0.0, "step_batch": 0.0, "respawn": 0.0}
K = 40
torch.cuda.synchronize(); T0 = time.perf_counter()
for _ in range(K):
    t0 = time.perf_counter(); s = ctl.solve_batch(x, XT); t1 = time.perf_counter()
    x = dyn.step_batch(x, s.u0, bench.DT); t2 = time.perf_counter()
    done = (x[:, 1] < 0.0) | (x[:, 0] <= 1.01) | ((x[:, 1] < 1.0) & (np.abs(x[:, 4]) < 5.0)) | ~np.all(np.abs(x) <= 1e6, axis=1)
    t2b = time.perf_counter()
    if done.any():
        x[done] = X0[done]; ts = time.perf_counter(); torch.cuda.synchronize(); t["pre_sync"] = t.get("pre_sync", 0.0) + time.perf_counter() - ts; t2c = time.perf_counter(); ctl.reset_scenarios(done, X0)
    else:
        t2c = time.perf_counter()
    t3 = time.perf_counter()
    t["solve_batch"] += t1 - t0; t["step_batch"] += t2 - t1; t["respawn"] += t3 - t2
    t["done_mask"] = t.get("done_mask", 0.0) + t2b - t2; t["x_assign"] = t.get("x_assign", 0.0) + t2c - t2b
    t["reset_scenarios"] = t.get("reset_scenarios", 0.0) + t3 - t2c; t["n_done"] = t.get("n_done", 0) + int(done.sum()) / 1e3 * K
tot = time.perf_counter() - T0
print({k: round(1e3 * v / K, 3) for k, v in t.items()}, "total ms/step", round(1e3 * tot / K, 3))

import os, sys, json
sys.path.insert(0, "/root/repo")
import numpy as np, torch
import bench
from gp_mpc_rocket_landing_b200.monte_carlo import BatchedMonteCarlo
class A:
    batch, sqp_iters, qp_max_iter, convex_rows = 1024, 1, 1000, True
dyn, gp, ctl, _ = bench.build_problem(A, 0)
mc = BatchedMonteCarlo(dyn, ctl, drag_cda=0.25, respawn=False, target_mode="per_step", on_solve_failure="continue")
mc.reset(first=0, count=A.batch, seed=42)
tan = np.tan(np.deg2rad(30))
first_fail = {}
for k in range(200):
    x = mc.x.cpu().numpy().copy(); act = mc.active.cpu().numpy().astype(bool)
    sol = mc.step()
    qs = sol.qp_status.cpu().numpy(); qi = sol.qp_iterations.cpu().numpy()
    bad = act & (qs != 1)
    for b in np.nonzero(bad)[0]:
        if b not in first_fail:
            first_fail[b] = (k, qs[b], x[b].round(3).tolist())
    if k % 10 == 0:
        lat = np.hypot(x[act, 2], x[act, 3]); alt = x[act, 1]
        print(k, "active", act.sum(), "bad", bad.sum(), "statuses", dict(zip(*np.unique(qs[act], return_counts=True))), "mean it", qi[act].mean().round(1) if act.any() else 0,
              "max lat/alt ratio", (lat / np.maximum(alt, 1e-3)).max().round(3) if act.any() else 0, "alt med", np.median(alt).round(2) if act.any() else 0)
    if not act.any(): break
ff = sorted(first_fail.items(), key=lambda kv: kv[1][0])
print("first failures:", len(ff))
for b, v in ff[:15]:
    lat = np.hypot(v[2][2], v[2][3])
    print(b, v, "lat/alt", round(lat / max(v[2][1], 1e-3), 3))
c = mc.counts.cpu().numpy(); print(c)

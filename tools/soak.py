"""GPU soak run (not part of the product): STEPS closed-loop steps of the bench workload with respawn; prints outcome counts,
NaN checks and the step-time drift."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench


class A:
    batch, sqp_iters, qp_max_iter = int(os.environ.get("B", 4096)), 1, 1000
    convex_rows = bool(os.environ.get("CONVEX"))


dyn, gp, ctl, mc = bench.build_problem(A, 0)
n = int(os.environ.get("STEPS", 2000))
ts = []
t0 = time.time()
for k in range(n):
    mc.step()
    if k % 250 == 249:
        torch.cuda.synchronize()
        ts.append(time.time() - t0)
        t0 = time.time()
torch.cuda.synchronize()
c = mc.counts.cpu().numpy()
x = mc.x
out = {"steps": n, "solves": int(c[0]), "outcomes": {k: int(v) for k, v in zip(("success", "crash", "fuel_exhausted", "constraint_violation", "timeout", "divergence"), c[1:7])},
       "admm_iters_per_solve": float(c[7]) / max(1.0, float(c[0])), "nan_in_state": bool(torch.isnan(x).any().item()),
       "inf_in_state": bool(torch.isinf(x).any().item()), "active": int(mc.active.sum().item()),
       "seconds_per_250_steps": [round(t, 3) for t in ts],
       "qp_status_hist_last_step": {int(k): int(v) for k, v in zip(*torch.unique(ctl._batch.ws.status, return_counts=True))}}
print(json.dumps(out))

"""GPU run (not part of the product): closed-loop outcomes of the bench workload with and without the convex thrust-cone /
glide-slope rows in the QP (GPMPCConfig.convex_constraints), same scenarios, no respawn, until every run has ended or
`STEPS` control steps.  Prints one JSON line per mode."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from gp_mpc_rocket_landing_b200.monte_carlo import BatchedMonteCarlo

NAMES = ["success", "crash", "fuel_exhausted", "constraint_violation", "timeout", "divergence"]
MODES = [(False, None)] + [(True, float(g)) for g in os.environ.get("GAMMAS", "30").split(",")]
for convex, gamma in MODES:
    class A:
        batch, sqp_iters, qp_max_iter, convex_rows, gamma_gs = int(os.environ.get("B", 4096)), 1, 1000, convex, gamma
    dyn, gp, ctl, _ = bench.build_problem(A, 0)
    mc = BatchedMonteCarlo(dyn, ctl, drag_cda=0.25, respawn=False, target_mode=os.environ.get("TARGET", "per_step"), on_solve_failure="continue")
    mc.reset(first=0, count=A.batch, seed=42)
    t0 = time.time()
    steps = 0
    it_sum = 0
    for k in range(int(os.environ.get("STEPS", 320))):
        mc.step()
        steps += 1
        if k % 20 == 19 and int(mc.active.sum().item()) == 0:
            break
    torch.cuda.synchronize()
    c = mc.counts.cpu().numpy()
    xf = mc.x_final.cpu().numpy() if hasattr(mc, "x_final") else None
    out = {"convex_rows": convex, "gamma_gs_deg": gamma, "target_mode": mc.target_mode, "scenarios": A.batch, "steps": steps, "wall_s": time.time() - t0,
           "outcomes": {n: int(c[1 + i]) for i, n in enumerate(NAMES)}, "still_active": int(mc.active.sum().item()),
           "solves": int(c[0]), "mean_admm_iters_per_solve": float(c[7]) / max(1.0, float(c[0]))}
    out["success_rate"] = out["outcomes"]["success"] / max(1, sum(out["outcomes"].values()))
    print(json.dumps(out), flush=True)
    del dyn, gp, ctl, mc
    torch.cuda.empty_cache()

"""GPU A/B of the ADMM kernel generations on realistic QP data (not part of the product): the QPs of a closed-loop step are
solved from the same warm start by kernel_variant 2 (stage-parallel, second generation) and 0 (third generation: light
iterations + cyclic reduction); prints agreement (status, iterations, x, y, rho) and timings.  DOF=6 for the 6-DoF shape."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from gp_mpc_rocket_landing_b200 import _lib
from gp_mpc_rocket_landing_b200.mpc import engine as me

DOF = int(os.environ.get("DOF", 3))
if DOF == 3:
    import bench
    class A: pass
    a = A(); a.batch = int(os.environ.get("B", 4096)); a.sqp_iters = 1; a.qp_max_iter = 1000
    dyn, gp, ctl, mc = bench.build_problem(a, 0)
    for k in range(int(os.environ.get("STEPS", 5))):
        mc.step()
    st = ctl._batch
    qp = ctl._qp
else:
    import gp_mpc_rocket_landing_b200.gp as gpm
    import gp_mpc_rocket_landing_b200.mpc as mpc
    from gp_mpc_rocket_landing_b200 import dynamics, workloads
    B = int(os.environ.get("B", 1024))
    X, U, Dv, Dw = workloads.residual_dataset_6dof(1000, seed=0)
    sg = gpm.StructuredRocketGP(gpm.StructuredGPConfig(n_inducing=100, use_sparse=True, noise_variance=1e-3))
    np.random.seed(0)
    sg.add_data(X, U, Dv, Dw); sg.fit()
    dyn = dynamics.Rocket6DoFDynamics()
    ctl = mpc.GPMPC6DoF(dyn, sg, mpc.GPMPC6DoFConfig(N=30, max_sqp_iter=1, use_gp_uncertainty=False))
    X0, XT = workloads.sample_initial_conditions_6dof(0, B)
    x = gpm.engine.to_dev(X0); xt = gpm.engine.to_dev(XT)
    for _ in range(3):
        s = ctl.solve_batch(x, xt)
        x = dyn.step_batch(x, s.u0.contiguous(), 0.1)
    st = ctl._batch
    qp = ctl._qp
ws = st.ws
x0, y0, r0 = ws.x.clone(), ws.y.clone(), ws.rho.clone()


def solve(variant, **kw):
    S = _lib.AdmmSettings.reference_defaults(max_iter=1000, warm_x=1, warm_y=1, rho_from_input=1, kernel_variant=variant)
    for k, v in kw.items():
        setattr(S, k, v)
    ws.x.copy_(x0); ws.y.copy_(y0); ws.rho.copy_(r0)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); me.admm_solve(qp, st.AB, st.beq, st.q, S, ws); e1.record(); torch.cuda.synchronize()
    return dict(ms=e0.elapsed_time(e1), x=ws.x.cpu().numpy().copy(), y=ws.y.cpu().numpy().copy(), rho=ws.rho.cpu().numpy().copy(),
                status=ws.status.cpu().numpy().copy(), iters=ws.iters.cpu().numpy().copy(), obj=ws.obj.cpu().numpy().copy(),
                pri=ws.pri.cpu().numpy().copy(), dua=ws.dua.cpu().numpy().copy())


out = {"dof": DOF, "B": int(ws.x.shape[0])}
for variant in (2, 3):
    solve(variant)          # warm-up (attribute setup)
a2, a3 = solve(2), solve(3)
out["full_solve_ms"] = {"gen2": a2["ms"], "gen3": a3["ms"]}
out["status_equal"] = float(np.mean(a2["status"] == a3["status"]))
out["iters_equal"] = float(np.mean(a2["iters"] == a3["iters"]))
out["iters_mean"] = [float(a2["iters"].mean()), float(a3["iters"].mean())]
same = (a2["status"] == a3["status"]) & (a2["iters"] == a3["iters"])
scale = np.maximum(1.0, np.abs(a2["x"]).max(axis=1))
out["x_maxdiff_same_iters"] = float((np.abs(a2["x"] - a3["x"]).max(axis=1) / scale)[same].max()) if same.any() else None
ys = np.maximum(1.0, np.abs(a2["y"]).max(axis=1))
out["y_maxdiff_same_iters"] = float((np.abs(a2["y"] - a3["y"]).max(axis=1) / ys)[same].max()) if same.any() else None
out["rho_equal_same_iters"] = float(np.mean(np.isclose(a2["rho"], a3["rho"], rtol=1e-6)[same])) if same.any() else None
out["status_hist_gen3"] = {int(k): int((a3["status"] == k).sum()) for k in np.unique(a3["status"])}
bad = np.where(~same)[0][:5]
out["mismatch_examples"] = [(int(i), int(a2["status"][i]), int(a3["status"][i]), int(a2["iters"][i]), int(a3["iters"][i]),
                             float(a2["pri"][i]), float(a3["pri"][i])) for i in bad]
# fixed iteration counts, checks and adaptation off: per-iteration cost and set-up cost
t = {}
for variant, name in ((2, "gen2"), (3, "gen3")):
    t[name] = {n: min(solve(variant, max_iter=n, check_every=0, adaptive_rho=0)["ms"] for _ in range(2)) for n in (0, 1, 50, 100, 200)}
    t[name]["per_iteration_us_batch"] = (t[name][200] - t[name][100]) * 10
out["fixed_iters_ms"] = t
f2 = solve(2, max_iter=100, check_every=0, adaptive_rho=0); f3 = solve(3, max_iter=100, check_every=0, adaptive_rho=0)
out["x_maxdiff_100_fixed_iters"] = float((np.abs(f2["x"] - f3["x"]).max(axis=1) / scale).max())
out["y_maxdiff_100_fixed_iters"] = float((np.abs(f2["y"] - f3["y"]).max(axis=1) / ys).max())
print(json.dumps(out))
# cost of the checking iterations and of refactorisations: 100 iterations that cannot terminate (eps 1e-14)
t2 = {}
for variant, name in ((2, "gen2"), (3, "gen3")):
    t2[name] = {
        "no_checks": min(solve(variant, max_iter=100, check_every=0, adaptive_rho=0)["ms"] for _ in range(2)),
        "checks_every_25": min(solve(variant, max_iter=100, check_every=25, adaptive_rho=0, eps_abs=1e-14, eps_rel=1e-14)["ms"] for _ in range(2)),
        "checks_every_5": min(solve(variant, max_iter=100, check_every=5, adaptive_rho=0, eps_abs=1e-14, eps_rel=1e-14)["ms"] for _ in range(2)),
        "checks_25_adaptive_tol_1.0001": min(solve(variant, max_iter=100, check_every=25, adaptive_rho=1, adaptive_rho_tolerance=1.0001, eps_abs=1e-14, eps_rel=1e-14)["ms"] for _ in range(2)),
    }
print(json.dumps({"check_and_refactor_cost_ms": t2}))

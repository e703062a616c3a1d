"""Diagnostic (not part of the product): how much of the ADMM launch is tail?  Collects the per-QP iteration counts of a
few closed-loop steps and simulates the CTA schedule (greedy, `slots` QPs resident) for three launch orders: scenario
index (what the kernel does today), sorted by the previous step's iteration count (known before the launch), and sorted
by the true count (lower bound for any order)."""
import heapq, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch


def makespan(iters, slots, setup):
    h = [0.0] * slots
    heapq.heapify(h)
    for it in iters:
        t = heapq.heappop(h)
        heapq.heappush(h, t + setup + it)
    return max(h)


def report(name, hist, slots, setup):
    out = []
    for prev, cur in zip(hist[:-1], hist[1:]):
        idx = makespan(cur, slots, setup)
        by_prev = makespan(cur[np.argsort(-prev, kind="stable")], slots, setup)
        best = makespan(np.sort(cur)[::-1], slots, setup)
        # exact-continuation phases: every QP runs at most cap iterations per launch, unfinished ones relaunch (set-up again)
        ph = 0.0
        rest = cur.copy()
        for cap in (150, 350, 1000):
            live = rest[rest > 0]
            if len(live) == 0:
                break
            ph += makespan(np.minimum(live, cap), slots, setup)
            rest = np.maximum(rest - cap, 0)
        out.append((idx, by_prev, best, float(cur.sum() + setup * len(cur)) / slots, int(cur.max()), ph))
    a = np.array(out)
    q = np.quantile(np.concatenate(hist), [0.5, 0.9, 0.99, 0.999])
    print(json.dumps({"case": name, "iteration_quantiles_50_90_99_999": q.tolist(),
                      "share_at_max": float((np.concatenate(hist) >= 1000).mean()),
                      "corr_prev_cur": float(np.mean([np.corrcoef(p_, c_)[0, 1] for p_, c_ in zip(hist[:-1], hist[1:])])),
                      "makespan_three_capped_phases_150_350_1000": a[:, 5].mean()}))
    print(json.dumps({"case": name, "slots": slots, "setup_iters_equiv": setup,
                      "makespan_index_order": a[:, 0].mean(), "makespan_sorted_by_previous": a[:, 1].mean(),
                      "makespan_sorted_by_true": a[:, 2].mean(), "perfect_balance": a[:, 3].mean(),
                      "max_iters_mean": a[:, 4].mean()}))


def three_dof():
    import bench
    class A: pass
    a = A(); a.batch = 4096; a.sqp_iters = 1; a.qp_max_iter = 1000
    dyn, gp, ctl, mc = bench.build_problem(a, 0)
    hist, cold = [], []
    for k in range(80):
        hw = ctl._batch.has_warm.cpu().numpy().copy() if ctl._batch is not None else np.zeros(4096, dtype=np.int32)
        mc.step()
        if k >= 20:
            hist.append(ctl._batch.ws.iters.cpu().numpy().astype(np.float64).copy())
            cold.append(hw == 0)
    report("3dof B=4096", hist, 444, 15.0)   # set-up 0.6 ms ~ 15 iterations of 39.6 us
    # cold-started QPs (respawned scenarios: no warm start, default rho) first, the rest in index order
    res = []
    for cur, c in zip(hist, cold):
        order = np.concatenate([np.nonzero(c)[0], np.nonzero(~c)[0]])
        res.append((makespan(cur, 444, 15.0), makespan(cur[order], 444, 15.0), makespan(np.sort(cur)[::-1], 444, 15.0),
                    float(cur.sum() + 15.0 * len(cur)) / 444, int(c.sum()), float(cur[c].mean()) if c.any() else 0.0,
                    float(cur[~c].mean())))
    a = np.array(res)
    print(json.dumps({"case": "3dof cold-first order", "makespan_index_order": a[:, 0].mean(), "makespan_cold_first": a[:, 1].mean(),
                      "makespan_sorted_by_true": a[:, 2].mean(), "perfect_balance": a[:, 3].mean(),
                      "cold_qps_per_step": a[:, 4].mean(), "mean_iters_cold": a[:, 5].mean(), "mean_iters_warm": a[:, 6].mean()}))


def six_dof():
    import gp_mpc_rocket_landing_b200.gp as gp
    import gp_mpc_rocket_landing_b200.mpc as mpc
    from gp_mpc_rocket_landing_b200 import dynamics, workloads
    X, U, Dv, Dw = workloads.residual_dataset_6dof(1000, seed=0)
    sg = gp.StructuredRocketGP(gp.StructuredGPConfig(n_inducing=100, use_sparse=True, noise_variance=1e-3))
    sg.add_data(X, U, Dv, Dw); sg.fit()
    dyn = dynamics.Rocket6DoFDynamics()
    ctl = mpc.GPMPC6DoF(dyn, sg, mpc.GPMPC6DoFConfig(N=30, max_sqp_iter=1, use_gp_uncertainty=False))
    X0, XT = workloads.sample_initial_conditions_6dof(0, 1024)
    x = gp.engine.to_dev(X0); xt = gp.engine.to_dev(XT)
    hist = []
    for k in range(23):
        s = ctl.solve_batch(x, xt)
        x = dyn.step_batch(x, s.u0.contiguous(), 0.1)
        if k >= 3:
            hist.append(ctl._batch.ws.iters.cpu().numpy().astype(np.float64).copy())
    report("6dof B=1024", hist, 148, 37.0)   # set-up 1.28 ms / 34.1 us per iteration (generation 3)


if os.environ.get("ONLY6") is None:
    three_dof()
if os.environ.get("ONLY3") is None:
    six_dof()

"""BASELINE config 4: 6-DoF quaternion GP-MPC, N = 30, sparse structured GP (100 inducing points, 1000 residual points),
1024 scenarios, closed loop (plant = the same 14-state model), S SQP iterations per solve.  Prints one JSON line:
solves/s and per-phase times from CUDA events; --cpu adds the oracle (oracle/mpc6.py + dense OSQP restatement) timed on
ONE solve of one scenario."""
import argparse, json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


class _A:
    pass


def measure(batch=1024, steps=20, warmup=3, sqp_iters=1, inducing=100, cpu=False, variant=0):
    a = _A()
    a.batch, a.steps, a.warmup, a.sqp_iters, a.inducing, a.cpu = batch, steps, warmup, sqp_iters, inducing, cpu
    import torch
    import gp_mpc_rocket_landing_b200.gp as gp
    import gp_mpc_rocket_landing_b200.mpc as mpc
    from gp_mpc_rocket_landing_b200 import dynamics, workloads
    ge = gp.engine
    X, U, Dv, Dw = workloads.residual_dataset_6dof(1000, seed=0)
    sg = gp.StructuredRocketGP(gp.StructuredGPConfig(n_inducing=a.inducing, use_sparse=True, noise_variance=1e-3))
    np.random.seed(0)   # the inducing points are drawn from the global numpy RNG (sparse_gp.py:122-148)
    sg.add_data(X, U, Dv, Dw); sg.fit()
    dyn = dynamics.Rocket6DoFDynamics()
    ctl = mpc.GPMPC6DoF(dyn, sg, mpc.GPMPC6DoFConfig(N=30, max_sqp_iter=a.sqp_iters, qp_kernel_variant=variant))
    X0, XT = workloads.sample_initial_conditions_6dof(0, a.batch)
    x = ge.to_dev(X0); xt = ge.to_dev(XT)
    its = []
    def step():
        nonlocal x
        s = ctl.solve_batch(x, xt)
        x = dyn.step_batch(x, s.u0.contiguous(), 0.1)
        return s
    for _ in range(a.warmup):
        step()
    torch.cuda.synchronize()
    from gp_mpc_rocket_landing_b200.mpc import engine as me
    ctl.timers = me.PhaseTimers()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        s = step()
        its.append(float(s.qp_iterations.double().mean().item()))
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / a.steps
    line = {"config": "BASELINE configs[3]: 6-DoF quaternion GP-MPC N=30, sparse structured GP (M=%d, n=1000), closed loop" % a.inducing,
            "scenarios": a.batch, "sqp_iters_per_solve": a.sqp_iters, "ms_per_step": ms,
            "solves_per_s": a.batch / (ms * 1e-3), "mean_admm_iters_per_solve": float(np.mean(its)),
            "qp": "n=524 m=958, admm_stage_kernel<14,3,30,3,256,1> (one QP per SM, 220 KB shared memory)",
            "solved_fraction_last_step": float((s.qp_status == 1).double().mean().item()), "dtype": "f64"}
    line["phase_ms_per_step"] = {k: round(v[0] / a.steps, 3) for k, v in ctl.timers.totals().items()}
    if a.cpu:
        from oracle import gp as ogp, mpc as ompc, mpc6 as o6
        p = o6.Rocket6DoFParams()
        Zv = ogp.features_batch(ogp.features_6dof_trans, X, U); Zw = ogp.features_batch(ogp.features_6dof_rot, X, U)
        fv = [ogp.sparse_fit(0, Zv, Dv[:, i], np.asarray(sg.gp_v.gps[0]._Z), np.ones(13), 1.0, 1e-3) for i in range(3)]
        fw = [ogp.sparse_fit(0, Zw, Dw[:, i], np.asarray(sg.gp_omega.gps[0]._Z), np.ones(12), 1.0, 1e-3) for i in range(3)]
        def predict(xx, uu):
            zv = ogp.features_6dof_trans(xx, uu)[None]; zw = ogp.features_6dof_rot(xx, uu)[None]
            return (np.array([ogp.sparse_predict(f, zv)[0][0] for f in fv]), np.array([ogp.sparse_predict(f, zw)[0][0] for f in fw]))
        t0 = time.perf_counter()
        U0 = np.zeros((30, 3)); U0[:, 0] = X0[0, 0]
        Xo, Do = o6.rollout_gp(X0[0], U0, 0.1, p, predict)
        aff = np.zeros((30, 14)); aff[:, 4:7] = Do[:, :3] * 0.1; aff[:, 11:14] = Do[:, 3:] * 0.1
        ABo, rhs = o6.stage_blocks(30, Xo, U0, 0.1, p, aff)
        A, l, u = o6.dense_constraints(30, ABo, X0[0], rhs)
        Pd, q = o6.build_cost(30, np.tile(XT[0], (31, 1)))
        r = ompc.OSQPOracle(Pd, q, A, l, u, ompc.OSQPSettings(max_iter=1000)).solve()
        dt = time.perf_counter() - t0
        line["cpu_baseline"] = {"value": 1.0 / dt, "unit": "solves/s", "cores": os.cpu_count(), "kind": "port",
                                "sample": "1 solve of 1 scenario (numpy/scipy oracle, dense KKT), %.2f s, %d ADMM iterations" % (dt, r.iters)}
    return line


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=1024)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--sqp-iters", type=int, default=1)
    ap.add_argument("--inducing", type=int, default=100)
    ap.add_argument("--cpu", action="store_true")
    ap.add_argument("--variant", type=int, default=0)
    a = ap.parse_args()
    print(json.dumps(measure(a.batch, a.steps, a.warmup, a.sqp_iters, a.inducing, a.cpu, a.variant)), flush=True)


if __name__ == "__main__":
    main()

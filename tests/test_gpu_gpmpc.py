"""GPU parity: the composed 3-DoF GP-MPC path (`GPMPC.solve` / `solve_batch`) and the RTI path (`OSQPRTIMPC.step`)
through the public API against the oracle's restatement of gp_mpc.py:229-369 / osqp_rti.py:403-599."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def setup():
    from gp_mpc_rocket_landing_b200 import workloads
    from gp_mpc_rocket_landing_b200.dynamics import Rocket3DoFDynamics
    from gp_mpc_rocket_landing_b200.gp import Simple3DoFGP
    from oracle import gp as ogp
    X, U, D = workloads.residual_dataset(300, seed=0)
    gp = Simple3DoFGP(noise_variance=1e-2, use_sparse=False)
    gp.add_data(X, U, D); gp.fit()
    o = ogp.Simple3DoFGPOracle(noise_variance=1e-2)
    o.add_data(X, U, D); o.fit()
    return Rocket3DoFDynamics(), gp, o


def test_simple3dof_predict_api(setup):
    _, gp, o = setup
    from gp_mpc_rocket_landing_b200 import workloads
    X0, _ = workloads.sample_initial_conditions(0, 5)
    u = np.array([2.0, 0.1, -0.2])
    for b in range(5):
        m, v = gp.predict(X0[b], u)
        mo, vo = o.predict(X0[b], u)
        np.testing.assert_allclose(m, mo, rtol=1e-9, atol=1e-12)   # GP mean/variance: 1e-9 relative (north_star)
        np.testing.assert_allclose(v, vo, rtol=1e-9)
    mb, vb = gp.predict_batch(X0, np.tile(u, (5, 1)))
    m4, v4 = gp.predict(X0[4], u)
    assert np.array_equal(mb[4], m4) and np.array_equal(vb[4], v4)   # batch = loop of singles, bitwise
    empty = type(gp)(use_sparse=False)
    m, v = empty.predict(X0[0], u)
    assert np.array_equal(m, np.zeros(3)) and np.array_equal(v, np.ones(3) * 0.1)


@pytest.mark.parametrize("use_gp", [True, False])
def test_gpmpc_solve_batch_matches_oracle(setup, use_gp):
    from gp_mpc_rocket_landing_b200 import workloads
    from gp_mpc_rocket_landing_b200.mpc import GPMPC, GPMPCConfig
    from oracle import mpc as ompc
    dyn, gp, o = setup
    N, S, B = 20, 3, 5
    X0, XT = workloads.sample_initial_conditions(0, B)
    cfg = GPMPCConfig(N=N, dt=0.1, max_sqp_iter=S, qp_max_iter=1000, use_gp_mean=use_gp, use_gp_uncertainty=use_gp)
    mpc = GPMPC(dyn, gp if use_gp else None, cfg)
    mpc.setup()
    oracles = [ompc.GPMPCOracle(o.predict, N=N, dt=0.1, settings=ompc.OSQPSettings(max_iter=1000), max_sqp_iter=S,
                                use_gp_mean=use_gp) for _ in range(B)]
    p = ompc.Rocket3DoFParams()
    x = X0.copy()
    for step in range(3):   # closed loop: exercises the warm starts (U unshifted, duals kept when solved)
        sol = mpc.solve_batch(x, XT)
        covs = mpc.last_covariances()
        assert (covs is not None) == use_gp
        for b in range(B):
            r = oracles[b].solve(x[b], XT[b])
            assert int(sol.iterations[b]) == r.iterations, (step, b)
            assert bool(sol.success[b]) == r.success
            assert int(sol.qp_status[b]) == r.qp_status[-1]
            assert int(sol.qp_iterations[b]) == sum(r.qp_iters), (step, b, sol.qp_iterations[b], r.qp_iters)
            # QP u0 within the OSQP tolerance eps_abs + eps_rel*|u| of the CPU restatement (observed ~1e-7)
            assert np.max(np.abs(sol.u0[b] - r.U_opt[0])) < 1e-4 + 1e-4 * np.max(np.abs(r.U_opt[0]))
            np.testing.assert_allclose(sol.U_opt[b], r.U_opt, rtol=0, atol=2e-5)
            np.testing.assert_allclose(sol.X_opt[b], r.X_opt, rtol=0, atol=2e-5)
            if np.isfinite(r.cost):
                assert sol.cost[b] == pytest.approx(r.cost, rel=1e-5)
            if use_gp:
                np.testing.assert_allclose(covs[b], r.covariances, rtol=1e-5, atol=1e-16)
        x = np.stack([ompc.step_3dof(x[b], sol.u0[b], 0.1, p) for b in range(B)])
    # single-scenario API = B=1 view
    mpc1 = GPMPC(dyn, gp if use_gp else None, cfg)
    s1 = mpc1.solve(X0[2], XT[2])
    mpc2 = GPMPC(dyn, gp if use_gp else None, cfg)
    sb = mpc2.solve_batch(X0, XT)
    np.testing.assert_array_equal(s1.U_opt, sb.U_opt[2])
    assert s1.u0.shape == (3,) and s1.status in ("Converged", "Max iterations")


def test_osqp_rti_step_matches_oracle(setup):
    from gp_mpc_rocket_landing_b200 import workloads
    from gp_mpc_rocket_landing_b200.mpc import FastRTI3DoF, OSQPRTIConfig, OSQPRTIMPC
    from oracle import mpc as ompc
    dyn, _, _ = setup
    p = ompc.Rocket3DoFParams()
    X0, XT = workloads.sample_initial_conditions(0, 4)
    for cls, lin, max_iter in ((FastRTI3DoF, "analytic", 50), (FastRTI3DoF, "analytic", 500), (OSQPRTIMPC, "fd", 500)):
        cfg = OSQPRTIConfig(N=15, osqp_max_iter=max_iter)
        ctl = cls(dyn, cfg)
        ctl.initialize_batch(X0, XT)
        ors = []
        for b in range(4):
            oc = ompc.OSQPRTIOracle(p, N=15, dt=0.1, lin=lin, settings=ompc.OSQPSettings(max_iter=max_iter))
            oc.initialize(X0[b], XT[b]); ors.append(oc)
        x = X0.copy()
        for step in range(4):
            sol = ctl.step_batch(x)
            for b in range(4):
                r = ors[b].step(x[b])
                assert bool(sol.success[b]) == r.success and int(sol.status[b]) == r.status, (cls, step, b)
                assert int(sol.osqp_iterations[b]) == r.osqp_iterations
                np.testing.assert_allclose(sol.u0[b], r.u0, rtol=0, atol=2e-5)
                np.testing.assert_allclose(sol.X_opt[b], r.X_opt, rtol=0, atol=5e-5)
            x = np.stack([ompc.step_3dof(x[b], sol.u0[b], 0.1, p) for b in range(4)])
    # single-controller API
    one = FastRTI3DoF(dyn, OSQPRTIConfig(N=15, osqp_max_iter=500))
    one.initialize(X0[0], XT[0])
    s = one.step(X0[0])
    assert s.u0.shape == (3,) and s.X_opt.shape == (16, 7) and isinstance(s.success, bool)


def test_rti_closed_loop_matches_the_reference_orchestration(setup):
    """a22 on the device: `FastRTI3DoF.step_batch` fed the states of tests/golden/rti_loop.npz -- the reference's own
    FastRTI3DoF.step loop around a solver-backed `osqp` stand-in -- takes the same branch at every step (max-iter fall-back
    on the first step at max_iter 50, pattern growth 654 -> 734 non-zeros absorbed by the fixed superset pattern, primal
    infeasible README state falling back at every step), the same ADMM iteration counts, and returns the same controls /
    trajectories to the ADMM tolerance."""
    from conftest import load_golden
    from gp_mpc_rocket_landing_b200.mpc import FastRTI3DoF, OSQPRTIConfig
    dyn, _, _ = setup
    g = load_golden("rti_loop.npz")
    for tags in (("vertical_it50",), ("lateral_it1000", "readme_infeasible_it1000")):
        mi = int(g[f"{tags[0]}_max_iter"])
        ctl = FastRTI3DoF(dyn, OSQPRTIConfig(N=20, osqp_max_iter=mi))
        ctl.initialize_batch(np.stack([g[f"{t}_x0"] for t in tags]), np.stack([g[f"{t}_xt"] for t in tags]))
        for k in range(6):
            sol = ctl.step_batch(np.stack([g[f"{t}_x"][k] for t in tags]))
            for b, t in enumerate(tags):
                assert bool(sol.success[b]) == bool(g[f"{t}_success"][k]), (t, k)
                assert int(sol.osqp_iterations[b]) == int(g[f"{t}_iters"][k]), (t, k)
                np.testing.assert_allclose(sol.u0[b], g[f"{t}_u0"][k], rtol=0, atol=2e-5)
                np.testing.assert_allclose(sol.X_opt[b], g[f"{t}_X_opt"][k], rtol=0, atol=5e-5)
                np.testing.assert_allclose(sol.U_opt[b], g[f"{t}_U_opt"][k], rtol=0, atol=5e-5)
                assert np.isinf(sol.cost[b]) == np.isinf(g[f"{t}_cost"][k])
                if np.isfinite(sol.cost[b]):
                    assert sol.cost[b] == pytest.approx(float(g[f"{t}_cost"][k]), rel=1e-4, abs=1e-3)


def test_gpmpc_solve_matches_the_reference_sqp_skeleton():
    """a17 on the device: tests/golden/gpmpc_skeleton.npz is the reference's own `GPMPC.solve` loop (only `_solve_qp` replaced
    by the OSQP-form QP, see oracle/gen_golden.py) over three consecutive control steps with the warm start carried.  The
    device controller with the reference's 10 SQP iterations takes the same path: no convergence (`success` False, 10
    iterations), the same total ADMM iteration count per solve, the same trajectories to the ADMM tolerance."""
    from conftest import load_golden
    from gp_mpc_rocket_landing_b200.dynamics import Rocket3DoFDynamics
    from gp_mpc_rocket_landing_b200.gp import Simple3DoFGP
    from gp_mpc_rocket_landing_b200.mpc import GPMPC, GPMPCConfig
    g = load_golden("gpmpc_skeleton.npz")
    gp = Simple3DoFGP(noise_variance=float(g["noise"]), use_sparse=False)
    gp.add_data(g["X"], g["U"], g["D"]); gp.fit()
    ctl = GPMPC(Rocket3DoFDynamics(), gp, GPMPCConfig(N=20, dt=0.1, max_sqp_iter=10, qp_max_iter=1000))
    ctl.setup()
    k0 = 0
    for s in range(3):
        sol = ctl.solve(g["x"][s], g["xt"][s], U_ref=g["U_ref"])
        n = int(g["qp_n"][s])
        assert sol.success == bool(g["success"][s]) and sol.iterations == int(g["iterations"][s])
        assert sol.status == "Max iterations"
        b = ctl._single
        assert int(b.qp_iters[0].item()) == int(g["qp_iters"][k0:k0 + n].sum())
        k0 += n
        np.testing.assert_allclose(sol.U_opt, g["U_opt"][s], rtol=0, atol=2e-5)
        np.testing.assert_allclose(sol.X_opt, g["X_opt"][s], rtol=0, atol=2e-5)
        assert sol.cost == pytest.approx(float(g["cost"][s]), rel=1e-5)


def test_per_stage_bound_tightening_extension(setup):
    """GPMPCConfig.tighten_state_bounds (extension; the reference prepares the tightening from Sigma_k and applies it at
    k = 0 only, gp_mpc.py:177-215,414): per-scenario boxes lb + kappa sqrt(diag Sigma_k), ub - kappa sqrt(diag Sigma_k) for
    the states of stages k >= 1, read by the ADMM kernel through GpmpcAdmmSettings.bound_stride.  Checked: the boxes
    against the covariances; the device solution against the dense OSQP restatement of the same QP with the same boxes."""
    from scipy.stats import norm
    from gp_mpc_rocket_landing_b200 import workloads
    from gp_mpc_rocket_landing_b200.mpc import GPMPC, GPMPCConfig, engine as me
    from oracle import mpc as ompc
    dyn, gp, _ = setup
    N, B = 20, 4
    X0, XT = workloads.sample_initial_conditions(0, B)
    X0[:, 4] = -49.5      # descent rate next to its -50 bound: the tightened box binds on the first stages
    cfg = GPMPCConfig(N=N, dt=0.1, max_sqp_iter=1, qp_max_iter=4000, tighten_state_bounds=True, confidence_level=0.99)
    mpc = GPMPC(dyn, gp, cfg); mpc.setup()
    sol = mpc.solve_batch(X0, XT)
    st = mpc._batch
    covs = mpc.last_covariances()
    kappa = norm.ppf(0.99)
    lb0, ub0 = mpc._qp.lb.cpu().numpy(), mpc._qp.ub.cpu().numpy()
    lbt, ubt = st.lb_t.cpu().numpy(), st.ub_t.cpu().numpy()
    for b in range(B):
        for k in range(N + 1):
            sl = slice(k * 10, k * 10 + 7)
            back = kappa * np.sqrt(np.diag(covs[b, k])) if k >= 1 else np.zeros(7)
            fin_l, fin_u = lb0[sl] > -1e20, ub0[sl] < 1e20
            np.testing.assert_allclose(lbt[b, sl][fin_l], (lb0[sl] + back)[fin_l], rtol=1e-14)
            np.testing.assert_allclose(ubt[b, sl][fin_u], (ub0[sl] - back)[fin_u], rtol=1e-14)
            assert np.array_equal(lbt[b, sl][~fin_l], lb0[sl][~fin_l])
            if k < N:
                assert np.array_equal(lbt[b, k * 10 + 7:k * 10 + 10], lb0[k * 10 + 7:k * 10 + 10])   # controls untouched
    assert np.all(sol.qp_status == 1)
    AB, beq, q = st.AB.cpu().numpy(), st.beq.cpu().numpy(), st.q.cpu().numpy()
    Pd = mpc._qp.Pdiag.cpu().numpy()
    z = st.ws.x.cpu().numpy()
    for b in range(2):
        A, l, u = me.dense_constraints(N, AB[b], beq[b], lbt[b], ubt[b])
        r = ompc.OSQPOracle(Pd, q[b], A, l, u, ompc.OSQPSettings(max_iter=4000)).solve()
        assert r.status == 1
        obj_d = 0.5 * z[b] @ (Pd * z[b]) + q[b] @ z[b]
        obj_o = 0.5 * r.x @ (Pd * r.x) + q[b] @ r.x
        assert abs(obj_d - obj_o) < 1e-3 * max(1.0, abs(obj_o))
        assert np.max(np.abs(z[b] - r.x)) < 5e-3 * max(1.0, np.max(np.abs(r.x)))      # two ADMM runs at eps = 1e-4
        assert np.all(z[b] >= lbt[b] - 1e-2) and np.all(z[b] <= ubt[b] + 1e-2)   # ADMM primal tolerance (eps_rel |z|)
        # the tightened box is active: the velocity of stage 1 sits above the ORIGINAL bound by about the back-off
        assert z[b][10 + 4] >= lb0[14] + 0.5 * kappa * np.sqrt(covs[b, 1][4, 4])
    # off (default) = the reference's fixed box
    ref = GPMPC(dyn, gp, GPMPCConfig(N=N, dt=0.1, max_sqp_iter=1, qp_max_iter=4000)); ref.setup()
    ref.solve_batch(X0, XT)
    assert ref._batch.lb_t is None

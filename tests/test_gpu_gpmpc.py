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

"""GPU parity: the 6-DoF quaternion GP-MPC step (BASELINE configs[3]: 14 states, N = 30, sparse structured GP) through
the C-ABI against oracle/mpc6.py (CPU restatement of the in-repo model statement; dynamics parity is unpinned -- see
that file's header) and the dense OSQP restatement.  Tolerances are written next to each check."""
import ctypes as C

import numpy as np
import pytest

from conftest import relerr

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def env():
    import gp_mpc_rocket_landing_b200.gp as gp
    import gp_mpc_rocket_landing_b200.mpc as mpc
    from gp_mpc_rocket_landing_b200 import dynamics, workloads
    gp.engine.torch_()
    return gp, mpc, dynamics, workloads


def _points(rng, n):
    X = np.zeros((n, 14))
    X[:, 0] = rng.uniform(1.2, 2.0, n); X[:, 1] = rng.uniform(5, 40, n); X[:, 2:4] = rng.normal(0, 3, (n, 2))
    X[:, 4] = rng.normal(-3, 1, n); X[:, 5:7] = rng.normal(0, .5, (n, 2))
    q = np.concatenate([np.ones((n, 1)), rng.normal(0, 0.2, (n, 3))], 1)
    X[:, 7:11] = q / np.linalg.norm(q, axis=1, keepdims=True)
    X[:, 11:14] = rng.normal(0, 0.2, (n, 3))
    U = np.stack([rng.uniform(1.5, 5, n), rng.uniform(-1, 1, n), rng.uniform(-1, 1, n)], 1)
    return X, U


@pytest.mark.parametrize("rk4", [False, True])
def test_step_and_linearize_match_oracle(env, rk4):
    from oracle import mpc6 as o6
    gp, mpc, dynamics, _ = env
    dyn = dynamics.Rocket6DoFDynamics(dynamics.Rocket6DoFConfig(use_rk4=rk4))
    p = o6.Rocket6DoFParams()
    X, U = _points(np.random.default_rng(3), 64)
    kw = dict(substeps=4 if rk4 else 1)
    Xn = dyn.step_batch(X, U, 0.1)
    ref = np.stack([o6.step6(X[i], U[i], 0.1, p, **kw) for i in range(64)])
    np.testing.assert_allclose(Xn, ref, rtol=1e-14, atol=1e-15)           # same operation order, no FMA contraction
    assert np.max(np.abs(np.linalg.norm(Xn[:, 7:11], axis=1) - 1.0)) < 1e-15
    AB = dyn.linearize_batch(X, U, 0.1).cpu().numpy()
    for i in range(0, 64, 7):
        A, B = o6.linearize_fd6(X[i], U[i], 0.1, p, **kw)
        # forward differences with eps = 1e-6 amplify the last-bit differences of the step by 1e6
        np.testing.assert_allclose(AB[i][:, :14], A, rtol=0, atol=5e-9)
        np.testing.assert_allclose(AB[i][:, 14:], B, rtol=0, atol=5e-9)
    A1, B1 = dyn.linearize(X[0], U[0], 0.1)
    assert A1.shape == (14, 14) and B1.shape == (14, 3) and np.array_equal(A1, AB[0][:, :14])


def _structured_gp(env, n=400, M=40):
    """StructuredRocketGP on the synthetic 6-DoF residual set + the per-output oracle fits on the SAME inducing points
    (read back after the reference-style re-draw)."""
    from oracle import gp as ogp
    gp, mpc, dynamics, workloads = env
    X, U, Dv, Dw = workloads.residual_dataset_6dof(n, seed=1)
    sg = gp.StructuredRocketGP(gp.StructuredGPConfig(n_inducing=M, use_sparse=True, noise_variance=1e-3))
    sg.add_data(X, U, Dv, Dw)
    np.random.seed(7)   # the reference draws the inducing points from the global numpy RNG (sparse_gp.py:122-148)
    sg.fit()
    Zv = ogp.features_batch(ogp.features_6dof_trans, X, U)
    Zw = ogp.features_batch(ogp.features_6dof_rot, X, U)
    fits_v = [ogp.sparse_fit(0, Zv, Dv[:, i], np.asarray(sg.gp_v.gps[0]._Z), np.ones(13), 1.0, 1e-3) for i in range(3)]
    fits_w = [ogp.sparse_fit(0, Zw, Dw[:, i], np.asarray(sg.gp_omega.gps[0]._Z), np.ones(12), 1.0, 1e-3) for i in range(3)]

    def predict(x, u):
        zv = ogp.features_6dof_trans(x, u)[None]; zw = ogp.features_6dof_rot(x, u)[None]
        dv = np.array([ogp.sparse_predict(f, zv)[0][0] for f in fits_v])
        dw = np.array([ogp.sparse_predict(f, zw)[0][0] for f in fits_w])
        return dv, dw

    def predict_full(x, u):
        zv = ogp.features_6dof_trans(x, u)[None]; zw = ogp.features_6dof_rot(x, u)[None]
        pv = [ogp.sparse_predict(f, zv) for f in fits_v]
        pw = [ogp.sparse_predict(f, zw) for f in fits_w]
        return (np.array([r[0][0] for r in pv]), np.array([r[0][0] for r in pw]),
                np.array([r[1][0] for r in pv]), np.array([r[1][0] for r in pw]))
    predict.full = predict_full
    return sg, predict


def test_rollout_with_structured_sparse_gp(env):
    """gp_mpc.py:139-175,277-281 on 14 states: the fused rollout (GP means evaluated in the kernel from the inducing
    points) against the sequential CPU loop; sparse posterior means agree to 1e-8 relative (K_uu + 1e-6 I, cond ~ 1e6)."""
    from oracle import mpc6 as o6
    gp, mpc, dynamics, workloads = env
    sg, predict = _structured_gp(env)
    dyn = dynamics.Rocket6DoFDynamics()
    ctl = mpc.GPMPC6DoF(dyn, sg, mpc.GPMPC6DoFConfig(N=12))
    ctl.setup()
    X0, _ = workloads.sample_initial_conditions_6dof(0, 5)
    rng = np.random.default_rng(0)
    U = np.zeros((5, 12, 3)); U[:, :, 0] = X0[:, :1] + 0.1 * rng.normal(size=(5, 12)); U[:, :, 1:] = 0.05 * rng.normal(size=(5, 12, 2))
    st = ctl._state_for(5, False)
    st.U.copy_(gp.engine.to_dev(U))
    ctl._rollout(st, gp.engine.to_dev(X0), st.U, st.X, st.dmean)
    Xg, Dg = st.X.cpu().numpy(), st.dmean.cpu().numpy()
    p = o6.Rocket6DoFParams()
    for b in range(5):
        Xo, Do = o6.rollout_gp(X0[b], U[b], 0.1, p, predict)
        scale = np.maximum(np.abs(Do), 1e-3)
        assert np.max(np.abs(Dg[b] - Do) / scale) < 1e-7
        np.testing.assert_allclose(Xg[b], Xo, rtol=1e-9, atol=1e-10)
    # nominal rollout (no GP): bit-level agreement with the plain step
    ctl0 = mpc.GPMPC6DoF(dyn, None, mpc.GPMPC6DoFConfig(N=12)); ctl0.setup()
    s0 = ctl0._state_for(5, False); s0.U.copy_(gp.engine.to_dev(U))
    ctl0._rollout(s0, gp.engine.to_dev(X0), s0.U, s0.X, None)
    Xn, _ = o6.rollout_gp(X0[0], U[0], 0.1, p, None)
    np.testing.assert_allclose(s0.X[0].cpu().numpy(), Xn, rtol=1e-14, atol=1e-15)


def test_assembly_and_solve_match_oracle(env):
    """One SQP iteration of the 6-DoF GP-MPC at the BASELINE horizon N = 30: QP data against oracle/mpc6.py (row order and
    signs of osqp_rti.py:260-372, affine GP term of gp_mpc.py:309-317,402-411), then the batched ADMM against the dense
    OSQP restatement on the same data: same status / iteration count, z within 1e-6, u0 within eps_abs + eps_rel |u0|."""
    from oracle import mpc as ompc
    from oracle import mpc6 as o6
    gp, mpc, dynamics, workloads = env
    sg, predict = _structured_gp(env)
    dyn = dynamics.Rocket6DoFDynamics()
    N, B = 30, 3
    ctl = mpc.GPMPC6DoF(dyn, sg, mpc.GPMPC6DoFConfig(N=N, max_sqp_iter=1, qp_max_iter=4000))
    X0, XT = workloads.sample_initial_conditions_6dof(0, B)
    sol = ctl.solve_batch(X0, XT)
    st = ctl._batch
    assert sol.X_opt.shape == (B, N + 1, 14) and sol.U_opt.shape == (B, N, 3) and sol.u0.shape == (B, 3)
    p = o6.Rocket6DoFParams()
    AB, beq, q = st.AB.cpu().numpy(), st.beq.cpu().numpy(), st.q.cpu().numpy()
    for b in range(B):
        U0 = np.zeros((N, 3)); U0[:, 0] = X0[b, 0]
        Xo, Do = o6.rollout_gp(X0[b], U0, 0.1, p, predict)
        gp_aff = np.zeros((N, 14)); gp_aff[:, 4:7] = Do[:, :3] * 0.1; gp_aff[:, 11:14] = Do[:, 3:] * 0.1
        ABo, rhs = o6.stage_blocks(N, Xo, U0, 0.1, p, gp_aff)
        np.testing.assert_allclose(AB[b], ABo, rtol=0, atol=5e-8)   # FD noise: ulp(|x| ~ 40) / eps = 7e-9 per bit
        np.testing.assert_allclose(beq[b][:14], X0[b], rtol=0, atol=0)
        np.testing.assert_allclose(beq[b][14:].reshape(N, 14), rhs, rtol=0, atol=2e-6)  # |x| ~ 40 times the FD noise
        Pd, qo = o6.build_cost(N, np.tile(XT[b], (N + 1, 1)))
        np.testing.assert_allclose(q[b], qo, rtol=1e-15, atol=0)
        np.testing.assert_allclose(ctl._qp.Pdiag.cpu().numpy(), Pd, rtol=0, atol=0)
        # solver parity on the device-assembled data
        A, l, u = o6.dense_constraints(N, AB[b], beq[b][:14], beq[b][14:].reshape(N, 14))
        np.testing.assert_allclose(ctl._qp.lb.cpu().numpy(), np.maximum(l[14 * (N + 1):], -1e30))
        s = ompc.OSQPSettings(max_iter=4000)
        orc = ompc.OSQPOracle(Pd, q[b], A, l, u, s)
        orc.warm_start(x=np.concatenate([np.concatenate([Xo[k], U0[k]]) for k in range(N)] + [Xo[N]]))
        r = orc.solve()
        assert st.ws.status[b].item() == r.status == 1, (b, st.ws.status[b].item(), r.status)
        assert st.ws.iters[b].item() == r.iters, (b, st.ws.iters[b].item(), r.iters)
        z = st.ws.x[b].cpu().numpy()
        assert np.max(np.abs(z - r.x)) < 1e-6 * max(1.0, np.max(np.abs(r.x)))
        u0 = r.x[14:17]
        assert np.max(np.abs(sol.u0[b] - u0)) < 1e-4 + 1e-4 * np.max(np.abs(u0))
        assert np.all(sol.U_opt[b] >= o6.U_MIN - 1e-2) and np.all(sol.U_opt[b] <= o6.U_MAX + 1e-2)  # ADMM primal tolerance
    # a second call starts from the stored warm start (gp_mpc.py:355-357) and needs no more ADMM iterations than the first
    it1 = st.qp_iters.cpu().numpy().copy()
    sol2 = ctl.solve_batch(sol.X_opt[:, 1], XT)
    assert np.all(sol2.qp_status == 1) and sol2.qp_iterations.sum() <= it1.sum()
    # single-scenario API: same numbers as row 0 of a fresh batch
    ctl1 = mpc.GPMPC6DoF(dyn, sg, mpc.GPMPC6DoFConfig(N=N, max_sqp_iter=1, qp_max_iter=4000))
    one = ctl1.solve(X0[0], XT[0])
    np.testing.assert_allclose(one.u0, sol.u0[0], rtol=1e-9, atol=1e-12)


def test_linear_uncertainty_propagation_14_state(env):
    """Step 2 of GPMPC.solve on the 14-state model (gp_mpc.py:283-289 -> uncertainty_prop.py:117-177): covariance chain
    along the GP-mean rollout, with the structured sparse GP's variances as process noise.  The chain kernel alone is
    checked first on random stage matrices (1e-13: same products, summation order of a plain triple loop)."""
    from oracle import mpc6 as o6
    gp, mpc, dynamics, workloads = env
    E = gp.engine
    torch = E.torch_()
    rng = np.random.default_rng(11)
    B, N, nx = 5, 9, 14
    A = rng.normal(size=(B, N, nx, 17)); A[..., :nx] = np.eye(nx) + 0.1 * A[..., :nx]
    qd = rng.uniform(0, 1e-3, size=(B, N, nx))
    covs = torch.zeros((B, N + 1, nx, nx), dtype=torch.float64, device="cuda")
    act = torch.ones((B,), dtype=torch.int32, device="cuda"); act[3] = 0
    Ad, qdd = E.to_dev(A), E.to_dev(qd)   # named: the buffers must outlive the asynchronous launch
    E.call("gpmpc_propagate_cov", nx, B, N, E.ptr(Ad), 17, nx * 17, E.ptr(qdd), 1e-6, E.ptr(covs), E.ptr(act), E.stream())
    covs = covs.cpu().numpy()
    for b in range(B):
        S = np.eye(nx) * 1e-6
        for k in range(N):
            if b != 3:
                np.testing.assert_allclose(covs[b, k], S, rtol=1e-13, atol=1e-13 * np.max(np.abs(S)))
            S = A[b, k, :, :nx] @ S @ A[b, k, :, :nx].T + np.diag(qd[b, k])
    assert np.all(covs[3] == 0.0)   # inactive scenario untouched
    # through the controller
    sg, predict = _structured_gp(env)
    dyn = dynamics.Rocket6DoFDynamics()
    N, B = 12, 3
    ctl = mpc.GPMPC6DoF(dyn, sg, mpc.GPMPC6DoFConfig(N=N, max_sqp_iter=1))
    X0, XT = workloads.sample_initial_conditions_6dof(0, B)
    ctl.solve_batch(X0, XT)
    cg = ctl.last_covariances()
    assert cg.shape == (B, N + 1, 14, 14)
    p = o6.Rocket6DoFParams()
    for b in range(B):
        U0 = np.zeros((N, 3)); U0[:, 0] = X0[b, 0]
        _, co = o6.propagate_linear6(X0[b], U0, 0.1, p, predict.full)
        # FD Jacobian noise (5e-8 absolute on entries of A) and the sparse-GP variance parity (1e-7) enter linearly
        assert np.max(np.abs(cg[b] - co)) < 1e-6 * np.max(np.abs(co)), np.max(np.abs(cg[b] - co)) / np.max(np.abs(co))
        # positive semi-definite: the quaternion normalisation inside the step makes A_k singular along q
        ev = np.linalg.eigvalsh(cg[b, -1])
        assert ev[0] > -1e-12 * ev[-1]
    off = mpc.GPMPC6DoF(dyn, sg, mpc.GPMPC6DoFConfig(N=N, max_sqp_iter=1, use_gp_uncertainty=False))
    off.solve_batch(X0, XT)
    assert off.last_covariances() is None

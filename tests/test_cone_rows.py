"""Convex thrust / glide-slope rows of the ADMM QP (second half; reference sub-problem constraints
gp_mpc.py:417-429 with the ConstraintParams defaults of constraints.py:35-50).

CPU: the polyhedral sets are INNER approximations of the reference's constraints, and the stage embedding (7 auxiliary
"states" per stage, include/gpmpc_b200.h GpmpcCone3) solves the same problem as the rows written directly.
GPU: the augmentation kernels against the oracle arrays, the controller with `convex_constraints=True` against
`GPMPCOracle(cone=...)` (same QP status / ADMM iteration counts, solution within the OSQP tolerance), and the constraint
satisfaction of every solved scenario of a batch."""
import numpy as np
import pytest

from oracle import mpc as ompc


def test_polyhedral_sets_are_inner_approximations():
    cone = ompc.ConeRows()
    rng = np.random.default_rng(0)
    u = rng.uniform([-1, -6, -6], [6, 6, 6], size=(400000, 3))
    b = cone.facet_b
    ok = (u[:, 0] >= cone.T_min) & (u[:, 0] + b * np.abs(u[:, 1]) <= cone.T_max) & (u[:, 0] + b * np.abs(u[:, 2]) <= cone.T_max)
    nrm = np.linalg.norm(u[ok], axis=1)
    assert ok.sum() > 1000 and nrm.max() <= cone.T_max and nrm.min() >= cone.T_min      # gp_mpc.py:419-421
    # the corner of the base square touches the ball: the approximation is not needlessly small
    w = (cone.T_max - cone.T_min) / b
    assert np.hypot(cone.T_min, np.sqrt(2) * w) == pytest.approx(cone.T_max, rel=1e-12)
    tan_g = 2 * cone.half_tan
    for e in ((1.0, 0.0), (0.3, -2.0), (-1.0, 1e-12), (0.0, 0.0)):
        n = ompc.glide_normals(*e)
        pts = rng.uniform(-40, 40, size=(200000, 2)); alt = rng.uniform(-5, 40, size=200000)
        inside = np.all(pts @ n.T - cone.half_tan * alt[:, None] <= 0, axis=1)
        assert inside.sum() > 1000
        assert np.all(alt[inside] ** 2 * tan_g**2 >= (pts[inside] ** 2).sum(1) - 1e-9) and alt[inside].min() >= 0   # :425-427
        rho = np.hypot(*e)
        if rho > 1e-9:   # the vertex in the nominal radial direction lies ON the cone: full radius available there
            v = np.array(e) / rho * (10 * tan_g)
            assert np.max(n @ v - cone.half_tan * 10) == pytest.approx(0.0, abs=1e-12)


def _problem(seed, N=12):
    rng = np.random.default_rng(seed)
    p = ompc.Rocket3DoFParams()
    x0 = np.array([2.0, 30 + rng.normal(), 3 * rng.normal(), 3 * rng.normal(), -3 + rng.normal(), 0.5 * rng.normal(),
                   0.5 * rng.normal()])
    U = np.tile([x0[0] * p.g0, 0, 0], (N, 1)) + 0.1 * rng.normal(size=(N, 3))
    X = np.zeros((N + 1, 7)); X[0] = x0
    for k in range(N):
        X[k + 1] = ompc.step_3dof(X[k], U[k], 0.1, p)
    xt = x0.copy(); xt[4:] = 0; xt[1] = x0[1] - 2
    Pd, q = ompc.build_cost(N, np.tile(xt, (N + 1, 1)))
    A, l, u = ompc.build_constraints(N, X, U, x0, 0.1, p, "analytic", None, mode="gpmpc")
    return N, X, U, Pd, q, A, l, u


@pytest.mark.parametrize("seed", [0, 1, 2])
def test_stage_embedding_solves_the_directly_constrained_qp(seed):
    cone = ompc.ConeRows(T_max=2.6)        # tight enough to be active around the hover thrust of 2
    N, X, U, Pd, q, A, l, u = _problem(seed)
    S = ompc.OSQPSettings(max_iter=40000, eps_abs=1e-9, eps_rel=1e-9)
    # (i) rows written directly on z = [x_0,u_0,...,x_N]
    C, hi = ompc.cone_direct_rows(N, X, cone)
    n, _ = ompc.qp_dims(N)
    lo_u = l.copy()
    for k in range(N):
        lo_u[7 * (N + 1) + k * 10 + 7] = max(lo_u[7 * (N + 1) + k * 10 + 7], cone.T_min)
    r1 = ompc.OSQPOracle(Pd, q, np.vstack([A, C]), np.r_[lo_u, np.full(len(hi), -np.inf)], np.r_[u, hi], S).solve()
    # (ii) the augmented stage form the device solves
    AB = np.stack([A[7 + k * 7:14 + k * 7, k * 10:(k + 1) * 10] for k in range(N)])
    ABc, beqc, qc, zc = ompc.augment_cone_qp(N, AB, l[:7 * (N + 1)], q, X, U, cone)
    Pdc, lbc, ubc = ompc.cone_shared_data(N, cone)
    Ac, lc, uc = ompc.dense_stage_constraints(ompc.NXC, 3, N, ABc, beqc, lbc, ubc)
    r2 = ompc.OSQPOracle(Pdc, qc, Ac, lc, uc, S).solve()
    assert r1.status == r2.status == ompc.SOLVED
    z2 = np.concatenate([np.r_[r2.x[k * 17:k * 17 + 7], r2.x[k * 17 + 14:k * 17 + 17]] for k in range(N)] + [r2.x[N * 17:N * 17 + 7]])
    np.testing.assert_allclose(z2, r1.x, rtol=0, atol=2e-6)
    # the auxiliary variables ARE the row values, and the rows bind somewhere (otherwise the test proves nothing)
    xi = np.stack([r2.x[k * 17 + 7:k * 17 + 14] for k in range(1, N + 1)])
    np.testing.assert_allclose(xi.ravel(), C @ r1.x, rtol=0, atol=5e-6)
    assert np.sum(np.abs(C @ r1.x - hi) < 1e-5) >= 1
    assert np.isinf(hi[4:7]).all() and np.isfinite(hi[11:14]).all()      # stage 1 carries no glide rows
    # the warm start the device writes is consistent: xi of the nominal trajectory
    assert np.allclose(zc[17 + 7:17 + 14], ompc.cone_values(cone, X[1], U[0]))


@pytest.mark.gpu
def test_augmentation_kernels_match_oracle():
    import torch
    from gp_mpc_rocket_landing_b200 import _lib
    from gp_mpc_rocket_landing_b200.gp import engine as ge
    cone = ompc.ConeRows()
    B, N = 5, 20
    probs = [_problem(s, N) for s in range(B)]
    f64 = dict(dtype=torch.float64, device="cuda")
    AB = np.stack([np.stack([pr[5][7 + k * 7:14 + k * 7, k * 10:(k + 1) * 10] for k in range(N)]) for pr in probs])
    beq = np.stack([pr[6][:7 * (N + 1)] for pr in probs]); q = np.stack([pr[4] for pr in probs])
    X = np.stack([pr[1] for pr in probs]); U = np.stack([pr[2] for pr in probs])
    X[3, :, 2:4] = 0.0          # a scenario on the axis: the default orientation
    n_c, _ = ompc.cone_dims(N)
    ABc = torch.full((B, N, 14, 17), np.nan, **f64); beqc = torch.full((B, 14 * (N + 1)), np.nan, **f64)
    qc = torch.full((B, n_c), np.nan, **f64); zc = torch.full((B, n_c), np.nan, **f64)
    c = _lib.Cone3(1, cone.T_max, cone.facet_b, cone.half_tan)
    act = torch.ones(B, dtype=torch.int32, device="cuda"); act[4] = 0
    d = [ge.to_dev(a) for a in (AB, beq, q, X, U)]      # keep the device inputs alive across the call
    ge.call("gpmpc_qp_augment_cone_3dof", _lib.C.byref(c), B, N, *[ge.ptr(t) for t in d], ge.ptr(ABc), ge.ptr(beqc),
            ge.ptr(qc), ge.ptr(zc), ge.ptr(act), ge.stream())
    for b in range(4):
        want = ompc.augment_cone_qp(N, AB[b], beq[b], q[b], X[b], U[b], cone)
        for got, w in zip((ABc[b], beqc[b], qc[b], zc[b]), want):
            np.testing.assert_allclose(got.cpu().numpy(), w, rtol=0, atol=1e-14)
    assert torch.isnan(ABc[4]).all() and torch.isnan(zc[4]).all()        # inactive scenario untouched
    z = torch.zeros((B, 10 * N + 7), **f64)
    ge.call("gpmpc_qp_extract_cone_3dof", B, N, ge.ptr(zc), ge.ptr(z), ge.ptr(act), ge.stream())
    for b in range(4):
        np.testing.assert_array_equal(z[b].cpu().numpy(), ompc.traj_to_vec(X[b], U[b]))


@pytest.mark.gpu
def test_controller_with_convex_rows_matches_oracle_and_satisfies_the_cones():
    import torch
    from gp_mpc_rocket_landing_b200 import workloads
    from gp_mpc_rocket_landing_b200.dynamics import Rocket3DoFDynamics
    from gp_mpc_rocket_landing_b200.gp import Simple3DoFGP
    from gp_mpc_rocket_landing_b200.mpc import ConstraintParams, GPMPC, GPMPCConfig
    from oracle import gp as ogp
    Xd, Ud, Dd = workloads.residual_dataset(300, seed=0)
    gp = Simple3DoFGP(noise_variance=1e-2, use_sparse=False); gp.add_data(Xd, Ud, Dd); gp.fit()
    o = ogp.Simple3DoFGPOracle(noise_variance=1e-2); o.add_data(Xd, Ud, Dd); o.fit()
    cp = ConstraintParams(T_max=3.0)       # active from the first step (braking wants more than 3)
    cone = ompc.ConeRows(T_min=cp.T_min, T_max=cp.T_max, gamma_gs_deg=cp.gamma_gs)
    B = 64
    X0, _ = workloads.sample_initial_conditions(0, B)
    X0[1, 2:4] = [0.52 * X0[1, 1], 0.0]    # close to the edge of the glide-slope cone (tan 30 deg = 0.577): facets bind
    X0[2, 2:4] = [0.56 * X0[2, 1], 0.0]    # too close: no admissible thrust keeps the descent inside -> primal infeasible
    feas = np.ones(B, dtype=bool); feas[2] = False
    ctl = GPMPC(Rocket3DoFDynamics(), gp, GPMPCConfig(N=20, dt=0.1, max_sqp_iter=2, qp_max_iter=4000, convex_constraints=True),
                constraint_params=cp)
    ctl.setup()
    idx = [0, 1, 2, 17, 63]
    ors = {b: ompc.GPMPCOracle(o.predict, N=20, dt=0.1, settings=ompc.OSQPSettings(max_iter=4000), max_sqp_iter=2, cone=cone)
           for b in idx}
    x = torch.as_tensor(X0, device="cuda").clone()
    p = ompc.Rocket3DoFParams()
    for step in range(2):
        sol = ctl.solve_batch(x, per_step_target=True)
        # prologue, rollout, variance, cov | assemble, augment, ADMM, extract, update | re-propagation (3) | GP means + 5
        assert ctl.last_kernel_launches == 4 + 5 + 3 + 6
        Uo = sol.U_opt.cpu().numpy(); Xo = sol.X_opt.cpu().numpy(); qs = sol.qp_status.cpu().numpy()
        qi = sol.qp_iterations.cpu().numpy(); xh = x.cpu().numpy()
        assert (qs[feas] == 1).all()
        if step == 0:
            assert qs[2] == -3 and np.isinf(sol.cost.cpu().numpy()[2])     # OSQP_PRIMAL_INFEASIBLE -> nominal kept, cost inf
        # every scenario: thrust inside the ball and above T_min, positions of stages >= 2 inside the glide-slope cone
        # (to the OSQP primal tolerance eps_abs + eps_rel max(|Ax|, |z|) with |z| up to the 500 m altitude bound)
        tol = 1e-4 + 1e-4 * 500.0
        nrm = np.linalg.norm(Uo[feas], axis=2)
        assert nrm.max() <= cp.T_max + tol and Uo[feas][:, :, 0].min() >= cp.T_min - tol
        lat = np.hypot(Xo[feas][:, 2:, 2], Xo[feas][:, 2:, 3]); alt = Xo[feas][:, 2:, 1]
        assert np.all(lat <= alt * np.tan(cp.gamma_gs_rad) + tol)
        assert nrm.max() > cp.T_max - 0.05          # and the thrust facets do bind
        for b in idx:
            xt = xh[b].copy(); xt[4:7] = 0; xt[1] = max(0.5, xh[b][1] - 2.0)
            r = ors[b].solve(xh[b], xt)
            if b == 2 and step == 0:
                assert r.qp_status == [-3] and int(qi[b]) == sum(r.qp_iters), (qi[b], r.qp_iters)
                continue
            if b == 2:
                continue
            assert r.qp_status == [1, 1] and int(qi[b]) == sum(r.qp_iters), (step, b, qi[b], r.qp_iters)
            np.testing.assert_allclose(Uo[b], r.U_opt, rtol=0, atol=5e-5)
            np.testing.assert_allclose(Xo[b], r.X_opt, rtol=0, atol=5e-5)
        xn = np.stack([ompc.step_3dof(xh[b], Uo[b, 0], 0.1, p) for b in range(B)])
        x = torch.as_tensor(xn, device="cuda")

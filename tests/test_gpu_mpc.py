"""GPU parity: MPC-side kernels through the C-ABI against the oracle (oracle/mpc.py) and the golden vectors the
real reference assembler / propagator produced (tests/golden/qp_assembly.npz, propagation.npz)."""
import numpy as np
import pytest

from conftest import load_golden, relerr

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def E():
    from gp_mpc_rocket_landing_b200.gp import engine as ge
    from gp_mpc_rocket_landing_b200.mpc import engine as me
    ge.torch_()
    return ge, me


def _dyn(me, dt=0.1):
    from oracle import mpc as ompc
    p = ompc.Rocket3DoFParams()
    return me.make_dyn(p.alpha, p.g_vec, dt), p


def _assemble(ge, me, N, X_lin, U_lin, x_init, x_ref, lin_mode=0, rhs_mode=0, c_gp=None):
    import torch
    dyn, _ = _dyn(me)
    qp = me.QPData.reference(N)
    B = X_lin.shape[0]
    n, m = me.qp_dims(N)
    AB = torch.zeros((B, N, 7, 10), dtype=torch.float64, device="cuda")
    beq = torch.zeros((B, 7 * (N + 1)), dtype=torch.float64, device="cuda")
    q = torch.zeros((B, n), dtype=torch.float64, device="cuda")
    me.qp_assemble(dyn, qp, ge.to_dev(X_lin), ge.to_dev(U_lin), ge.to_dev(x_init), ge.to_dev(x_ref), AB, beq, q,
                   c_gp=None if c_gp is None else ge.to_dev(c_gp), lin_mode=lin_mode, rhs_mode=rhs_mode)
    return qp, AB, beq, q


def test_qp_assembly_matches_reference(E):
    ge, me = E
    g = load_golden("qp_assembly.npz")
    x0, xt = g["x0"], g["xt"]
    for tag, N, lin in (("fast20", 20, 0), ("fast15", 15, 0), ("fd20", 20, 1)):
        X_lin = np.array([(1 - k / N) * x0 + (k / N) * xt for k in range(N + 1)])
        U_lin = np.zeros((N, 3)); U_lin[:, 0] = x0[0] * 1.0
        x_ref = np.tile(xt, (N + 1, 1))
        cases = [(X_lin, U_lin, x0, "A", "l", "u"), (g[f"{tag}_Xl"], g[f"{tag}_Ul"], g[f"{tag}_xc"], "A2", "l2", "u2")]
        for Xl, Ul, xc, kA, kl, ku in cases:
            qp, AB, beq, q = _assemble(ge, me, N, Xl[None], Ul[None], xc[None], x_ref, lin_mode=lin)
            A, l, u = me.dense_constraints(N, AB[0].cpu().numpy(), beq[0].cpu().numpy(), qp.lb.cpu().numpy(), qp.ub.cpu().numpy())
            # matrix entries: same scalar formulas, same |.|>1e-10 filter -> identical pattern, values to 1 ulp
            assert np.array_equal(A != 0, g[f"{tag}_{kA}"] != 0)
            np.testing.assert_allclose(A, g[f"{tag}_{kA}"], rtol=3e-16, atol=0)
            lr = np.maximum(g[f"{tag}_{kl}"], -1e30); ur = np.minimum(g[f"{tag}_{ku}"], 1e30)
            # right-hand sides c_k = f - A x - B u: 10-term sums in a different association than numpy's matmul
            np.testing.assert_allclose(l, lr, rtol=0, atol=2e-13)
            np.testing.assert_allclose(u, ur, rtol=0, atol=2e-13)
            np.testing.assert_array_equal(q[0].cpu().numpy(), g[f"{tag}_q"])
        np.testing.assert_array_equal(qp.Pdiag.cpu().numpy(), np.diag(g[f"{tag}_P"]))


def test_gpmpc_subproblem_assembly_matches_the_reference_arguments(E):
    """a18 on the device: `qp_assemble` in the GP-MPC deviation form (rhs_mode 1, gp_mpc.py:402-411) with the affine GP term,
    against the first sub-problem of tests/golden/gpmpc_skeleton.npz -- built in the fixture from the A_list / B_list / c_list
    that the reference's own `GPMPC.solve` hands to `_solve_qp`."""
    ge, me = E
    g = load_golden("gpmpc_skeleton.npz")
    N = 20
    c_gp = g["qp0_c"][:, 4:7] / 0.1            # the kernel takes the GP means d and forms c = [0,0,0,0,d dt] itself
    qp, AB, beq, q = _assemble(ge, me, N, g["qp0_X"][None], g["qp0_U"][None], g["qp0_x0"][None], g["qp0_xref"], lin_mode=0,
                               rhs_mode=1, c_gp=c_gp[None])
    A, l, u = me.dense_constraints(N, AB[0].cpu().numpy(), beq[0].cpu().numpy(), qp.lb.cpu().numpy(), qp.ub.cpu().numpy())
    assert np.array_equal(A != 0, g["qp0_A"] != 0)
    np.testing.assert_allclose(A, g["qp0_A"], rtol=3e-16, atol=0)
    np.testing.assert_allclose(l, np.maximum(g["qp0_l"], -1e30), rtol=0, atol=2e-13)
    np.testing.assert_allclose(u, np.minimum(g["qp0_u"], 1e30), rtol=0, atol=2e-13)
    np.testing.assert_array_equal(q[0].cpu().numpy(), g["qp0_q"])
    np.testing.assert_array_equal(qp.Pdiag.cpu().numpy(), g["qp0_Pd"])


def _oracle_qp(N, x0, xt, seed=0, perturb=0.0):
    from oracle import mpc as ompc
    rng = np.random.default_rng(seed)
    p = ompc.Rocket3DoFParams()
    X_lin = np.array([(1 - k / N) * x0 + (k / N) * xt for k in range(N + 1)]) + perturb * rng.normal(size=(N + 1, 7))
    U_lin = np.zeros((N, 3)); U_lin[:, 0] = x0[0]
    U_lin += perturb * rng.normal(size=(N, 3))
    Pd, q = ompc.build_cost(N, np.tile(xt, (N + 1, 1)))
    A, l, u = ompc.build_constraints(N, X_lin, U_lin, x0, 0.1, p, "analytic")
    return Pd, q, A, l, u, X_lin, U_lin


@pytest.mark.parametrize("variant", [0, 1, 3])  # 0: stage-parallel kernel (admm_stage.cu), 1: first generation, 3: admm3.cu
@pytest.mark.parametrize("max_iter", [50, 4000])
def test_admm_matches_osqp_restatement(E, max_iter, variant):
    """Same algorithm, different linear algebra (block-tridiagonal reduced KKT vs dense LU): statuses, iteration
    counts and rho histories must agree, iterates to ~1e-7 relative."""
    import torch
    from oracle import mpc as ompc
    ge, me = E
    N = 20
    cfg = ompc.SimConfig()
    x0s = np.stack([ompc.sample_initial_condition(42 + i, cfg) for i in range(6)])
    x0s = np.vstack([x0s, [[2.0, 500, 50, 0, -75, 10, 0]]])  # README x0: violates the velocity box -> infeasible
    B = x0s.shape[0]
    xts = np.zeros_like(x0s); xts[:, 0] = x0s[:, 0]
    Xl = np.zeros((B, N + 1, 7)); Ul = np.zeros((B, N, 3)); ref = []
    for b in range(B):
        Pd, q, A, l, u, Xl[b], Ul[b] = _oracle_qp(N, x0s[b], xts[b], seed=b, perturb=0.01)
        s = ompc.OSQPSettings(max_iter=max_iter)
        ref.append((Pd, q, A, l, u, ompc.OSQPOracle(Pd, q, A, l, u, s).solve()))
    x_ref = np.stack([np.tile(xts[b], (N + 1, 1)) for b in range(B)])
    qp, AB, beq, q = _assemble(ge, me, N, Xl, Ul, x0s, x_ref)
    ws = me.AdmmWorkspace(B, N)
    S = me.AdmmSettings.reference_defaults(max_iter=max_iter, kernel_variant=variant)
    me.admm_solve(qp, AB, beq, q, S, ws)
    torch.cuda.synchronize()
    st, it = ws.status.cpu().numpy(), ws.iters.cpu().numpy()
    x, y, rho = ws.x.cpu().numpy(), ws.y.cpu().numpy(), ws.rho.cpu().numpy()
    for b in range(B):
        r = ref[b][5]
        assert st[b] == r.status, (b, st[b], r.status)
        assert it[b] == r.iters, (b, it[b], r.iters)
        assert rho[b] == pytest.approx(r.rho, rel=1e-6)
        scale = max(1.0, np.max(np.abs(r.x)))
        assert np.max(np.abs(x[b] - r.x)) < 1e-6 * scale
        assert np.max(np.abs(y[b] - r.y)) < 1e-6 * max(1.0, np.max(np.abs(r.y)))
        assert ws.obj[b].item() == pytest.approx(r.obj_val, rel=1e-6, abs=1e-6)
    if max_iter == 4000:
        assert (st[:6] == 1).all() and st[6] in (-3, 3)
        # Solver-independent check: re-solve on the GPU at eps = 1e-9 and verify the KKT conditions of the inequality
        # QP in plain numpy (sufficient for optimality: the QP is convex).  Then u0 at the reference tolerance
        # (eps_abs = eps_rel = 1e-4) must be within that tolerance of the optimum, with the same (strongly) active set.
        ws2 = me.AdmmWorkspace(B, N)
        S2 = me.AdmmSettings.reference_defaults(max_iter=20000, eps_abs=1e-9, eps_rel=1e-9, kernel_variant=variant)
        me.admm_solve(qp, AB, beq, q, S2, ws2)
        torch.cuda.synchronize()
        x2, y2, st2 = ws2.x.cpu().numpy(), ws2.y.cpu().numpy(), ws2.status.cpu().numpy()
        for b in range(6):
            Pd, qv, A, l, u, r = ref[b]
            assert st2[b] == 1
            prim, stat, comp = ompc.kkt_residuals(Pd, qv, A, l, u, x2[b], y2[b])
            ys = max(1.0, np.max(np.abs(y2[b])))
            assert prim < 1e-6 and stat < 1e-9 * ys * 10 and comp < 1e-8 * ys * 10, (b, prim, stat, comp)
            tol = 1e-4 + 1e-4 * np.max(np.abs(x2[b]))   # OSQP: eps_abs + eps_rel * |.|_inf
            assert np.max(np.abs(x[b][7:10] - x2[b][7:10])) < tol
            # same active set: GPU vs CPU restatement at every multiplier above noise, and both vs the optimum on the
            # strongly active rows (OSQP convention: y < 0 lower bound active, y > 0 upper bound active)
            ytol = 1e-6 * max(1.0, np.max(np.abs(r.y)))
            assert np.array_equal(y[b] < -ytol, r.y < -ytol) and np.array_equal(y[b] > ytol, r.y > ytol)
            lo, up = ompc.active_set(y[b], l, u, 1e-3 * ys)
            lo2, up2 = ompc.active_set(y2[b], l, u, 1e-3 * ys)
            assert np.array_equal(lo, lo2) and np.array_equal(up, up2) and lo2.sum() + up2.sum() > 0


def _random_stage_qp(nx, nu, N, rng):
    """A feasible stage-structured QP in the reference's row order (osqp_rti.py:260-372) for any state dimension."""
    ns = nx + nu
    n, neq = N * ns + nx, nx * (N + 1)
    AB = np.zeros((N, nx, ns))
    for k in range(N):
        AB[k, :, :nx] = np.eye(nx) + 0.1 * rng.normal(size=(nx, nx)) * (rng.random((nx, nx)) < 0.3)
        AB[k, :, nx:] = 0.1 * rng.normal(size=(nx, nu))
    x0 = rng.normal(size=nx)
    beq = np.concatenate([x0, 0.01 * rng.normal(size=nx * N)])
    Pd = np.concatenate([np.tile(np.concatenate([rng.uniform(0.5, 10.0, nx), np.full(nu, 0.01)]), N),
                         rng.uniform(5.0, 100.0, nx)])
    q = rng.normal(size=n)
    lb, ub = np.full(n, -1e30), np.full(n, 1e30)
    for k in range(N):
        lb[k * ns + nx:(k + 1) * ns], ub[k * ns + nx:(k + 1) * ns] = -0.5, 0.5           # input box (active)
        lb[k * ns + 1:k * ns + 4], ub[k * ns + 1:k * ns + 4] = -3.0, 3.0                  # a few state bounds
    A = np.zeros((neq + n, n))
    A[:nx, :nx] = np.eye(nx)
    for k in range(N):
        r = nx + k * nx
        A[r:r + nx, k * ns:(k + 1) * ns] = AB[k]
        A[r:r + nx, (k + 1) * ns:(k + 1) * ns + nx] = -np.eye(nx)
    A[neq:, :] = np.eye(n)
    l = np.concatenate([beq, lb]); u = np.concatenate([beq, ub])
    return AB, beq, Pd, q, lb, ub, A, l, u


@pytest.mark.parametrize("nx,N,variant", [(14, 30, 0), (14, 30, 2), (14, 20, 0), (14, 20, 2), (14, 8, 0), (14, 8, 1), (7, 11, 0),
                                          (7, 20, 3)])
def test_admm_generic_stage_qp(E, nx, N, variant):
    """The solver on random stage-structured QPs of the 6-DoF size (nx = 14, N = 30: BASELINE configs[3]) and on a
    run-time horizon, against the dense CPU restatement: same status / iterations / rho, iterates to 1e-6."""
    import torch
    from oracle import mpc as ompc
    ge, me = E
    rng = np.random.default_rng(100 * nx + N)
    B = 3
    probs = [_random_stage_qp(nx, 3, N, rng) for _ in range(B)]
    for b in range(1, B):  # Pdiag and the box are shared by all scenarios in the C-ABI
        probs[b] = probs[b][:2] + probs[0][2:3] + probs[b][3:4] + probs[0][4:6] + probs[b][6:7] + (
            np.concatenate([probs[b][1], probs[0][4]]), np.concatenate([probs[b][1], probs[0][5]]))
    AB = ge.to_dev(np.stack([p[0] for p in probs])); beq = ge.to_dev(np.stack([p[1] for p in probs]))
    q = ge.to_dev(np.stack([p[3] for p in probs]))
    qp = me.QPData(N, ge.to_dev(probs[0][2]), ge.to_dev(probs[0][4]), ge.to_dev(probs[0][5]), None, None, None)
    ws = me.AdmmWorkspace(B, N, nx=nx, nu=3)
    S = me.AdmmSettings.reference_defaults(max_iter=2000, kernel_variant=variant)
    me.admm_solve(qp, AB, beq, q, S, ws)
    torch.cuda.synchronize()
    for b in range(B):
        _, _, Pd, qv, _, _, A, l, u = probs[b]
        r = ompc.OSQPOracle(Pd, qv, A, l, u, ompc.OSQPSettings(max_iter=2000)).solve()
        assert ws.status[b].item() == r.status == 1 and ws.iters[b].item() == r.iters, (b, ws.iters[b].item(), r.iters)
        assert ws.rho[b].item() == pytest.approx(r.rho, rel=1e-6)
        assert np.max(np.abs(ws.x[b].cpu().numpy() - r.x)) < 1e-6 * max(1.0, np.max(np.abs(r.x)))
        assert np.max(np.abs(ws.y[b].cpu().numpy() - r.y)) < 1e-6 * max(1.0, np.max(np.abs(r.y)))


@pytest.mark.parametrize("variant", [0, 1, 3])
def test_admm_warm_start_and_inactive(E, variant):
    """x/y/rho persist across solves like the OSQP object does (osqp_rti.py:522-527); inactive scenarios are
    untouched."""
    import torch
    from oracle import mpc as ompc
    ge, me = E
    N = 15
    x0 = np.array([2.0, 30.0, 1.0, -2.0, -3.0, 0.2, 0.1]); xt = np.array([2.0, 0, 0, 0, 0, 0, 0.0])
    Pd, q, A, l, u, Xl, Ul = _oracle_qp(N, x0, xt)
    s = ompc.OSQPSettings(max_iter=200)
    o = ompc.OSQPOracle(Pd, q, A, l, u, s)
    r1 = o.solve()
    o.warm_start(x=r1.x)
    r2 = o.solve()
    qp, AB, beq, qd = _assemble(ge, me, N, np.stack([Xl, Xl]), np.stack([Ul, Ul]), np.stack([x0, x0]), np.tile(xt, (N + 1, 1)))
    ws = me.AdmmWorkspace(2, N)
    S = me.AdmmSettings.reference_defaults(max_iter=200, kernel_variant=variant)
    me.admm_solve(qp, AB, beq, qd, S, ws)
    assert ws.iters.cpu().tolist() == [r1.iters, r1.iters]
    S2 = me.AdmmSettings.reference_defaults(max_iter=200, warm_x=1, warm_y=1, rho_from_input=1, kernel_variant=variant)
    active = torch.tensor([1, 0], dtype=torch.int32, device="cuda")
    x_before = ws.x.clone()
    me.admm_solve(qp, AB, beq, qd, S2, ws, active=active)
    assert ws.iters[0].item() == r2.iters and ws.status[0].item() == r2.status
    assert np.max(np.abs(ws.x[0].cpu().numpy() - r2.x)) < 1e-6 * max(1, np.max(np.abs(r2.x)))
    assert torch.equal(ws.x[1], x_before[1])


def test_rollout_and_propagation_match_reference(E):
    """gp_mpc.py:277-281 rollout + uncertainty_prop.py:117-177 against the golden output of the reference's own
    UncertaintyPropagator (14-state embedding) and the oracle."""
    import torch
    from oracle import gp as ogp, mpc as ompc
    ge, me = E
    g = load_golden("propagation.npz")
    dyn, p = _dyn(me)
    Z = ge.features(0, g["X"], g["U"])
    f = ge.ExactFactor(0, np.ones(11), 1.0, float(g["noise"])).fit(Z)
    model = ge.ExactModelDev(f).fit_targets(g["D"])
    N = g["Useq"].shape[0]
    rng = np.random.default_rng(0)
    B = 9
    x0 = np.tile(g["x0"], (B, 1)); U = np.tile(g["Useq"], (B, 1, 1))
    x0[1:] += 0.05 * rng.normal(size=(B - 1, 7)); U[1:] += 0.05 * rng.normal(size=(B - 1, N, 3))
    X = torch.zeros((B, N + 1, 7), dtype=torch.float64, device="cuda")
    dmean = torch.zeros((B, N, 3), dtype=torch.float64, device="cuda")
    q_pad = ge.pad_to(B * N, 128)
    Ks = torch.zeros((q_pad, f.n_pad), dtype=torch.float64, device="cuda")
    ms = model.struct()
    me.rollout(ms, dyn, ge.to_dev(x0), ge.to_dev(U), X, dmean, Ks)
    sumsq = f.sumsq_inplace(Ks)
    dvar = torch.zeros((q_pad, 3), dtype=torch.float64, device="cuda")
    ge.call("gpmpc_predict_var", ge.ptr(sumsq), None, B * N, 3, f.sv, 1e-10, ge.ptr(model.ystd), ge.ptr(dvar), 3, 0, ge.stream())
    covs = torch.zeros((B, N + 1, 7, 7), dtype=torch.float64, device="cuda")
    me.propagate_cov(dyn, X, ge.to_dev(U), dvar[: B * N].contiguous(), covs)
    Xh, covh = X.cpu().numpy(), covs.cpu().numpy()
    np.testing.assert_allclose(Xh[0], g["means"], rtol=1e-9, atol=1e-11)
    np.testing.assert_allclose(covh[0], g["covs"], rtol=1e-8, atol=1e-18)
    o = ogp.Simple3DoFGPOracle(noise_variance=float(g["noise"]))
    o.add_data(g["X"], g["U"], g["D"]); o.fit()
    for b in (1, 5, 8):
        means, cv, dm, dv = ompc.propagate_linear(x0[b], U[b], np.eye(7) * 1e-6, 0.1, p, o.predict)
        np.testing.assert_allclose(Xh[b], means, rtol=1e-9, atol=1e-11)
        np.testing.assert_allclose(covh[b], cv, rtol=1e-8, atol=1e-18)
        np.testing.assert_allclose(dmean[b].cpu().numpy(), dm, rtol=1e-8, atol=1e-11)
        np.testing.assert_allclose(dvar[b * N:(b + 1) * N].cpu().numpy(), dv, rtol=1e-8)
    # nominal rollout (no model) = plain Euler steps
    Xn = torch.zeros_like(X)
    me.rollout(None, dyn, ge.to_dev(x0), ge.to_dev(U), Xn)
    xs = x0[3].copy()
    for k in range(N):
        xs = ompc.step_3dof(xs, U[3, k], 0.1, p)
    np.testing.assert_allclose(Xn[3, N].cpu().numpy(), xs, rtol=1e-15, atol=1e-15)


def test_mpc_update_semantics(E):
    """mode 1: gp_mpc.py:336-345 (+ :478-482 failure -> nominal, which 'converges' with cost=inf);
    mode 0: osqp_rti.py:532-552 (accept + shift, or keep everything)."""
    import torch
    ge, me = E
    N = 5
    n, _ = me.qp_dims(N)
    qp = me.QPData.reference(N)
    rng = np.random.default_rng(1)
    B = 4
    X = rng.normal(size=(B, N + 1, 7)); U = rng.normal(size=(B, N, 3))
    z = np.zeros((B, n))
    for b in range(B):
        for k in range(N):
            z[b, k * 10:k * 10 + 7] = X[b, k]; z[b, k * 10 + 7:k * 10 + 10] = U[b, k]
        z[b, N * 10:] = X[b, N]
    z[1] += 1e-5; z[2] += 1e-3
    status = np.array([1, 1, 1, -2], dtype=np.int32)
    x_ref = np.zeros((N + 1, 7))
    t = lambda a, dt=None: ge.to_dev(a, dt)
    Xd, Ud = t(X), t(U)
    Xp, Up = torch.zeros_like(Xd), torch.zeros_like(Ud)
    cost = torch.zeros(B, dtype=torch.float64, device="cuda")
    conv = torch.zeros(B, dtype=torch.int32, device="cuda"); act = torch.ones(B, dtype=torch.int32, device="cuda")
    nit = torch.zeros(B, dtype=torch.int32, device="cuda")
    me.mpc_update(qp, 1, t(z), t(status, torch.int32), t(x_ref), Xd, Ud, Xp, Up, cost, conv, act, nit)
    assert conv.cpu().tolist() == [1, 1, 0, 1] and act.cpu().tolist() == [0, 0, 1, 0] and nit.cpu().tolist() == [1] * 4
    c = cost.cpu().numpy()
    assert np.isinf(c[3]) and np.isfinite(c[:3]).all()
    Xn, Un = z[0][: N * 10].reshape(N, 10)[:, :7], z[0][: N * 10].reshape(N, 10)[:, 7:]
    expect = np.sum(me.Q_DIAG * Xn**2) + np.sum(me.R_DIAG * Un**2) + np.sum(me.QF_DIAG * z[0][N * 10:] ** 2)
    assert c[0] == pytest.approx(expect, rel=1e-13)
    np.testing.assert_allclose(Xd[2].cpu().numpy(), X[2] + 1e-3, rtol=0, atol=1e-15)
    np.testing.assert_array_equal(Xd[3].cpu().numpy(), X[3])
    # shifted warm start of a solved scenario
    np.testing.assert_array_equal(Xp[2, :-1].cpu().numpy(), Xd[2, 1:].cpu().numpy())
    np.testing.assert_array_equal(Xp[2, -1].cpu().numpy(), Xd[2, -1].cpu().numpy())
    np.testing.assert_array_equal(Up[2, :-1].cpu().numpy(), Ud[2, 1:].cpu().numpy())
    np.testing.assert_array_equal(Up[2, -1].cpu().numpy(), Ud[2, -1].cpu().numpy())
    # mode 0: failure keeps X, U, X_prev
    Xd2, Xp2 = t(X), t(X * 2)
    me.mpc_update(qp, 0, t(z), t(status, torch.int32), t(x_ref), Xd2, t(U), Xp2, t(U), cost, conv, None, None)
    assert conv.cpu().tolist() == [1, 1, 1, 0]
    np.testing.assert_array_equal(Xd2[3].cpu().numpy(), X[3]); np.testing.assert_array_equal(Xp2[3].cpu().numpy(), 2 * X[3])


def test_unscented_propagation_matches_reference(E):
    """UncertaintyPropagator(method="unscented") (uncertainty_prop.py:179-264) on the device against the golden output of
    the reference class and the oracle.  alpha = 1e-3 gives w_m[0] ~ -1e6: last-bit differences of the GP means (1e-13
    relative, different summation order of K* alpha) are amplified to ~1e-7 in the mean and ~1e-7 relative in Sigma, hence
    the tolerances 2e-6 absolute / 1e-5 relative."""
    from oracle import gp as ogp, mpc as ompc
    from gp_mpc_rocket_landing_b200.dynamics import Rocket3DoFDynamics
    from gp_mpc_rocket_landing_b200.gp import Simple3DoFGP
    from gp_mpc_rocket_landing_b200.mpc.uncertainty_prop import UncertaintyPropagator
    g = load_golden("propagation_unscented.npz")
    gp = Simple3DoFGP(noise_variance=float(g["noise"]), use_sparse=False)
    gp.add_data(g["X"], g["U"], g["D"]); gp.fit()
    up = UncertaintyPropagator(Rocket3DoFDynamics(), gp, method="unscented")
    r = up.propagate(g["x0"], g["Useq"], np.eye(7) * 1e-6, dt=0.1)
    np.testing.assert_allclose(r.means, g["means"], rtol=0, atol=2e-6)
    np.testing.assert_allclose(r.covariances, g["covs"], rtol=1e-5, atol=1e-5 * np.abs(g["covs"]).max())
    # batch: perturbed initial states / controls against the oracle, scenario by scenario
    rng = np.random.default_rng(4)
    B, N = 5, g["Useq"].shape[0]
    x0 = np.tile(g["x0"], (B, 1)); U = np.tile(g["Useq"], (B, 1, 1))
    x0[1:] += 0.05 * rng.normal(size=(B - 1, 7)); U[1:] += 0.05 * rng.normal(size=(B - 1, N, 3))
    means, covs, dmean, dvar = up.propagate_batch(x0, U, None, 0.1)
    means, covs = means.cpu().numpy(), covs.cpu().numpy()
    np.testing.assert_allclose(means[0], r.means, rtol=0, atol=0)
    o = ogp.Simple3DoFGPOracle(noise_variance=float(g["noise"]))
    o.add_data(g["X"], g["U"], g["D"]); o.fit()
    for b in (1, 4):
        mo, co = ompc.propagate_unscented(x0[b], U[b], np.eye(7) * 1e-6, 0.1, ompc.Rocket3DoFParams(), o.predict)
        np.testing.assert_allclose(means[b], mo, rtol=0, atol=2e-6)
        np.testing.assert_allclose(covs[b], co, rtol=1e-5, atol=1e-5 * np.abs(co).max())
    # no GP: the transform of the nominal model alone
    r0 = UncertaintyPropagator(Rocket3DoFDynamics(), None, method="unscented").propagate(g["x0"], g["Useq"], None, dt=0.1)
    mo, co = ompc.propagate_unscented(g["x0"], g["Useq"], None, 0.1, ompc.Rocket3DoFParams(),
                                      lambda x, u: (np.zeros(3), np.zeros(3)))
    np.testing.assert_allclose(r0.means, mo, rtol=0, atol=1e-9)
    np.testing.assert_allclose(r0.covariances, co, rtol=1e-7, atol=1e-16)
    with pytest.raises(Exception):
        UncertaintyPropagator(Rocket3DoFDynamics(), gp, method="no_such_method").propagate(g["x0"], g["Useq"])


def test_monte_carlo_propagation_matches_reference(E):
    """UncertaintyPropagator(method="monte_carlo") (uncertainty_prop.py:266-319): with the reference's own particles and
    GP-posterior draws injected (golden: the reference class, seeded global RNG, replayed) the device path reproduces its
    sample means and covariances; with its own Philox draws the moments agree with the linear method within Monte Carlo
    error."""
    from gp_mpc_rocket_landing_b200.dynamics import Rocket3DoFDynamics
    from gp_mpc_rocket_landing_b200.gp import Simple3DoFGP
    from gp_mpc_rocket_landing_b200.mpc.uncertainty_prop import UncertaintyPropagator
    g = load_golden("propagation_mc.npz")
    gp = Simple3DoFGP(noise_variance=float(g["noise"]), use_sparse=False)
    gp.add_data(g["X"], g["U"], g["D"]); gp.fit()
    up = UncertaintyPropagator(Rocket3DoFDynamics(), gp, method="monte_carlo")
    N, S = g["xi"].shape[0], g["xi"].shape[1]
    B = 3     # three copies of the scenario: every warp must give the same answer
    x0 = np.tile(g["x0"], (B, 1)); U = np.tile(g["Useq"], (B, 1, 1))
    p0 = np.tile(g["particles0"][None], (B, 1, 1)); xi = np.tile(g["xi"][None].transpose(0, 1, 2, 3), (B, 1, 1, 1))
    means, covs, dmean, dvar = up._propagate_monte_carlo_batch(x0, U, float(g["sigma0"]), 0.1, particles0=p0, xi=xi)
    means, covs = means.cpu().numpy(), covs.cpu().numpy()
    for b in range(B):
        np.testing.assert_allclose(means[b], g["means"], rtol=1e-10, atol=1e-11)
        np.testing.assert_allclose(covs[b][1:], g["covs"][1:], rtol=1e-7, atol=1e-15)
    np.testing.assert_array_equal(means[0], means[2])
    assert np.all(dvar.cpu().numpy() > 0) and dmean.shape == (B, N, 3)
    # own random numbers: reproducible per seed, and statistically consistent with the linear propagation
    r1 = up._propagate_monte_carlo_batch(x0[:1], U[:1], 1e-6, 0.1, n_samples=400, seed=7)
    r2 = up._propagate_monte_carlo_batch(x0[:1], U[:1], 1e-6, 0.1, n_samples=400, seed=7)
    assert np.array_equal(r1[0].cpu().numpy(), r2[0].cpu().numpy())
    lin = UncertaintyPropagator(Rocket3DoFDynamics(), gp, method="linear").propagate(g["x0"], g["Useq"], None, dt=0.1)
    m_mc, c_mc = r1[0][0].cpu().numpy(), r1[1][0].cpu().numpy()
    sd = np.sqrt(np.maximum(np.diagonal(lin.covariances, axis1=1, axis2=2), 1e-12))
    assert np.all(np.abs(m_mc - lin.means) < 6.0 * sd / np.sqrt(400) + 1e-6)
    vmc, vlin = np.diagonal(c_mc, axis1=1, axis2=2)[-1, 4:7], np.diagonal(lin.covariances, axis1=1, axis2=2)[-1, 4:7]
    # same order of magnitude only: the particles see the GP MEAN's dependence on the state (a drag-like residual contracts
    # the velocity spread), which the linear method's nominal Jacobian ignores -- in the reference as well
    assert np.all(vmc < 3.0 * vlin) and np.all(vmc > 0.3 * vlin)
    one = up.propagate(g["x0"], g["Useq"], np.eye(7) * 1e-6, dt=0.1)      # the reference signature (own draws)
    assert one.means.shape == (N + 1, 7) and one.covariances.shape == (N + 1, 7, 7)


def test_tube_recursion_batched(E):
    """gpmpc_propagate_tube (e_{k+1} = |A_k| e_k + w_k) against the host loop of uncertainty_prop.py:435-470, and the
    GP-variance tubes of safety/tube_mpc.py:118-183 (w = n_sigma sqrt(var)) against the same loop fed the oracle variances."""
    from oracle import gp as ogp, mpc as ompc
    from gp_mpc_rocket_landing_b200 import workloads
    from gp_mpc_rocket_landing_b200.dynamics import Rocket3DoFDynamics
    from gp_mpc_rocket_landing_b200.gp import Simple3DoFGP
    from gp_mpc_rocket_landing_b200.mpc.uncertainty_prop import TubeBasedRobustness
    ge, me = E
    rng = np.random.default_rng(6)
    B, N = 6, 12
    X0, _ = workloads.sample_initial_conditions(0, B)
    X = np.zeros((B, N + 1, 7)); U = np.zeros((B, N, 3))
    p = ompc.Rocket3DoFParams()
    for b in range(B):
        X[b, 0] = X0[b]
        for k in range(N):
            U[b, k] = [X0[b, 0] + 0.2 * rng.normal(), 0.2 * rng.normal(), 0.2 * rng.normal()]
            X[b, k + 1] = ompc.step_3dof(X[b, k], U[b, k], 0.1, p)
    dyn = Rocket3DoFDynamics()
    tb = TubeBasedRobustness(dyn, d_max=0.3)
    T = tb.compute_tube_batch(X, U, 0.1).cpu().numpy()
    for b in range(B):
        np.testing.assert_allclose(T[b], tb.compute_tube(X[b], U[b], 0.1), rtol=1e-13, atol=1e-16)
    Xr, Ur, Dr = workloads.residual_dataset(200, seed=2)
    gp = Simple3DoFGP(noise_variance=1e-2, use_sparse=False); gp.add_data(Xr, Ur, Dr); gp.fit()
    o = ogp.Simple3DoFGPOracle(noise_variance=1e-2); o.add_data(Xr, Ur, Dr); o.fit()
    Tg, var = tb.propagate_with_gp_batch(X, U, 0.1, gp, n_sigma=2.0)
    Tg, var = Tg.cpu().numpy(), var.cpu().numpy()
    for b in (0, 3, 5):
        e = np.zeros(7)
        for k in range(N):
            A, _ = ompc.linearize_3dof(X[b, k], U[b, k], 0.1, p)
            _, v = o.predict(X[b, k], U[b, k])
            np.testing.assert_allclose(var[b, k], v, rtol=1e-8)
            w = np.zeros(7); w[4:7] = 2.0 * np.sqrt(v)
            e = np.abs(A) @ e + w
            np.testing.assert_allclose(Tg[b, k + 1], e, rtol=1e-8, atol=1e-14)

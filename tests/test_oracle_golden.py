"""CPU: the numpy oracle (oracle/) against golden vectors produced by the REAL reference
(oracle/gen_golden.py, run in the build container).  This is what pins the oracle."""
import numpy as np
import pytest

from conftest import load_golden, relerr
from oracle import gp as ogp
from oracle import mpc as ompc

KINDS = {"se": ogp.SE, "m32": ogp.M32, "m52": ogp.M52}


def test_kernels_match_reference():
    g = load_golden("kernels.npz")
    for name, kind in KINDS.items():
        # identical operation order -> bit-for-bit apart from BLAS dot order: allow 4 ulp
        np.testing.assert_allclose(ogp.gram(kind, g["X1"], None, g["ls"], float(g["sv"])), g[f"{name}_self"], rtol=1e-15, atol=1e-15)
        np.testing.assert_allclose(ogp.gram(kind, g["X1"], g["X2"], g["ls"], float(g["sv"])), g[f"{name}_cross"], rtol=1e-15, atol=1e-15)
    gr = ogp.se_ard_gradients(g["X1"], g["X2"], g["ls"], float(g["sv"]))
    for i, nm in enumerate(g["se_grad_names"]):
        np.testing.assert_allclose(gr[str(nm)], g["se_grad"][i], rtol=1e-14, atol=1e-15)


# n1e-4 is the reference default (cond(K+s2I) ~ 9.2e6): even numpy-vs-numpy (batched GEMM K* here vs the
# reference's one-query-at-a-time GEMV) differs by 1.07e-9 on the mean, so the 1e-9 bar is applied on the
# well-conditioned fixture and scaled by cond/1e6 (-> 1e-8) on the default one.
@pytest.mark.parametrize("tag,noise,tol", [("n1e-2", 1e-2, 1e-9), ("n1e-4", 1e-4, 1e-8)])
def test_simple3dof_exact_matches_reference(tag, noise, tol):
    g = load_golden("simple3dof_exact.npz")
    o = ogp.Simple3DoFGPOracle(noise_variance=noise)
    o.add_data(g["X"], g["U"], g["D"]); o.fit()
    np.testing.assert_allclose(ogp.features_batch(ogp.features_3dof, g["X"], g["U"]), g["Z_train"], rtol=0, atol=0)
    m, v = o.predict_batch(g["Xq"], g["Uq"])
    assert relerr(m, g[f"{tag}_mean"]) < tol
    assert relerr(v, g[f"{tag}_var"]) < tol
    for i, f in enumerate(o.fits):
        assert relerr(f.lml, g[f"{tag}_lml"][i]) < 1e-12
        assert f.y_mean == g[f"{tag}_ymean"][i] and f.y_std == g[f"{tag}_ystd"][i]
    np.testing.assert_allclose(np.diag(o.fits[0].L), g[f"{tag}_Ldiag"], rtol=1e-12)
    m1, v1 = o.predict(g["Xq"][0], g["Uq"][0])
    assert relerr(m1, g[f"{tag}_mean"][0]) < tol and relerr(v1, g[f"{tag}_var"][0]) < tol


def test_exact_gp_all_kernels():
    g = load_golden("exact_gp.npz")
    for name, kind in KINDS.items():
        f = ogp.exact_fit(kind, g["X"], g["y"], g["ls"], float(g["sv"]), float(g["noise"]))
        m, v, s = ogp.exact_predict(f, g["Xt"])
        assert relerr(m, g[f"{name}_mean"]) < 1e-10 and relerr(v, g[f"{name}_var"]) < 1e-10
        assert relerr(s, g[f"{name}_std"]) < 1e-10
        np.testing.assert_allclose(f.L, g[f"{name}_L"], rtol=1e-10, atol=1e-13)
        assert relerr(f.lml, g[f"{name}_lml"]) < 1e-12
        mc, cov = ogp.exact_predict(f, g["Xt"], return_cov=True)
        np.testing.assert_allclose(cov, g[f"{name}_cov"], rtol=1e-8, atol=1e-12)
    f = ogp.exact_fit(ogp.SE, g["X"], g["y"], g["ls"], float(g["sv"]), float(g["noise"]), normalize_y=False)
    m, v, _ = ogp.exact_predict(f, g["Xt"])
    assert relerr(m, g["se_nonorm_mean"]) < 1e-10 and relerr(v, g["se_nonorm_var"]) < 1e-10
    assert relerr(f.lml, g["se_nonorm_lml"]) < 1e-12


def test_exact_gp_jitter_escalation():
    g = load_golden("exact_gp.npz")
    f = ogp.exact_fit(ogp.SE, g["jit_X"], g["jit_y"], g["jit_ls"], float(g["sv"]), float(g["jit_noise"]))
    assert f.jitter > 0.0, "fixture must exercise the jitter branch (exact_gp.py:163-175)"
    np.testing.assert_allclose(np.diag(f.L), g["jit_Ldiag"], rtol=1e-9)
    assert relerr(f.lml, g["jit_lml"]) < 1e-9


def test_exact_lml_gradient_closed_form_vs_central_differences():
    """The closed-form LML gradient (kernel derivatives pinned by test_kernels_match_reference) against central
    differences of the reference LML restatement -- what scipy's L-BFGS-B differences in exact_gp.py:357-421."""
    rng = np.random.default_rng(0)
    X = rng.normal(size=(60, 4))
    y = np.sin(X[:, 0]) + 0.1 * rng.normal(size=60)
    th = np.log(np.array([1.3, 0.8, 1.1, 2.0, 0.6, 3e-2]))

    def fit(t):
        return ogp.exact_fit(0, X, y, np.exp(t[1:5]), np.exp(t[0]), np.exp(t[5]))

    g = ogp.exact_lml_gradient(fit(th))
    h = 1e-5
    fd = np.array([(fit(th + h * e).lml - fit(th - h * e).lml) / (2 * h) for e in np.eye(6)])
    np.testing.assert_allclose(g, fd, rtol=1e-7, atol=1e-7)


def test_sparse_gp():
    g = load_golden("sparse_gp.npz")
    for method in ("fitc", "vfe"):
        f = ogp.sparse_fit(ogp.SE, g["X"], g["y"], g["Z"], g["ls"], float(g["sv"]), float(g["noise"]), method)
        m, v, _ = ogp.sparse_predict(f, g["Xt"])
        assert relerr(m, g[f"{method}_mean"]) < 1e-9 and relerr(v, g[f"{method}_var"]) < 1e-9
        np.testing.assert_allclose(f.L_uu, g[f"{method}_Luu"], rtol=1e-9, atol=1e-13)
        np.testing.assert_allclose(f.L_B, g[f"{method}_LB"], rtol=1e-9, atol=1e-12)
        assert relerr(f.lml, g[f"{method}_lml"]) < 1e-10
    o = ogp.Simple3DoFGPOracle(noise_variance=1e-4, use_sparse=True, Z=g["s3_Z"])
    o.add_data(g["s3_X"], g["s3_U"], g["s3_D"]); o.fit()
    m, v = o.predict_batch(g["s3_Xq"], g["s3_Uq"])
    assert relerr(m, g["s3_mean"]) < 1e-8 and relerr(v, g["s3_var"]) < 1e-8


def test_fast_gp():
    g = load_golden("fast_gp.npz")
    L, alpha = ogp.fast_fit(g["X"], g["Y"], g["ls"], float(g["variance"]), float(g["noise"]))
    m, v = ogp.fast_predict(g["X"], L, alpha, g["ls"], float(g["variance"]), g["Xt"])
    np.testing.assert_allclose(m, g["mean"], rtol=1e-10, atol=1e-12)
    np.testing.assert_allclose(v, g["var"], rtol=1e-9, atol=1e-13)


def test_features():
    g = load_golden("features.npz")
    np.testing.assert_allclose(ogp.features_batch(ogp.features_3dof, g["X14"][:, :7], g["U"]), g["f3"], rtol=0, atol=0)
    np.testing.assert_allclose(ogp.features_batch(ogp.features_6dof_trans, g["X14"], g["U"]), g["ft"], rtol=0, atol=0)
    np.testing.assert_allclose(ogp.features_batch(ogp.features_6dof_rot, g["X14"], g["U"]), g["fr"], rtol=0, atol=0)


def test_qp_assembly_matches_reference():
    g = load_golden("qp_assembly.npz")
    p = ompc.Rocket3DoFParams()
    x0, xt = g["x0"], g["xt"]
    for tag, N, lin in (("fast20", 20, "analytic"), ("fast15", 15, "analytic"), ("fd20", 20, "fd")):
        n, m = ompc.qp_dims(N)
        Pd, q = ompc.build_cost(N, np.tile(xt, (N + 1, 1)))
        np.testing.assert_array_equal(np.diag(Pd), g[f"{tag}_P"])
        np.testing.assert_array_equal(q, g[f"{tag}_q"])
        X_lin = np.array([(1 - k / N) * x0 + (k / N) * xt for k in range(N + 1)])
        U_lin = np.zeros((N, 3)); U_lin[:, 0] = x0[0] * 1.0
        A, l, u = ompc.build_constraints(N, X_lin, U_lin, x0, 0.1, p, lin)
        assert A.shape == (m, n) == g[f"{tag}_A"].shape
        np.testing.assert_array_equal(A, g[f"{tag}_A"])
        np.testing.assert_array_equal(l, g[f"{tag}_l"]); np.testing.assert_array_equal(u, g[f"{tag}_u"])
        assert int(np.count_nonzero(A)) == int(g[f"{tag}_nnzA"])
        A2, l2, u2 = ompc.build_constraints(N, g[f"{tag}_Xl"], g[f"{tag}_Ul"], g[f"{tag}_xc"], 0.1, p, lin)
        np.testing.assert_array_equal(A2, g[f"{tag}_A2"])
        np.testing.assert_array_equal(l2, g[f"{tag}_l2"]); np.testing.assert_array_equal(u2, g[f"{tag}_u2"])
    assert ompc.qp_dims(20) == (207, 354) and ompc.qp_dims(15) == (157, 269)
    assert int(g["fast20_nnzA"]) == 654 and int(g["fast20_nnzP"]) == 186
    np.testing.assert_array_equal(g["fast20_settings"], [50, 1e-4, 1e-4, 3, 0.0, 1.0])


def test_propagation_matches_reference():
    g = load_golden("propagation.npz")
    o = ogp.Simple3DoFGPOracle(noise_variance=float(g["noise"]))
    o.add_data(g["X"], g["U"], g["D"]); o.fit()
    means, covs, _, _ = ompc.propagate_linear(g["x0"], g["Useq"], np.eye(7) * 1e-6, 0.1, ompc.Rocket3DoFParams(), o.predict)
    np.testing.assert_allclose(means, g["means"], rtol=1e-11, atol=1e-12)
    np.testing.assert_allclose(covs, g["covs"], rtol=1e-9, atol=1e-18)


def test_unscented_propagation_matches_reference():
    """uncertainty_prop.py:179-264 through the reference's own class (14-state embedding) against the 7-state
    restatement: identical term order, so the match is to the last bit although w_m[0] ~ -1e6."""
    g = load_golden("propagation_unscented.npz")
    o = ogp.Simple3DoFGPOracle(noise_variance=float(g["noise"]))
    o.add_data(g["X"], g["U"], g["D"]); o.fit()
    means, covs = ompc.propagate_unscented(g["x0"], g["Useq"], np.eye(7) * 1e-6, 0.1, ompc.Rocket3DoFParams(), o.predict)
    np.testing.assert_allclose(means, g["means"], rtol=0, atol=1e-9)
    np.testing.assert_allclose(covs, g["covs"], rtol=1e-7, atol=1e-16)
    assert float(g["cov_cross_max"]) < 1e-12     # the inert block of the 14-state class never couples with the first 7


def test_monte_carlo_sampling_and_landing():
    g = load_golden("monte_carlo.npz")
    cfg = ompc.SimConfig()
    ics = np.stack([ompc.sample_initial_condition(42 + i, cfg) for i in range(64)])
    np.testing.assert_array_equal(ics, g["ics"])
    ok = np.array([ompc.check_landing(s, 2.0) for s in g["states"]])
    np.testing.assert_array_equal(ok, g["landing_ok"])
    assert ok.any() and (~ok).any()


def test_mc_loop_matches_reference_run_single():
    """oracle.mpc.run_single against the reference's own `MonteCarloSimulator.run_single` (monte_carlo.py:401-583) driven
    with the same deterministic stub controllers: outcome class, final state, number of plant steps and fuel for 32 seeds,
    both controller kinds (`.solve`: per-step target + success check; `.step`: fixed target, flag ignored)."""
    g = load_golden("mc_loop.npz")
    assert set(np.unique(g["solve_outcome"])) == {1, 2, 3, 4, 5, 6}     # every LandingOutcome class is exercised
    for kind, tm, pol in (("solve", "per_step", "divergence"), ("step", "fixed", "continue")):
        for i in range(32):
            gains, fail_step = ompc.pd_test_gains(i)
            k = [0]
            targets = []

            def solve(x, xt):
                targets.append(xt.copy())
                ok = k[0] != fail_step
                k[0] += 1
                return ompc.pd_test_law(x, xt, gains), ok

            code, xf, nst, _ = ompc.run_single(solve, g["x0"][i], dt=0.1, max_time=20.0, target_mode=tm,
                                               on_solve_failure=pol)
            assert code == g[kind + "_outcome"][i], (kind, i)
            assert nst == g[kind + "_nsteps"][i], (kind, i)
            np.testing.assert_array_equal(xf, g[kind + "_final"][i])
            assert g["x0"][i][0] - xf[0] == g[kind + "_fuel"][i]
            if kind == "solve":
                t = np.array(targets[:3])
                np.testing.assert_array_equal(t, g["solve_targets"][i][:len(t)])


ARBITER_K = 4.0   # both implementations must sit within ARBITER_K * cond(K + noise I) * eps of the long-double posterior


@pytest.mark.parametrize("tag,noise", [("n1e-4", 1e-4), ("n1e-2", 1e-2)])
def test_reference_and_oracle_within_conditioning_of_longdouble_truth(tag, noise):
    """at the reference's default hyper-parameters cond(K + noise I) ~ 9e6, so the reference's OWN FP64
    posterior is only accurate to ~cond*eps = 2e-9 (measured: mean 3.8e-9, variance 1.8e-10 against the x87 long-double
    arbiter) -- north_star's 1e-9 "vs the reference" cannot be met by any second FP64 implementation there.  The bar
    that can: every implementation within ARBITER_K*cond*eps of the arbiter.  Here: the reference (golden) and the numpy
    oracle; tests/test_gpu_gp.py applies the same bound to both CUDA variance paths.  Also pins the arbiter itself
    (`longdouble_exact_predict`) against its committed output on a subset of the queries."""
    g = load_golden("simple3dof_exact.npz")
    a = load_golden("simple3dof_arbiter.npz")
    bound = ARBITER_K * float(a[f"{tag}_cond"]) * np.finfo(float).eps
    tm, tv = a[f"{tag}_truth_mean"], a[f"{tag}_truth_var"]
    assert relerr(g[f"{tag}_mean"], tm) < bound and relerr(g[f"{tag}_var"], tv) < bound
    fits = [ogp.exact_fit(ogp.SE, g["Z_train"], g["D"][:, p], np.ones(11), 1.0, noise) for p in range(3)]
    mo, vo = ogp.multi_exact_predict(fits, g["Z_query"])
    assert relerr(mo, tm) < bound and relerr(vo, tv) < bound
    if tag == "n1e-4":
        assert float(a[f"{tag}_cond"]) > 5e6 and relerr(g[f"{tag}_mean"], tm) > 1e-9   # the reference itself misses 1e-9
    # the arbiter reproduces its fixture (4 queries, one output: the long-double Cholesky at n = 1000 takes ~2 s)
    m, v = ogp.longdouble_exact_predict(fits[1], g["Z_query"][:4])
    np.testing.assert_array_equal(m.astype(np.float64), tm[:4, 1])
    np.testing.assert_array_equal(v.astype(np.float64), tv[:4, 1])


def test_safety_filter_matches_reference():
    """oracle.safety against the reference's own PredictiveSafetyFilter (safety_filter.py:184-361, gradient path) on 2 x 48
    seeded cases: same decisions, filtered controls and margins (the reference ran on a 14-state embedding)."""
    from oracle import safety as osf
    g = load_golden("safety_filter.npz")
    for tag in ("a25", "a100"):
        s = osf.SafetySetup(K=g[f"{tag}_K"], P=g[f"{tag}_P"], x_eq=g[f"{tag}_x_eq"], u_eq=g[f"{tag}_u_eq"],
                            alpha=float(g[f"{tag}_alpha"]))
        assert float(g[f"{tag}_Kfull_rest"]) == 0.0          # the inert states never feed back
        mod = g[f"{tag}_modified"]
        assert mod.any() and (~mod).any() and (tag != "a25" or (g[f"{tag}_safe"] & mod).any())
        for i in range(g[f"{tag}_X"].shape[0]):
            u, m, ok, margin, bv = osf.safety_filter(s, g[f"{tag}_X"][i], g[f"{tag}_U"][i])
            assert m == bool(mod[i]) and ok == bool(g[f"{tag}_safe"][i]), (tag, i)
            np.testing.assert_allclose(u, g[f"{tag}_u_safe"][i], rtol=1e-9, atol=1e-12)
            np.testing.assert_allclose(margin, g[f"{tag}_margin"][i], rtol=1e-8, atol=1e-9)
            np.testing.assert_allclose(bv, g[f"{tag}_backup_value"][i], rtol=1e-8, atol=1e-9)


def test_monte_carlo_propagation_matches_reference():
    """uncertainty_prop.py:266-319 through the reference's own class (14-state embedding, seeded global RNG) against the
    restatement fed with the same particles and posterior noise."""
    g = load_golden("propagation_mc.npz")
    o = ogp.Simple3DoFGPOracle(noise_variance=float(g["noise"]))
    o.add_data(g["X"], g["U"], g["D"]); o.fit()
    means, covs = ompc.propagate_monte_carlo(g["x0"], g["Useq"], np.eye(7) * float(g["sigma0"]), 0.1, ompc.Rocket3DoFParams(),
                                             o.predict, g["particles0"], g["xi"])
    np.testing.assert_allclose(means, g["means"], rtol=1e-12, atol=1e-13)
    np.testing.assert_allclose(covs, g["covs"], rtol=1e-9, atol=1e-16)


def test_reference_gp_package_under_oracle_ref_agrees_with_the_restatement():
    """oracle/_ref (placed by oracle/make_ref.py from /root/reference/src/gp; travels to the GPU box, never committed) is the
    GP leg of bench.py's CPU arms: the reference's own Simple3DoFGP and the numpy restatement give the same predictions
    on the bench fixture (bit-compatible operation order; the 1e-4 noise fixture has cond 9e6, hence 1e-8)."""
    from oracle import ref_gp
    from gp_mpc_rocket_landing_b200 import workloads
    ref = ref_gp.load()
    if ref is None:
        pytest.skip("oracle/_ref is absent (run python oracle/make_ref.py where /root/reference exists)")
    X, U, D = workloads.residual_dataset(300, seed=0)
    g = ref.Simple3DoFGP(use_sparse=False, noise_variance=1e-4); g.add_data(X, U, D); g.fit()
    o = ogp.Simple3DoFGPOracle(noise_variance=1e-4); o.add_data(X, U, D); o.fit()
    rng = np.random.default_rng(5)
    for _ in range(5):
        x = np.array([2.0, 30, 0, 0, -3, 0, 0]) + rng.normal(size=7) * [0.1, 5, 2, 2, 1, 0.3, 0.3]
        u = np.array([2.0, 0, 0]) + 0.3 * rng.normal(size=3)
        (m1, v1), (m2, v2) = g.predict(x, u), o.predict(x, u)
        np.testing.assert_allclose(m1, m2, rtol=1e-8, atol=1e-12)
        np.testing.assert_allclose(v1, v2, rtol=1e-8, atol=1e-14)


@pytest.mark.parametrize("tag", ["vertical_it50", "lateral_it1000", "readme_infeasible_it1000"])
def test_rti_orchestration_matches_the_reference_closed_loop(tag):
    """a22: the reference's own FastRTI3DoF.step loop (osqp_rti.py:403-599) run in closed loop around a solver-backed stand-in
    for `osqp` (oracle/gen_golden.py::rti_loop_fixture) -- first step at max_iter 50 falls back to the previous trajectory,
    a lateral scenario whose sparsity pattern grows from 654 to 734 non-zeros, and the README initial state whose QP is
    primal infeasible at every step (:545-552 fall-back).  `OSQPRTIOracle` must take the same branches with the same iterates."""
    g = load_golden("rti_loop.npz")
    oc = ompc.OSQPRTIOracle(ompc.Rocket3DoFParams(), N=20, dt=0.1, lin="analytic",
                            settings=ompc.OSQPSettings(max_iter=int(g[f"{tag}_max_iter"])))
    oc.initialize(g[f"{tag}_x0"], g[f"{tag}_xt"])
    for k in range(6):
        r = oc.step(g[f"{tag}_x"][k])
        assert r.success == bool(g[f"{tag}_success"][k]) and r.osqp_iterations == int(g[f"{tag}_iters"][k]), (tag, k)
        np.testing.assert_allclose(r.u0, g[f"{tag}_u0"][k], rtol=0, atol=1e-11)
        np.testing.assert_allclose(r.X_opt, g[f"{tag}_X_opt"][k], rtol=0, atol=1e-10)
        np.testing.assert_allclose(r.U_opt, g[f"{tag}_U_opt"][k], rtol=0, atol=1e-10)
        if r.success:
            assert r.cost == pytest.approx(float(g[f"{tag}_cost"][k]), rel=1e-10)
        else:
            assert np.isinf(g[f"{tag}_cost"][k]) and np.isinf(r.cost)
    if tag == "lateral_it1000":
        assert g[f"{tag}_nnzA"].tolist() == [654] + [734] * 5


def test_gpmpc_sqp_skeleton_matches_the_reference_solve():
    """a17: the reference's own `GPMPC.solve` (gp_mpc.py:229-369) over three consecutive control steps with only `_solve_qp`
    (CasADi/IPOPT, absent) replaced by the OSQP-form QP built from the A / B / c lists the reference hands over
    (oracle/gen_golden.py::gpmpc_skeleton_fixture).  `GPMPCOracle` must produce the same sequence: 10 SQP iterations without
    meeting the 1e-4 step-size test (`success` False, as the reference), the same QP status / iteration count at each of
    the 30 QPs, the same iterates."""
    g = load_golden("gpmpc_skeleton.npz")
    o = ogp.Simple3DoFGPOracle(noise_variance=float(g["noise"])); o.add_data(g["X"], g["U"], g["D"]); o.fit()
    ctl = ompc.GPMPCOracle(o.predict, N=20, dt=0.1, settings=ompc.OSQPSettings(max_iter=1000), max_sqp_iter=10)
    k0 = 0
    for s in range(3):
        r = ctl.solve(g["x"][s], g["xt"][s])
        n = int(g["qp_n"][s])
        assert r.success == bool(g["success"][s]) and r.iterations == int(g["iterations"][s]) == n
        assert list(r.qp_status) == g["qp_status"][k0:k0 + n].tolist()
        assert list(r.qp_iters) == g["qp_iters"][k0:k0 + n].tolist()
        k0 += n
        np.testing.assert_allclose(r.U_opt, g["U_opt"][s], rtol=0, atol=1e-8)
        np.testing.assert_allclose(r.X_opt, g["X_opt"][s], rtol=0, atol=1e-8)
        assert r.cost == pytest.approx(float(g["cost"][s]), rel=1e-9)


def test_gpmpc_subproblem_rows_match_the_reference_arguments():
    """a18: the first sub-problem of the skeleton fixture, built there from the A_list / B_list / c_list the reference's
    `GPMPC.solve` passes to `_solve_qp` (deviation form of gp_mpc.py:402-411 in absolute variables) -- against
    `build_constraints(mode="gpmpc")`, which derives the same rows from (X_nom, U_nom, c) by itself."""
    g = load_golden("gpmpc_skeleton.npz")
    p = ompc.Rocket3DoFParams()
    A, l, u = ompc.build_constraints(20, g["qp0_X"], g["qp0_U"], g["qp0_x0"], 0.1, p, "analytic", g["qp0_c"], mode="gpmpc")
    np.testing.assert_array_equal(A != 0, g["qp0_A"] != 0)
    np.testing.assert_allclose(A, g["qp0_A"], rtol=0, atol=1e-15)
    np.testing.assert_allclose(l[:147], g["qp0_l"][:147], rtol=0, atol=1e-13)
    np.testing.assert_array_equal(l[147:], g["qp0_l"][147:]); np.testing.assert_array_equal(u[147:], g["qp0_u"][147:])
    Pd, q = ompc.build_cost(20, g["qp0_xref"])
    np.testing.assert_array_equal(Pd, g["qp0_Pd"]); np.testing.assert_array_equal(q, g["qp0_q"])


def test_reference_gpmpc_solve_from_oracle_ref_reproduces_the_skeleton_fixture():
    """bench.py's CPU leg at S = 10 runs the reference's own `GPMPC.solve` out of oracle/_ref (oracle/ref_controller.py).
    It is the construction that generated gpmpc_skeleton.npz, so it must reproduce it exactly."""
    from oracle import ref_controller, ref_gp
    ref, gm = ref_gp.load(), ref_gp.load_gpmpc()
    if ref is None or gm is None:
        pytest.skip("oracle/_ref is absent (run python oracle/make_ref.py where /root/reference exists)")
    g = load_golden("gpmpc_skeleton.npz")
    gp = ref.Simple3DoFGP(use_sparse=False, noise_variance=float(g["noise"])); gp.add_data(g["X"], g["U"], g["D"]); gp.fit()
    c = ref_controller.ReferenceGPMPC(gm, gp.predict)
    for s in range(3):
        u0, ok, it = c.solve(g["x"][s], g["xt"][s])
        assert (ok, it) == (bool(g["success"][s]), int(g["iterations"][s]))
        np.testing.assert_array_equal(u0, g["U_opt"][s][0])
    assert [i for _, i in c.qp_log] == g["qp_iters"].tolist()

"""GPU parity: GP kernels through the C-ABI (libgpmpc_b200.so) against the oracle and the golden vectors
generated from the real reference.  FP64 tolerances are written next to each check; the 1e-9 relative bar of
north_star applies on well-conditioned fixtures and is scaled by cond(K)/1e6 on the reference-default one
(sigma_n^2 = 1e-4, cond ~ 9e6; even numpy-vs-numpy differs by 1.07e-9 there -- SURVEY H1)."""
import numpy as np
import pytest

from conftest import load_golden, relerr

pytestmark = pytest.mark.gpu

KINDS = {"se": 0, "m32": 1, "m52": 2}


@pytest.fixture(scope="module")
def eng():
    from gp_mpc_rocket_landing_b200.gp import engine
    engine.torch_()
    return engine


def test_gram_matches_reference(eng):
    g = load_golden("kernels.npz")
    sv = float(g["sv"])
    for name, kind in KINDS.items():
        K = eng.gram(kind, g["X1"], None, g["ls"], sv).cpu().numpy()
        Kc = eng.gram(kind, g["X1"], g["X2"], g["ls"], sv).cpu().numpy()
        # same operation order as kernels.py:205-262; only the 11-term dot differs (DMMA vs BLAS): few ulp
        np.testing.assert_allclose(K, g[f"{name}_self"], rtol=4e-15, atol=4e-15)
        np.testing.assert_allclose(Kc, g[f"{name}_cross"], rtol=4e-15, atol=4e-15)
    for i, nm in enumerate(g["se_grad_names"]):
        nm = str(nm)
        if nm == "log_signal_variance":
            G = eng.gram(0, g["X1"], g["X2"], g["ls"], sv).cpu().numpy()
        else:
            G = eng.gram_grad_lengthscale(g["X1"], g["X2"], g["ls"], sv, int(nm.split("_")[-1])).cpu().numpy()
        np.testing.assert_allclose(G, g["se_grad"][i], rtol=1e-13, atol=1e-15)


def test_gram_large_and_ragged(eng):
    """Sizes that are not multiples of the 128 tile, empty-ish inputs, and symmetry / unit diagonal."""
    from oracle import gp as ogp
    rng = np.random.default_rng(3)
    for n1, n2, D in ((1, 1, 1), (129, 257, 11), (300, 5, 13), (1000, 640, 12)):
        X1, X2 = rng.normal(size=(n1, D)), rng.normal(size=(n2, D))
        ls = rng.uniform(0.5, 2.0, D)
        for kind in (0, 1, 2):
            K = eng.gram(kind, X1, X2, ls, 1.7).cpu().numpy()
            np.testing.assert_allclose(K, ogp.gram(kind, X1, X2, ls, 1.7), rtol=1e-13, atol=1e-15)
    X = rng.normal(size=(500, 11))
    K = eng.gram(0, X, None, np.ones(11), 1.0).cpu().numpy()
    assert np.array_equal(K, K.T) or np.max(np.abs(K - K.T)) < 1e-15
    assert np.max(np.abs(np.diag(K) - 1.0)) < 1e-14


def test_features_match_reference(eng):
    g = load_golden("features.npz")
    f3 = eng.features(0, g["X14"][:, :7].copy(), g["U"]).cpu().numpy()
    ft = eng.features(1, g["X14"], g["U"]).cpu().numpy()
    fr = eng.features(2, g["X14"], g["U"]).cpu().numpy()
    # element-wise double arithmetic in the reference order; exp/atan2/asin may differ from libm by 1-2 ulp
    np.testing.assert_allclose(f3, g["f3"], rtol=2e-15, atol=1e-17)
    np.testing.assert_allclose(ft, g["ft"], rtol=2e-15, atol=1e-16)
    np.testing.assert_allclose(fr, g["fr"], rtol=2e-15, atol=1e-16)


def test_exact_fit_predict_all_kernels(eng):
    g = load_golden("exact_gp.npz")
    n = g["X"].shape[0]
    for name, kind in KINDS.items():
        f = eng.ExactFactor(kind, g["ls"], float(g["sv"]), float(g["noise"])).fit(g["X"])
        assert f.jitter == 0.0
        m = eng.ExactModelDev(f).fit_targets(g["y"])
        L = f.L.cpu().numpy()
        np.testing.assert_allclose(L[:n, :n], g[f"{name}_L"], rtol=1e-9, atol=1e-12)
        assert np.array_equal(L[n:, :n], np.zeros_like(L[n:, :n])) and np.array_equal(L[n:, n:], np.eye(L.shape[0] - n))
        assert relerr(m.alpha.cpu().numpy()[0, :n], g[f"{name}_alpha"]) < 1e-7   # alpha ~ cond * eps
        assert relerr(m.lml.cpu().numpy()[0], g[f"{name}_lml"]) < 1e-11
        assert m.ymean.item() == pytest.approx(float(g[f"{name}_ymean"]), rel=1e-15)
        assert m.ystd.item() == pytest.approx(float(g[f"{name}_ystd"]), rel=1e-15)
        mean, var = m.predict(g["Xt"])
        assert relerr(mean.cpu().numpy()[:, 0], g[f"{name}_mean"]) < 1e-9
        assert relerr(var.cpu().numpy()[:, 0], g[f"{name}_var"]) < 1e-9
    f = eng.ExactFactor(0, g["ls"], float(g["sv"]), float(g["noise"])).fit(g["X"])
    m = eng.ExactModelDev(f).fit_targets(g["y"], normalize=False)
    mean, var = m.predict(g["Xt"])
    assert relerr(mean.cpu().numpy()[:, 0], g["se_nonorm_mean"]) < 1e-9
    assert relerr(var.cpu().numpy()[:, 0], g["se_nonorm_var"]) < 1e-9
    assert relerr(m.lml.cpu().numpy()[0], g["se_nonorm_lml"]) < 1e-11


def test_exact_fit_jitter_escalation(eng):
    """exact_gp.py:163-175: the fixture is not PD at noise alone; the same jitter step must be selected."""
    from oracle import gp as ogp
    g = load_golden("exact_gp.npz")
    fo = ogp.exact_fit(ogp.SE, g["jit_X"], g["jit_y"], g["jit_ls"], float(g["sv"]), float(g["jit_noise"]))
    f = eng.ExactFactor(0, g["jit_ls"], float(g["sv"]), float(g["jit_noise"])).fit(g["jit_X"])
    assert f.jitter == fo.jitter > 0.0
    m = eng.ExactModelDev(f).fit_targets(g["jit_y"])
    n = g["jit_X"].shape[0]
    np.testing.assert_allclose(np.diag(f.L.cpu().numpy())[:n], g["jit_Ldiag"], rtol=1e-6)
    assert relerr(m.lml.cpu().numpy()[0], g["jit_lml"]) < 1e-6


@pytest.mark.parametrize("tag,noise,tol", [("n1e-2", 1e-2, 1e-9), ("n1e-4", 1e-4, 1e-8)])
def test_simple3dof_exact_1000pt(eng, tag, noise, tol):
    """BASELINE config 1 GP: 1000 synthetic residual points, 3 outputs sharing one factor."""
    g = load_golden("simple3dof_exact.npz")
    Z = eng.features(0, g["X"], g["U"])
    np.testing.assert_allclose(Z.cpu().numpy(), g["Z_train"], rtol=2e-15, atol=1e-17)
    f = eng.ExactFactor(0, np.ones(11), 1.0, noise).fit(Z)
    m = eng.ExactModelDev(f).fit_targets(g["D"])
    Zq = eng.features(0, g["Xq"], g["Uq"])
    mean, var = m.predict(Zq)
    assert relerr(mean.cpu().numpy(), g[f"{tag}_mean"]) < tol
    assert relerr(var.cpu().numpy(), g[f"{tag}_var"]) < tol
    assert relerr(m.lml.cpu().numpy(), g[f"{tag}_lml"]) < 1e-10
    np.testing.assert_allclose(np.diag(f.L.cpu().numpy())[:1000], g[f"{tag}_Ldiag"], rtol=1e-9)
    np.testing.assert_allclose(f.L.cpu().numpy()[999, :1000], g[f"{tag}_Lrow999"], rtol=1e-7, atol=1e-12)


@pytest.mark.parametrize("tag,noise,tol", [("n1e-2", 1e-2, 1e-9), ("n1e-4", 1e-4, 1e-8)])
def test_explicit_inverse_variance_path(eng, tag, noise, tol):
    """The horizon pipeline's one-launch variance (|L^-1 k*|^2 against the explicit inverse factor) against the
    reference's own predictive variance (golden) and against the blocked in-place solve; Ks must stay untouched."""
    import torch
    g = load_golden("simple3dof_exact.npz")
    f = eng.ExactFactor(0, np.ones(11), 1.0, noise).fit(eng.to_dev(g["Z_train"]))
    m = eng.ExactModelDev(f).fit_targets(g["D"])
    W = f.ensure_linv()
    Lh, Wh = f.L.cpu().numpy()[:1000, :1000], W.cpu().numpy()
    assert np.max(np.abs(np.triu(Wh, 1))) == 0.0                      # exactly lower triangular
    assert np.max(np.abs(Wh[:1000, :1000] @ Lh - np.eye(1000))) < 1e-9
    Ks, Q = f.cross(eng.to_dev(g["Z_query"]))
    Ks0 = Ks.clone()
    ss_inv = f.sumsq_linv(Ks)
    assert torch.equal(Ks, Ks0)
    ss_trsm = f.sumsq_inplace(Ks)
    ystd = m.ystd.cpu().numpy()
    var_inv = np.maximum(1.0 - ss_inv.cpu().numpy()[:Q], 1e-10)[:, None] * ystd[None, :] ** 2
    var_trsm = np.maximum(1.0 - ss_trsm.cpu().numpy()[:Q], 1e-10)[:, None] * ystd[None, :] ** 2
    assert relerr(var_inv, g[f"{tag}_var"]) < tol
    assert relerr(var_inv, var_trsm) < tol
    # the arbiter (tests/test_oracle_golden.py::test_reference_and_oracle_within_conditioning_of_longdouble_truth holds the
    # reference to the same bound): mean and BOTH variance paths within 4 cond eps of the long-double posterior.  On the
    # default-hyper fixture (cond 9.2e6) that is 8.2e-9; the reference's own result is 3.8e-9 away from the truth there,
    # which is why the comparison WITH the reference above is held to 1e-8, not 1e-9.
    a = load_golden("simple3dof_arbiter.npz")
    bound = 4.0 * float(a[f"{tag}_cond"]) * np.finfo(float).eps
    mean, var_api = m.predict(eng.to_dev(g["Z_query"]))
    assert relerr(mean.cpu().numpy(), a[f"{tag}_truth_mean"]) < bound
    assert relerr(var_api.cpu().numpy(), a[f"{tag}_truth_var"]) < bound
    assert relerr(var_inv, a[f"{tag}_truth_var"]) < bound
    assert relerr(var_trsm, a[f"{tag}_truth_var"]) < bound


def test_update_with_mean_function_equals_fit(eng):
    """ExactGP.update with a prior mean (exact_gp.py:141-154: the stored targets have m(X) / y_std subtracted): extending
    the model must equal `fit` on the concatenated RAW data -- the old rows' targets are reconstructed with m(X) added back."""
    from gp_mpc_rocket_landing_b200.gp import ExactGP, SquaredExponentialARD
    rng = np.random.default_rng(3)
    D = 4
    X = rng.normal(size=(90, D)); y = 2.0 + np.sin(X[:, 0]) + 0.5 * X[:, 1] + 0.05 * rng.normal(size=90)
    Xq = rng.normal(size=(25, D))
    mean_fn = lambda Z: 2.0 + 0.5 * np.asarray(Z)[:, 1]   # noqa: E731
    inc = ExactGP(SquaredExponentialARD(D), noise_variance=1e-2, mean_function=mean_fn).fit(X[:60], y[:60])
    inc.update(X[60:], y[60:])
    full = ExactGP(SquaredExponentialARD(D), noise_variance=1e-2, mean_function=mean_fn).fit(X, y)
    assert inc.n_train == full.n_train == 90
    assert inc._y_mean == pytest.approx(full._y_mean, rel=1e-13) and inc._y_std == pytest.approx(full._y_std, rel=1e-13)
    np.testing.assert_allclose(inc.y_train, full.y_train, rtol=1e-12, atol=1e-13)
    pi, pf = inc.predict(Xq), full.predict(Xq)
    assert relerr(pi.mean, pf.mean) < 1e-10 and relerr(pi.variance, pf.variance) < 1e-9
    assert abs(inc.log_marginal_likelihood - full.log_marginal_likelihood) < 1e-9 * abs(full.log_marginal_likelihood)


def test_predict_batch_equals_singles_and_variance_floor(eng):
    """batch = loop of singles (bitwise: every query row is solved independently), floor 1e-10 * y_std^2."""
    rng = np.random.default_rng(5)
    X = rng.normal(size=(300, 11)); Y = rng.normal(size=(300, 3))
    f = eng.ExactFactor(0, np.ones(11), 1.0, 1e-6).fit(X)
    m = eng.ExactModelDev(f).fit_targets(Y)
    Xq = np.vstack([X[:5], rng.normal(size=(200, 11))])
    mean, var = m.predict(Xq)
    mean, var = mean.cpu().numpy(), var.cpu().numpy()
    for i in (0, 3, 77, 204):
        m1, v1 = m.predict(Xq[i:i + 1])
        assert np.array_equal(m1.cpu().numpy()[0], mean[i]) and np.array_equal(v1.cpu().numpy()[0], var[i])
    ystd = m.ystd.cpu().numpy()
    assert np.all(var >= 1e-10 * ystd**2 * (1 - 1e-15))


def test_refit_sweep_matches_sequential_fits(eng):
    """BASELINE config 3 (reduced): a batched hyper-parameter grid equals the loop of single fits (exact_gp.py:118-204),
    including per-candidate jitter escalation."""
    from gp_mpc_rocket_landing_b200.gp import refit
    from oracle import gp as ogp
    rng = np.random.default_rng(3)
    n, D, P = 300, 11, 3
    X = rng.normal(0, 0.5, (n, D))
    Y = np.stack([np.sin(X[:, 0] * 2) + 0.1 * rng.normal(size=n), X[:, 1] ** 2, X[:, 2] - X[:, 3]], axis=1)
    for kind in (0, 2):
        ls, sv, nz = refit.default_grid(D, n_lengthscales=3, noises=(1e-3, 1e-1))
        r = refit.refit_sweep(X, Y, ls, sv, nz, kind=kind)
        assert r["lml"].shape == (6, P)
        for g in range(6):
            for p in range(P):
                f = ogp.exact_fit(kind, X, Y[:, p], ls[g], sv[g], nz[g])
                assert relerr(r["lml"][g, p], f.lml) < 1e-9, (kind, g, p)
        assert r["best_shared"] == int(np.argmax(r["lml"].sum(1)))
    # duplicated rows + tiny noise: the first Cholesky fails, the candidate alone escalates its jitter
    g = load_golden("exact_gp.npz")
    fo = ogp.exact_fit(ogp.SE, g["jit_X"], g["jit_y"], g["jit_ls"], float(g["sv"]), float(g["jit_noise"]))
    sweep = eng.ExactSweep(0, g["jit_X"], g["jit_y"])
    out = sweep.fit(np.stack([g["jit_ls"], g["jit_ls"]]), float(g["sv"]), [float(g["jit_noise"]), 1e-2])
    assert out["jitter"][0] == fo.jitter > 0.0 and out["jitter"][1] == 0.0
    assert relerr(out["lml"].cpu().numpy()[0, 0], g["jit_lml"]) < 1e-6


@pytest.mark.parametrize("n,G", [(4000, 16), (8000, 2)])
def test_refit_sweep_config3_n4000_matches_scipy(eng, n, G):
    """BASELINE configs[2] at its measured sizes: n = 4000 (all 16 candidates) and n = 8000 (two of them), D = 11, P = 3 outputs, the grid of SURVEY 8(d)
    (shared lengthscale log-uniform in [0.1, 10] x noise in {1e-4 .. 1e-1}) -- the batched device sweep against
    `scipy.linalg.cholesky` LMLs of exact_gp.py:118-204 (same jitter escalation), candidate by candidate.  Tolerance 1e-7
    relative: cond(K + noise I) reaches 4e7 at lengthscale 10 / noise 1e-4 (two FP64 factorizations differ by cond x eps)."""
    import scipy.linalg as sla
    from gp_mpc_rocket_landing_b200.gp import refit
    from oracle import gp as ogp
    D, P = 11, 3
    rng = np.random.default_rng(1)
    X = rng.normal(0, 0.4, (n, D))
    Y = np.stack([np.sin(2 * X[:, 0]) + 0.05 * rng.normal(size=n), X[:, 1] * X[:, 2], np.cos(X[:, 3])], axis=1)
    ls, sv, nz = refit.default_grid(D)
    assert ls.shape[0] == 16
    if G < 16:      # n = 8000 (the largest size of BASELINE configs[2]): two candidates keep the CPU side to a few seconds
        pick = [5, 10]
        ls, sv, nz = ls[pick], sv[pick], nz[pick]
    r = refit.refit_sweep(X, Y, ls, sv, nz, kind=0)
    Yn = (Y - Y.mean(0)) / Y.std(0)
    want = np.zeros((G, P)); jit = np.zeros(G)
    for g in range(G):
        Kn = ogp.gram(ogp.SE, X, None, ls[g], sv[g]) + nz[g] * np.eye(n)
        j = 0.0
        while True:
            try:
                L = sla.cholesky(Kn + j * np.eye(n) if j else Kn, lower=True, check_finite=False)
                break
            except sla.LinAlgError:
                j = 1e-6 if j == 0.0 else j * 10
                assert j < 1.0
        jit[g] = j
        A = sla.cho_solve((L, True), Yn, check_finite=False)
        want[g] = -0.5 * np.einsum("np,np->p", Yn, A) - np.sum(np.log(np.diag(L))) - 0.5 * n * np.log(2 * np.pi)
    np.testing.assert_array_equal(r["jitter"], jit)
    np.testing.assert_allclose(r["lml"], want, rtol=1e-7, atol=0)
    assert r["best_shared"] == int(np.argmax(want.sum(1)))
    np.testing.assert_array_equal(r["best"], np.argmax(want, axis=0))


@pytest.mark.parametrize("n,D,noise,tol", [(300, 5, 1e-2, 1e-9), (1000, 11, 1e-2, 1e-8), (130, 17, 1e-3, 1e-8)])
def test_lml_gradient_matches_closed_form(eng, n, D, noise, tol):
    """gpmpc_lml_grad_se (all D + 2 components, P outputs on one factor) against the oracle's closed form, which
    tests/test_oracle_golden.py pins to central differences of the reference LML restatement.  tol: FP64, relative to the
    largest component; K^{-1} enters, so it scales with cond(K) like the variance checks above."""
    from oracle import gp as og
    rng = np.random.default_rng(n + D)
    X = rng.normal(size=(n, D))
    Y = np.stack([np.sin(X[:, 0]) + 0.3 * X[:, 1], np.cos(X[:, 2]) * X[:, 0], 0.1 * rng.normal(size=n)], axis=1)
    ls = np.exp(rng.normal(scale=0.3, size=D)) * np.sqrt(D)
    sv = 1.7
    f = eng.ExactFactor(0, ls, sv, noise).fit(X)
    m = eng.ExactModelDev(f).fit_targets(Y)
    assert f.jitter == 0.0
    g = m.lml_gradient().cpu().numpy()
    assert g.shape == (3, D + 2)
    for p in range(3):
        ref = og.exact_lml_gradient(og.exact_fit(0, X, Y[:, p], ls, sv, noise))
        assert np.abs(g[p] - ref).max() / np.abs(ref).max() < tol, (p, g[p], ref)


def test_optimize_hyperparameters_with_device_gradients(eng):
    """ExactGP.optimize_hyperparameters: L-BFGS-B fed the device gradient reaches the optimum the reference's
    finite-difference path reaches (exact_gp.py:357-421), with far fewer refits."""
    from gp_mpc_rocket_landing_b200.gp import ExactGP, SquaredExponentialARD
    rng = np.random.default_rng(5)
    X = rng.uniform(-2, 2, size=(200, 3))
    y = np.sin(2 * X[:, 0]) + 0.5 * X[:, 1] + 0.05 * rng.normal(size=200)
    res = {}
    for mode in (True, False):
        gp = ExactGP(SquaredExponentialARD(3), noise_variance=1e-2).fit(X, y)
        lml0 = gp.log_marginal_likelihood
        g = gp.log_marginal_likelihood_gradient()
        assert g.shape == (5,)
        out = gp.optimize_hyperparameters(n_restarts=1, analytic_gradients=mode)
        assert out["log_marginal_likelihood"] > lml0
        res[mode] = (out["log_marginal_likelihood"], gp.kernel.get_params().copy(), gp.noise_variance)
    assert abs(res[True][0] - res[False][0]) < 1e-3 * abs(res[False][0])
    # the finite-difference run stops on its noisier gradient a little earlier: same optimum to a few per cent
    np.testing.assert_allclose(res[True][1][:3], res[False][1][:3], rtol=0.1)
    assert res[True][0] >= res[False][0] - 1e-6 * abs(res[False][0])


@pytest.mark.parametrize("n0,adds", [(100, (1, 27, 1)), (250, (6, 130)), (126, (2, 3))])
def test_incremental_factor_append_equals_refit(eng, n0, adds):
    """ExactGP.update: the factor extended row-wise (ExactFactor.append) against a from-scratch fit on the concatenated
    data and against the oracle (exact_gp.py:118-204).  Crossing a 64-block, a 128-pad boundary (buffers grow) and a chunk
    of more than 128 points are covered.  The old rows of L are bit-identical; the new rows, alpha, LML and predictions
    agree to FP64 rounding scaled by cond(K) (1e-9 here, noise 1e-2)."""
    from oracle import gp as og
    from gp_mpc_rocket_landing_b200.gp import ExactGP, SquaredExponentialARD
    rng = np.random.default_rng(n0)
    D = 6
    ntot = n0 + sum(adds)
    X = rng.normal(size=(ntot, D))
    y = np.sin(X[:, 0]) + 0.2 * X[:, 1] ** 2 + 0.05 * rng.normal(size=ntot)
    Xq = rng.normal(size=(40, D))

    def kern():
        k = SquaredExponentialARD(D)
        k.set_params(np.log(np.concatenate([[1.4], np.linspace(1.5, 2.5, D)])))
        return k
    inc = ExactGP(kern(), noise_variance=1e-2).fit(X[:n0], y[:n0])
    L0 = inc._L.copy()
    n = n0
    for m in adds:
        inc.update(X[n:n + m], y[n:n + m])
        n += m
        assert inc.n_train == n
    full = ExactGP(kern(), noise_variance=1e-2).fit(X, y)
    Li, Lf = inc._L, full._L
    assert np.array_equal(Li[:n0, :n0], L0)            # existing rows untouched
    assert np.array_equal(np.triu(Li, 1), np.zeros_like(Li))
    np.testing.assert_allclose(Li, Lf, rtol=0, atol=1e-11 * np.abs(Lf).max())
    ref = og.exact_fit(0, X, y, inc.kernel.lengthscales, inc.kernel.signal_variance, 1e-2)
    assert relerr(inc._alpha, ref.alpha) < 1e-9
    assert abs(inc.log_marginal_likelihood - ref.lml) < 1e-9 * abs(ref.lml)
    pi, pf = inc.predict(Xq), full.predict(Xq)
    mo, vo = og.exact_predict(ref, Xq)[:2]
    assert relerr(pi.mean, mo) < 1e-9 and relerr(pi.variance, vo) < 1e-8
    assert relerr(pi.mean, pf.mean) < 1e-10 and relerr(pi.variance, pf.variance) < 1e-9
    assert inc.update(np.zeros((0, D)), np.zeros(0)) is inc and inc.n_train == ntot      # empty update: no-op
    # changed hyper-parameters: update falls back to a full refit with the new ones
    inc.kernel.set_params(inc.kernel.get_params() + 0.1)
    inc.update(X[:3] + 0.5, y[:3])
    assert inc.n_train == ntot + 3 and np.allclose(inc._model.f.ls_host, inc.kernel.lengthscales)


def test_blocked_potrs_large_n_few_rhs(eng):
    """gpmpc_potrs switches to the blocked multi-CTA substitution for n_pad >= 2048 with few right-hand sides (one launch
    per 64-block, all SMs on the update): batch of two factors x three right-hand sides against numpy's cho_solve, and the
    exact-GP alpha / LML at n = 2100 against the oracle (1e-8: cond(K) ~ n s2 / noise = 2e5; the backward error is 1e-13)."""
    from oracle import gp as og
    torch = eng.torch_()
    rng = np.random.default_rng(2)
    n, n_pad = 2100, 2176
    Ls, Ws = [], []
    A = torch.zeros((2, n_pad, n_pad), dtype=torch.float64, device="cuda")
    Kh = []
    for b in range(2):
        X = rng.normal(size=(n, 5))
        K = og.gram(0, X, None, np.full(5, 2.0 + b), 1.0) + 1e-2 * np.eye(n)
        Kh.append(K)
        A[b] = torch.eye(n_pad, dtype=torch.float64, device="cuda")
        A[b, :n, :n] = eng.to_dev(K)
    Winv, info = eng.potrf(A, batch=2)
    assert int(info.abs().sum().item()) == 0
    Bh = rng.normal(size=(2, 3, n))
    B = torch.zeros((2, 3, n_pad), dtype=torch.float64, device="cuda")
    B[:, :, :n] = eng.to_dev(Bh)
    eng.call("gpmpc_potrs", eng.ptr(A), n_pad, n_pad, n_pad * n_pad, eng.ptr(Winv), eng.ptr(B), 3, 2, eng.stream())
    Xs = B.cpu().numpy()
    for b in range(2):
        # backward error (the forward error scales with cond(K) ~ 1e5 for either solver): |K x - b| <= 1e-13 |K| |x|
        res = np.abs(Kh[b] @ Xs[b, :, :n].T - Bh[b].T)
        assert np.max(res / (np.abs(Kh[b]) @ np.abs(Xs[b, :, :n].T))) < 1e-13
        assert relerr(Xs[b, :, :n], np.linalg.solve(Kh[b], Bh[b].T).T) < 1e-7
        assert np.all(Xs[b, :, n:] == 0.0)
    X = rng.normal(size=(n, 5))
    Y = np.stack([np.sin(X[:, 0]), X[:, 1] * X[:, 2]], 1)
    f = eng.ExactFactor(0, np.full(5, 2.0), 1.0, 1e-2).fit(X)
    m = eng.ExactModelDev(f).fit_targets(Y)
    for p in range(2):
        ref = og.exact_fit(0, X, Y[:, p], np.full(5, 2.0), 1.0, 1e-2)
        assert relerr(m.alpha[p, :n].cpu().numpy(), ref.alpha) < 1e-8
        assert abs(m.lml[p].item() - ref.lml) < 1e-10 * abs(ref.lml)


def test_sweep_then_gradient_refinement_shared_hyperparameters(eng):
    """refit.sweep_and_refine: the best grid candidate of the batched sweep refined by L-BFGS-B on the summed LML of the
    three outputs with the device gradient (one fit + one gradient pass per evaluation).  The refined LML is at least the
    grid's, the gradient at the optimum vanishes (interior components), and the oracle agrees on the LML there."""
    from oracle import gp as og
    from gp_mpc_rocket_landing_b200.gp import refit
    rng = np.random.default_rng(8)
    n, D = 400, 4
    X = rng.uniform(-2, 2, size=(n, D))
    Y = np.stack([np.sin(2 * X[:, 0]) + 0.3 * X[:, 1], np.cos(X[:, 1]) * X[:, 2], 0.5 * X[:, 3] ** 2], 1)
    Y += 0.05 * rng.normal(size=Y.shape)
    r = refit.sweep_and_refine(X, Y)
    assert r["lml"].shape == (3,) and r["lml"].sum() >= r["grid_best_lml"].sum() - 1e-9
    assert r["lml"].sum() > r["grid_best_lml"].sum() + 1.0          # the grid is coarse: the refinement moves
    f = eng.ExactFactor(0, r["lengthscales"], r["signal_variance"], r["noise_variance"]).fit(X)
    m = eng.ExactModelDev(f).fit_targets(Y)
    g = m.lml_gradient().sum(dim=0).cpu().numpy()
    assert np.max(np.abs(g)) < 1e-2 * max(1.0, abs(r["lml"].sum()))
    for p in range(3):
        ref = og.exact_fit(0, X, Y[:, p], r["lengthscales"], r["signal_variance"], r["noise_variance"])
        assert abs(ref.lml - r["lml"][p]) < 1e-8 * abs(ref.lml)
    assert r["n_evaluations"] < 120      # the differenced optimiser needs (D + 3) fits per gradient

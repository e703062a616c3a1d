"""GPU parity: sparse (FITC / VFE) GP, FastGPPredictor family and StructuredRocketGP through the public API against the
golden vectors generated from the real reference (oracle/gen_golden.py) and the oracle.  Inducing points are injected
(the reference draws them with k-means from the global numpy RNG).  Tolerances: the sparse posterior goes
through K_uu + 1e-6 I (cond ~ 1e6), so factors agree to ~1e-9 and predictions to 1e-8 relative."""
import numpy as np
import pytest

from conftest import load_golden, relerr

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def G():
    import gp_mpc_rocket_landing_b200.gp as gp
    gp.engine.torch_()
    return gp


@pytest.mark.parametrize("method", ["fitc", "vfe"])
def test_sparse_gp_matches_reference(G, method):
    g = load_golden("sparse_gp.npz")
    k = G.SquaredExponentialARD(6, float(g["sv"]), g["ls"])
    gp = G.SparseGP(k, g["Z"].shape[0], float(g["noise"]), method, g["Z"].copy()).fit(g["X"], g["y"])
    np.testing.assert_allclose(gp._L_uu, g[f"{method}_Luu"], rtol=1e-8, atol=1e-12)
    np.testing.assert_allclose(gp._L_B, g[f"{method}_LB"], rtol=1e-7, atol=1e-10)
    assert relerr(gp._alpha, g[f"{method}_alpha"]) < 1e-7
    assert relerr(gp.log_marginal_likelihood, g[f"{method}_lml"]) < 1e-9
    if method == "fitc":
        np.testing.assert_allclose(gp._Lambda_diag, g["fitc_Lambda"], rtol=1e-7, atol=1e-12)
    pr = gp.predict(g["Xt"])
    assert relerr(pr.mean, g[f"{method}_mean"]) < 1e-8
    assert relerr(pr.variance, g[f"{method}_var"]) < 1e-7
    np.testing.assert_allclose(pr.std, np.sqrt(pr.variance), rtol=1e-15)
    m2, v2 = gp.predict_f(g["Xt"][:3])
    np.testing.assert_allclose(m2, pr.mean[:3], rtol=1e-13)
    # single query == batch row; before-fit error; update() refits on the stacked data with Z kept
    one = gp.predict(g["Xt"][5])
    assert one.mean.shape == (1,) and abs(one.mean[0] - pr.mean[5]) <= 1e-12 * max(1.0, abs(pr.mean[5]))
    with pytest.raises(RuntimeError):
        G.SparseGP(k, 10).predict(g["Xt"])
    n0 = gp.n_train
    gp.update(g["X"][:7] + 0.01, g["y"][:7])
    assert gp.n_train == n0 + 7 and gp.inducing_points.shape == g["Z"].shape


def test_sparse_matern_against_oracle(G):
    from oracle import gp as ogp
    g = load_golden("sparse_gp.npz")
    for kind, cls in ((1, G.Matern32), (2, G.Matern52)):
        f = ogp.sparse_fit(kind, g["X"], g["y"], g["Z"], g["ls"], 0.8, 1e-3, "fitc")
        mo, vo, _ = ogp.sparse_predict(f, g["Xt"])
        gp = G.SparseGP(cls(6, 0.8, g["ls"]), 40, 1e-3, "fitc", g["Z"].copy()).fit(g["X"], g["y"])
        pr = gp.predict(g["Xt"])
        assert relerr(pr.mean, mo) < 1e-8 and relerr(pr.variance, vo) < 1e-7
        assert relerr(gp.log_marginal_likelihood, f.lml) < 1e-9


def test_simple3dof_sparse_shared_factor(G):
    """Simple3DoFGP(use_sparse=True): 3 outputs on one (L_uu, L_B); Z injected after the reference's own re-draw."""
    g = load_golden("sparse_gp.npz")
    gp = G.Simple3DoFGP(n_inducing=50, use_sparse=True)
    gp.add_data(g["s3_X"], g["s3_U"], g["s3_D"])
    multi = gp.gp
    multi.share_inducing = False                  # keep the injected Z (sparse_gp.py:448-451 would re-draw it)
    for s in multi.gps:
        s._Z = g["s3_Z"].copy()
    gp.fit()
    assert multi._shared is not None              # one factorisation for the three outputs
    m, v = gp.predict_batch(g["s3_Xq"], g["s3_Uq"])
    assert relerr(m, g["s3_mean"]) < 1e-8 and relerr(v, g["s3_var"]) < 1e-7
    m1, v1 = gp.predict(g["s3_Xq"][3], g["s3_Uq"][3])
    np.testing.assert_allclose(m1, m[3], rtol=1e-12, atol=1e-15)
    np.testing.assert_allclose(v1, v[3], rtol=1e-12, atol=1e-18)
    # default construction draws Z with k-means on the host, as the reference does: runs and gives finite output
    np.random.seed(0)
    gp2 = G.Simple3DoFGP()                        # use_sparse=True, n_inducing=50 (structured_gp.py:423-428)
    gp2.add_data(g["s3_X"], g["s3_U"], g["s3_D"])
    mm, vv = gp2.predict(g["s3_Xq"][0], g["s3_Uq"][0])
    assert mm.shape == (3,) and np.all(np.isfinite(mm)) and np.all(vv > 0)


def test_fast_gp_predictor(G):
    g = load_golden("fast_gp.npz")
    fg = G.FastGPPredictor()
    ms = fg.fit(g["X"], g["Y"], g["ls"], float(g["variance"]), float(g["noise"]))
    assert ms > 0
    mean, var = fg.predict(g["Xt"])
    np.testing.assert_allclose(mean, g["mean"], rtol=1e-9, atol=1e-11)
    np.testing.assert_allclose(var, g["var"], rtol=1e-8, atol=1e-12)
    np.testing.assert_allclose(fg._alpha, g["alpha"], rtol=1e-8, atol=1e-10)
    np.testing.assert_allclose(np.diag(fg._L), g["Ldiag"], rtol=1e-10)
    m1, v1 = fg.predict_single(g["Xt"][4])
    np.testing.assert_allclose(m1, mean[4], rtol=1e-12, atol=1e-14)
    assert fg.predict(g["Xt"], return_var=False)[1] is None
    cg = G.create_fast_gp("cached")
    cg.fit(g["X"], g["Y"], g["ls"], float(g["variance"]), float(g["noise"]))
    a = cg.predict_single(g["Xt"][0]); b = cg.predict_single(g["Xt"][0])
    assert cg.get_cache_stats()["hits"] == 1 and np.array_equal(a[0], b[0])
    fg.update(g["Xt"][:3], g["mean"][:3])
    assert fg._n_train == 203


def test_sparse_gp_predictor_against_numpy(G):
    """fast_gp.py:338-462 restated in numpy here (explicit formulas), same `np.random.choice` draw."""
    from scipy.linalg import solve_triangular
    g = load_golden("fast_gp.npz")
    X, Y, Xt, ls = g["X"], g["Y"], g["Xt"], g["ls"]
    var0, noise, jit, m = 1.2, 0.01, 1e-6, 40

    def k(A, B):
        As, Bs = A / ls, B / ls
        d = (As**2).sum(1, keepdims=True) + (Bs**2).sum(1, keepdims=True).T - 2 * As @ Bs.T
        return var0 * np.exp(-0.5 * np.maximum(d, 0))

    np.random.seed(3)
    idx = np.random.choice(len(X), m, replace=False)
    Z = X[idx]
    Kmm = k(Z, Z) + jit * np.eye(m)
    Knm = k(X, Z)
    Lam = var0 - np.sum(Knm @ np.linalg.inv(Kmm) * Knm, axis=1) + noise
    Lm = np.linalg.cholesky(Kmm)
    A = np.linalg.cholesky(Kmm + Knm.T @ np.diag(1 / Lam) @ Knm + jit * np.eye(m))
    W = np.linalg.solve(A.T, np.linalg.solve(A, Knm.T @ np.diag(1 / Lam) @ Y))
    Ksm = k(Xt, Z)
    mean_ref = Ksm @ W
    v = solve_triangular(Lm, Ksm.T, lower=True); w = solve_triangular(A, Ksm.T, lower=True)
    var_ref = np.maximum(var0 - (v**2).sum(0) + (w**2).sum(0), 0)

    np.random.seed(3)
    sp = G.create_fast_gp("sparse", n_inducing=m)
    sp.fit(X, Y, ls, var0, noise)
    np.testing.assert_array_equal(sp._Z, Z)
    mean, var = sp.predict(Xt)
    assert relerr(mean, mean_ref) < 1e-7
    np.testing.assert_allclose(var[:, 0], var_ref, rtol=1e-6, atol=1e-9)
    assert np.array_equal(var[:, 0], var[:, 5])


def test_structured_rocket_gp_exact_and_sparse(G):
    g = load_golden("structured_gp.npz")
    sg = G.StructuredRocketGP(G.StructuredGPConfig(use_sparse=False, noise_variance=1e-3))
    z, _, pv, _ = sg.predict(g["X"][0], g["U"][0])               # no data: prior (structured_gp.py:250-253)
    assert np.all(z == 0) and np.all(pv == 0.1)
    sg.add_data(g["X"], g["U"], g["Dv"], g["Dw"])
    for i in range(12):
        dv, dw, vv, vw = sg.predict(g["X"][i] + 0.01, g["U"][i] - 0.01)
        assert relerr(dv, g["dv_mean"][i]) < 1e-8 and relerr(dw, g["dw_mean"][i]) < 1e-8
        assert relerr(vv, g["dv_var"][i]) < 1e-7 and relerr(vw, g["dw_var"][i]) < 1e-7
    dvb, dwb, vvb, vwb = sg.predict_batch(g["X"][:12] + 0.01, g["U"][:12] - 0.01)
    assert relerr(dvb, g["dv_mean"]) < 1e-8 and relerr(vwb, g["dw_var"]) < 1e-7
    d_mean, d_var = sg.get_full_residual(g["X"][0] + 0.01, g["U"][0] - 0.01)
    assert d_mean.shape == (14,) and np.all(d_mean[:4] == 0) and relerr(d_mean[4:7], g["dv_mean"][0]) < 1e-8
    # sparse default (BASELINE config 4 model: 100 inducing points): runs end to end, finite, positive variances
    np.random.seed(1)
    ss = G.StructuredRocketGP(G.StructuredGPConfig(n_inducing=100, use_sparse=True))
    ss.add_data(g["X"], g["U"], g["Dv"], g["Dw"])
    out = ss.predict_batch(g["X"][:20], g["U"][:20])
    assert all(np.all(np.isfinite(o)) for o in out) and np.all(out[2] > 0) and np.all(out[3] > 0)

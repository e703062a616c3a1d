"""Online data path (reference src/gp/online_update.py) against tests/golden/online_update.npz, which
oracle/gen_golden.py wrote by driving the reference's own DataBuffer / OnlineGPUpdater / ResidualCollector."""
import numpy as np
import pytest

from conftest import load_golden, relerr


def test_data_buffer_ring_matches_reference_deques():
    from gp_mpc_rocket_landing_b200.gp.online_update import DataBuffer
    g = load_golden("online_update.npz")
    buf = DataBuffer(max_size=12, feature_dim=3, target_dim=2)
    assert buf.is_empty and buf.get_features().shape == (0, 3) and buf.get_targets().shape == (0, 2)
    acc = [buf.add_if_novel(g["buf_F"][i], g["buf_T"][i], float(g["buf_nov"][i]), threshold=0.3, min_distance=0.05)
           for i in range(40)]
    assert np.array_equal(acc, g["buf_accepted"])
    assert np.array_equal(buf.get_features(), g["buf_Z"]) and np.array_equal(buf.get_targets(), g["buf_D"])
    rZ, rD = buf.get_recent(5)
    assert np.array_equal(rZ, g["buf_recent_Z"]) and np.array_equal(rD, g["buf_recent_D"])
    st = buf.get_statistics()
    np.testing.assert_allclose([st["size"], st["total_added"], st["total_rejected"], st["acceptance_rate"],
                                st["mean_novelty"]], g["buf_stats"], rtol=1e-15)
    assert buf.is_full and buf.evictions == st["total_added"] - 12
    buf.clear()
    assert buf.is_empty and buf.size == 0


def test_residual_collector_matches_reference():
    from gp_mpc_rocket_landing_b200.gp.online_update import ResidualCollector
    g = load_golden("online_update.npz")
    nominal = lambda x, u, dt: x + dt * np.concatenate([[0.0], x[4:7], u, 0.5 * x[7:14]])  # noqa: E731
    rc = ResidualCollector(nominal, max_samples=25)
    assert rc.get_training_data()[0].shape == (0, 14) and rc.get_statistics() == {"n_samples": 0}
    for i in range(30):
        rc.record(g["rc_X"][i], g["rc_U"][i], g["rc_Xn"][i], 0.1)
    for got, want in zip(rc.get_training_data(), (g["rc_states"], g["rc_controls"], g["rc_dv"], g["rc_dw"])):
        np.testing.assert_allclose(got, want, rtol=1e-15, atol=0)
    # the batch form (one call of a batched nominal model) stores the same rows
    rb = ResidualCollector(nominal, max_samples=25)
    rb.record_batch(g["rc_X"], g["rc_U"], g["rc_Xn"], 0.1,
                    batched_dynamics=lambda X, U, dt: np.stack([nominal(X[i], U[i], dt) for i in range(len(X))]))
    for got, want in zip(rb.get_training_data(), rc.get_training_data()):
        np.testing.assert_allclose(got, want, rtol=1e-15, atol=0)
    assert rb.n_samples == 25 and set(rb.get_statistics()) == {"n_samples", "d_v_mean", "d_v_std", "d_omega_mean", "d_omega_std"}


class _FakeGP:
    """Scheduling logic only: constant novelty, records how it was trained."""

    class kernel:
        signal_variance = 1.0

    supports_incremental_update = True     # the updater extends only ExactGP or a model that declares this

    def __init__(self):
        self._model = None
        self.n_train = 0
        self.calls = []

    def fit(self, Z, y):
        self._model = object(); self.n_train = len(y); self.calls.append(("fit", len(y))); return self

    def update(self, Z, y):
        self.n_train += len(y); self.calls.append(("update", len(y))); return self

    def predict(self, Z):
        class P: variance = np.full(len(Z), 0.5)
        return P


def test_updater_schedule_and_incremental_routing():
    from gp_mpc_rocket_landing_b200.gp.online_update import OnlineGPUpdater, OnlineUpdateConfig
    cfg = OnlineUpdateConfig(buffer_size=50, min_data_for_fit=8, update_interval=4, refit_interval=12, min_distance=None)
    gp = _FakeGP()
    up = OnlineGPUpdater(gp, cfg)
    assert up.update() == {"status": "no_data"}
    rng = np.random.default_rng(0)
    modes = []
    for i in range(60):
        assert up.add_observation(rng.normal(size=3), rng.normal(size=2), rng.normal())
        if up.should_update():
            modes.append(up.update()["mode"])
    assert up.update() == {"status": "skipped"}
    # first update fits; then the factor is extended by the 4 new rows until 12 new points force a refit; once the
    # 50-row ring starts evicting, the training prefix has moved and every update is a refit
    assert modes[0] == "refit" and modes[1] == "extended" and "refit" in modes[2:4]
    assert ("update", 4) in gp.calls and all(m == "refit" for m in modes[-2:])
    st = up.get_statistics()
    assert st["size"] == 50 and st["total_added"] == 60 and st["incremental_updates"] == modes.count("extended")
    # the batch gate accepts exactly what the row-by-row gate accepts
    a, b = OnlineGPUpdater(_FakeGP(), cfg), OnlineGPUpdater(_FakeGP(), cfg)
    X, U, D = rng.normal(size=(20, 3)), rng.normal(size=(20, 2)), rng.normal(size=20)
    X[5] = X[4]; U[5] = U[4]
    a.config = b.config = OnlineUpdateConfig(buffer_size=50, min_distance=0.01)
    one = np.array([a.add_observation(X[i], U[i], D[i]) for i in range(20)])
    assert np.array_equal(b.add_observations(X, U, D), one) and not one[5] and one.sum() == 19
    assert b.add_observations(np.zeros((0, 3)), np.zeros((0, 2)), np.zeros(0)).shape == (0,)   # empty batch: no-op


@pytest.mark.gpu
def test_online_updater_sequence_matches_reference():
    """Same stream of observations as the reference run: accept / reject flags, update schedule, final buffer and the
    refitted sparse GP's predictions (1e-7: FITC with K_uu + 1e-6 I, cond ~ 1e6)."""
    from gp_mpc_rocket_landing_b200.gp import (ExactGP, OnlineGPUpdater, OnlineUpdateConfig, SparseGP,
                                                SquaredExponentialARD)
    g = load_golden("online_update.npz")
    X, U, d = g["X"], g["U"], g["d"]
    gp = SparseGP(SquaredExponentialARD(6, 1.0, g["ls"]), n_inducing=20, noise_variance=1e-2,
                  inducing_points=g["Zind"].copy())
    cfg = OnlineUpdateConfig(buffer_size=60, min_data_for_fit=20, update_interval=10, novelty_threshold=0.05,
                             min_distance=0.01)
    up = OnlineGPUpdater(gp, cfg)
    accepted, updates = [], []
    for i in range(len(d)):
        accepted.append(up.add_observation(X[i], U[i], d[i]))
        if up.should_update():
            r = up.update()
            updates.append([i, r["n_data"], r["total_updates"], float(r["status"] == "success")])
    assert np.array_equal(accepted, g["accepted"])
    assert np.array_equal(np.array(updates), g["updates"])
    Zb, Db = up._buffer.get_data()
    assert np.array_equal(Zb, g["final_Z"]) and np.array_equal(Db, g["final_D"])
    p = gp.predict(g["Xq"])
    assert relerr(p.mean, g["pred_mean"]) < 1e-7 and relerr(p.variance, g["pred_var"]) < 1e-7
    st = up.get_statistics()
    assert [st["total_updates"], st["points_since_update"], st["size"], st["total_added"], st["total_rejected"]] == \
        list(g["stats"])
    # exact GP behind the same updater: updates extend the factor; the model equals a fresh fit on the buffer
    ex = ExactGP(SquaredExponentialARD(6, 1.0, g["ls"]), noise_variance=1e-2)
    ue = OnlineGPUpdater(ex, OnlineUpdateConfig(buffer_size=500, min_data_for_fit=20, update_interval=10,
                                                refit_interval=10**6, novelty_threshold=0.05, min_distance=0.01))
    flags = ue.add_observations(X[:40], U[:40], d[:40])
    assert flags.all()                                     # nothing fitted yet: no gating
    modes = [ue.update()["mode"]]
    for s0 in range(40, 160, 24):
        ue.add_observations(X[s0:s0 + 24], U[s0:s0 + 24], d[s0:s0 + 24])   # one batched variance query per call
        if ue.should_update():
            modes.append(ue.update()["mode"])
    modes.append(ue.update(force=True)["mode"])            # flush the rows accepted since the last scheduled update
    assert modes[0] == "refit" and set(modes[1:]) <= {"extended", "refit"} and "extended" in modes
    assert ue.incremental_updates == modes.count("extended") and modes.count("refit") == 1
    Ze, De = ue._buffer.get_data()
    fresh = ExactGP(SquaredExponentialARD(6, 1.0, g["ls"]), noise_variance=1e-2).fit(Ze, De[:, 0])
    assert ex.n_train == len(Ze)
    pe, pf = ex.predict(g["Xq"]), fresh.predict(g["Xq"])
    assert relerr(pe.mean, pf.mean) < 1e-10 and relerr(pe.variance, pf.variance) < 1e-9
    assert abs(ex.log_marginal_likelihood - fresh.log_marginal_likelihood) < 1e-9 * abs(fresh.log_marginal_likelihood)


def test_constraint_tightening_and_tube_host_classes():
    """uncertainty_prop.py:322-470 (host arithmetic): back-offs kappa sqrt(g' Sigma g) and the |A| tube recursion against
    direct evaluation; the batch form equals the per-stage loop."""
    from scipy.stats import norm
    from gp_mpc_rocket_landing_b200.mpc.uncertainty_prop import (ConstraintTightening, PropagatedUncertainty,
                                                                 TubeBasedRobustness)
    rng = np.random.default_rng(1)
    N, n = 6, 7
    M = rng.normal(size=(N + 1, n, n))
    covs = M @ M.transpose(0, 2, 1) * 1e-3
    pu = PropagatedUncertainty(means=rng.normal(size=(N + 1, n)), covariances=covs)
    ct = ConstraintTightening(0.95)
    assert abs(ct.kappa - norm.ppf(0.95)) < 1e-15
    a = rng.normal(size=n)
    assert abs(ct.tighten_linear_constraint(a, 0.3, pu.means[2], covs[2]) - (0.3 + ct.kappa * np.sqrt(a @ covs[2] @ a))) < 1e-15
    G = rng.normal(size=(3, n))
    bo = ct.compute_back_offs(pu, [list(G)] * N)
    assert bo.shape == (N, 3)
    np.testing.assert_allclose(bo[4, 1], ct.kappa * np.sqrt(G[1] @ covs[4] @ G[1]), rtol=1e-14)
    np.testing.assert_allclose(ct.compute_back_offs_batch(covs[None], G)[0], bo, rtol=1e-12)
    lo, hi = pu.get_confidence_bounds(3)
    np.testing.assert_allclose(hi - lo, 2 * norm.ppf(0.975) * np.sqrt(np.diag(covs[3])), rtol=1e-14)

    class Dyn:
        def linearize(self, x, u, dt=None):
            A = np.eye(n); A[1:4, 4:7] = dt * np.eye(3); A[4:7, 0] = -u * dt
            return A, np.zeros((n, 3))
    X, U = rng.normal(size=(N + 1, n)), rng.normal(size=(N, 3))
    W = TubeBasedRobustness(Dyn(), d_max=0.2).compute_tube(X, U, 0.1)
    w = np.zeros(n)
    for k in range(N):
        w = np.abs(Dyn().linearize(X[k], U[k], dt=0.1)[0]) @ w
        w[4:7] += 0.02
        np.testing.assert_allclose(W[k + 1], w, rtol=1e-15)
    assert np.all(W[0] == 0)

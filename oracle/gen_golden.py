#!/usr/bin/env python
"""Generate tests/golden/*.npz from the REAL reference (run in the build container only).

Imports /root/reference/src/{gp,mpc,experiments} through a ``sys.modules`` shim that bypasses
``src/__init__.py`` (which needs the absent third-party ``simdyn``) and ``src/mpc/__init__.py``
(which needs CasADi), and drives the reference's own code on seeded inputs.  The QP assembler
(osqp_rti.py) is run with a stub ``osqp`` module that only records what ``setup`` receives.
Outputs are small fixtures; the committed files are what the tests (and the GPU box) read --
/root/reference itself does not travel.

    python -m oracle.gen_golden
"""
from __future__ import annotations

import os
import sys
import types

os.environ["PYTHONDONTWRITEBYTECODE"] = "1"
sys.dont_write_bytecode = True

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))   # `from oracle import ...` below

REF = "/root/reference/src"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")


def install_shim():
    for name in ("src", "src.mpc", "src.experiments"):
        pkg = types.ModuleType(name)
        pkg.__path__ = [os.path.join(REF, *name.split(".")[1:])]
        sys.modules[name] = pkg

    class _StubOSQP:  # records the arguments of the reference's call sites (osqp_rti.py:464-478)
        last = None

        def setup(self, **kw):
            _StubOSQP.last = kw

        def update(self, **kw):
            pass

        def warm_start(self, **kw):
            pass

    osqp = types.ModuleType("osqp")
    osqp.OSQP = _StubOSQP
    sys.modules["osqp"] = osqp
    return _StubOSQP


def gp_fixture(n=1000, seed=0):
    """SURVEY section 8(d) config 1 synthetic residual data."""
    rng = np.random.default_rng(seed)
    m = rng.uniform(1.2, 2, n); alt = rng.uniform(0, 40, n)
    y = rng.normal(0, 3, n); z = rng.normal(0, 3, n)
    vx = rng.normal(-3, 1, n); vy = rng.normal(0, .5, n); vz = rng.normal(0, .5, n)
    X = np.stack([m, alt, y, z, vx, vy, vz], 1)
    U = np.stack([rng.uniform(.3, 5, n), rng.uniform(-1, 1, n), rng.uniform(-1, 1, n)], 1)
    rho = 1.225 * np.exp(-alt / 8500)
    v = X[:, 4:7]; sp = np.linalg.norm(v, axis=1)
    D = -0.5 * rho[:, None] * 0.5 * 0.5 * sp[:, None] * v / m[:, None] + rng.normal(0, 0.01, (n, 3))
    return X, U, D


def query_fixture(X, U, nq, seed):
    rng = np.random.default_rng(seed)
    idx = rng.integers(0, X.shape[0], nq)
    return X[idx] + rng.normal(0, 0.05, (nq, 7)), U[idx] + rng.normal(0, 0.05, (nq, 3))


def mc_loop_fixture(save):
    """12. the reference's own closed loop (monte_carlo.py:401-583 `run_single`) around deterministic stub controllers:
    pins the loop semantics -- termination rules, per-step target for `.solve` controllers, `success == False ->
    DIVERGENCE`, `.step` controllers whose flag is ignored, the for/else TIMEOUT -- that `gpmpc_mc_advance_3dof` and
    `oracle.mpc.run_single` restate.  Runs 32 seeds with max_time 20 s (200 steps) so that every outcome class occurs."""
    import src.experiments.monte_carlo as MC
    from oracle import mpc as ompc
    from oracle.mpc import EulerDynamics

    class Sol:
        def __init__(self, u, ok):
            self.u0, self.success = u, ok

    class SolveCtl:      # has .solve, no .step: the branch of monte_carlo.py:495-510
        def __init__(self, gains, fail_step):
            self.gains, self.fail_step, self.k, self.targets = gains, fail_step, 0, []

        def solve(self, x, x_target):
            self.targets.append(x_target.copy())
            ok = self.k != self.fail_step
            self.k += 1
            return Sol(ompc.pd_test_law(x, x_target, self.gains), ok)

    class StepCtl:       # has .initialize / .step: the branch of :492-494 (fixed target, flag ignored)
        def __init__(self, gains, fail_step):
            self.gains, self.fail_step, self.k = gains, fail_step, 0

        def initialize(self, x0, x_target):
            self.xt = x_target.copy()

        def step(self, x):
            ok = self.k != self.fail_step
            self.k += 1
            return Sol(ompc.pd_test_law(x, self.xt, self.gains), ok)

    cfg = MC.SimulationConfig(dt=0.1, max_time=20.0, altitude_mean=30.0, altitude_std=5.0, horizontal_std=3.0,
                              velocity_mean=np.array([-3.0, 0.0, 0.0]), velocity_std=np.array([1.0, 0.5, 0.5]))
    order = list(MC.LandingOutcome)
    out = {}
    for kind, cls in (("solve", SolveCtl), ("step", StepCtl)):
        oc, xf, nst, fu, x0s, tg = [], [], [], [], [], []
        for i in range(32):
            gains, fail_step = ompc.pd_test_gains(i)
            ctl = cls(gains, fail_step)
            sim = MC.MonteCarloSimulator(EulerDynamics(), controller=ctl, config=cfg)
            r = sim.run_single(i, seed=42 + i)
            oc.append(order.index(r.outcome) + 1)
            xf.append(r.states[-1]); nst.append(len(r.controls)); fu.append(r.fuel_used); x0s.append(r.initial_state)
            if kind == "solve":
                t = np.zeros((3, 7))
                t[:min(3, len(ctl.targets))] = np.array(ctl.targets[:3]).reshape(-1, 7)
                tg.append(t)
        out[kind + "_outcome"] = np.array(oc); out[kind + "_final"] = np.array(xf)
        out[kind + "_nsteps"] = np.array(nst); out[kind + "_fuel"] = np.array(fu)
        if kind == "solve":
            out["x0"] = np.array(x0s); out["solve_targets"] = np.array(tg)
    print("   mc_loop outcomes (.solve):", np.bincount(out["solve_outcome"], minlength=7)[1:],
          "(.step):", np.bincount(out["step_outcome"], minlength=7)[1:])
    save("mc_loop.npz", **out)


def rti_loop_fixture(save):
    """13. the reference's own RTI orchestration (osqp_rti.py:403-599: initialize / prepare / feedback / step, the status
    handling and fall-back of :532-552, the shifted warm start) run END TO END in closed loop, with the absent `osqp` module
    replaced by a solver-backed stand-in: `osqp.OSQP` = a thin adapter that hands the reference's setup / update / warm_start
    / solve calls to oracle.mpc.OSQPOracle and answers with OSQP 0.6's result object (x, y, info.status strings, info.iter,
    info.obj_val).  What this pins is everything AROUND the solver: which matrices the reference builds at every step, what
    it warm-starts from, what it does with each status -- the solver itself stays the restatement ("parity unpinned").
    `update(Ax=A.data)` carries only the values; the reference's sparsity pattern changes when a filtered entry (|a| <= 1e-10,
    :301,309) becomes non-zero, which real OSQP would reject -- the adapter reads the matrix the reference just built from
    an instrumented `_build_constraint_matrix`, i.e. behaves like a solver that is set up again on a pattern change."""
    import src.mpc.osqp_rti as R
    from oracle import mpc as ompc
    from oracle.mpc import EulerDynamics

    STATUS = {ompc.SOLVED: "solved", ompc.SOLVED_INACCURATE: "solved inaccurate", ompc.MAX_ITER: "maximum iterations reached",
              ompc.PRIMAL_INFEASIBLE: "primal infeasible", ompc.PRIMAL_INFEASIBLE_INACCURATE: "primal infeasible inaccurate",
              ompc.DUAL_INFEASIBLE: "dual infeasible", ompc.DUAL_INFEASIBLE_INACCURATE: "dual infeasible inaccurate"}
    last_A = {}

    class Adapter:
        def setup(self, P, q, A, l, u, **kw):
            st = ompc.OSQPSettings(max_iter=kw["max_iter"], eps_abs=kw["eps_abs"], eps_rel=kw["eps_rel"],
                                   scaling=kw["scaling"], warm_start=kw["warm_start"])
            self.o = ompc.OSQPOracle(np.asarray(P.todense()).diagonal().copy(), np.asarray(q, float), A.toarray(), l, u, st)

        def update(self, q=None, l=None, u=None, Ax=None):
            A = None
            if Ax is not None:
                A = last_A["A"].toarray()
                assert np.array_equal(last_A["A"].data, Ax)
            self.o.update(q=q, l=l, u=u, A=A)

        def warm_start(self, x=None, y=None):
            self.o.warm_start(x=x, y=y)

        def solve(self):
            r = self.o.solve()
            info = types.SimpleNamespace(status=STATUS[r.status], iter=r.iters, obj_val=r.obj_val, status_val=r.status)
            return types.SimpleNamespace(x=r.x, y=r.y, info=info)

    stub = sys.modules["osqp"].OSQP
    sys.modules["osqp"].OSQP = Adapter
    out = {}
    try:
        cases = (("vertical_it50", [2.0, 30, 0, 0, -3, 0, 0], 50), ("lateral_it1000", [2.0, 30, 3, -2, -3, 0.5, 0.2], 1000),
                 ("readme_infeasible_it1000", [2.0, 500, 50, 0, -75, 10, 0], 1000))
        for tag, x0, max_iter in cases:
            x0 = np.array(x0, float)
            xt = np.array([x0[0], 0, 0, 0, 0, 0, 0.0])
            dyn = EulerDynamics()
            ctl = R.FastRTI3DoF(dyn, R.OSQPRTIConfig(N=20, osqp_max_iter=max_iter))
            orig = ctl._build_constraint_matrix

            def rec(X, U, x, _orig=orig):
                A, l, u = _orig(X, U, x)
                last_A["A"] = A
                return A, l, u

            ctl._build_constraint_matrix = rec
            ctl.initialize(x0, xt)
            x = x0.copy()
            rows = {k: [] for k in ("x", "u0", "X_opt", "U_opt", "cost", "iters", "success", "nnzA")}
            for _ in range(6):
                sol = ctl.step(x)
                rows["x"].append(x.copy()); rows["u0"].append(np.array(sol.u0, float)); rows["X_opt"].append(np.array(sol.X_opt))
                rows["U_opt"].append(np.array(sol.U_opt)); rows["cost"].append(float(sol.cost))
                rows["iters"].append(int(sol.osqp_iterations)); rows["success"].append(bool(sol.success))
                rows["nnzA"].append(int(last_A["A"].nnz))
                x = dyn.step(x, sol.u0, 0.1)
            for k, v in rows.items():
                out[f"{tag}_{k}"] = np.array(v)
            out[f"{tag}_x0"] = x0; out[f"{tag}_xt"] = xt; out[f"{tag}_max_iter"] = max_iter
            print("rti_loop", tag, rows["success"], rows["iters"], rows["nnzA"])
    finally:
        sys.modules["osqp"].OSQP = stub
    save("rti_loop.npz", **out)


def gpmpc_skeleton_fixture(G, save):
    """14. the reference's own `GPMPC.solve` (gp_mpc.py:229-369) -- initial controls, GP-augmented rollout, uncertainty
    propagation, per-iteration linearisation list and affine GP terms, the step-size convergence test, warm-start store,
    `success` / `iterations` / `status` -- run over three consecutive control steps, with ONLY `_solve_qp` (the CasADi/IPOPT
    sub-problem, absent) replaced: the replacement receives exactly what the reference passes (x0, x_target, X_nom, U_nom,
    A_list, B_list, c_list, ...) and solves the OSQP-form QP of `north_star` (osqp_rti.py cost and box, deviation-form
    dynamics rows of gp_mpc.py:402-411 written in absolute variables) with oracle.mpc.OSQPOracle.  `casadi` is an empty
    stub module so that the reference's modules import.  7-state model inside the reference's 14-state arrays."""
    sys.modules.setdefault("casadi", types.ModuleType("casadi"))
    import src.mpc.gp_mpc as GM
    from oracle import mpc as ompc
    from oracle.mpc import Rocket3DoFParams, linearize_3dof, step_3dof
    X, U, Dres = gp_fixture(300, 21)
    gp = G.Simple3DoFGP(use_sparse=False, noise_variance=1e-3); gp.add_data(X, U, Dres); gp.fit()
    p = Rocket3DoFParams()
    N, dt = 20, 0.1

    class Dyn14:
        params = p

        def step(self, x, u, dt):
            xn = x.copy(); xn[:7] = step_3dof(x[:7], u, dt, p); return xn

        def linearize(self, x, u, dt=None):
            A = np.eye(14); B = np.zeros((14, 3)); A[:7, :7], B[:7] = linearize_3dof(x[:7], u, dt, p); return A, B

    class GP14:
        def predict(self, x, u):
            m, v = gp.predict(x[:7], u); return m, np.zeros(3), v, np.zeros(3)

    log = []
    first = {}
    state = {"y": None, "rho": None}

    def solve_qp(x0, x_target, X_nom, U_nom, A_list, B_list, c_list, uncertainty, X_ref, U_ref):
        n, m = ompc.qp_dims(N)
        x_ref = np.tile(x_target[:7], (N + 1, 1)) if X_ref is None else X_ref[:N + 1, :7]
        Pd, q = ompc.build_cost(N, x_ref)
        A = np.zeros((m, n)); l = np.zeros(m); u = np.zeros(m)
        A[:7, :7] = np.eye(7); l[:7] = u[:7] = x0[:7]
        row, idx = 7, 0
        for k in range(N):
            Ak, Bk, ck = A_list[k][:7, :7], B_list[k][:7], c_list[k][:7]
            A[row:row + 7, idx:idx + 7] = np.where(np.abs(Ak) > 1e-10, Ak, 0.0)
            A[row:row + 7, idx + 7:idx + 10] = np.where(np.abs(Bk) > 1e-10, Bk, 0.0)
            A[row:row + 7, idx + 10:idx + 17] = -np.eye(7)
            l[row:row + 7] = u[row:row + 7] = -(X_nom[k + 1, :7] - Ak @ X_nom[k, :7] - Bk @ U_nom[k] + ck)
            row += 7; idx += 10
        A[row:, :] = np.eye(n)
        lo = np.concatenate([np.tile(np.concatenate([ompc.X_MIN, ompc.U_MIN]), N), ompc.X_MIN])
        hi = np.concatenate([np.tile(np.concatenate([ompc.X_MAX, ompc.U_MAX]), N), ompc.X_MAX])
        l[row:] = lo; u[row:] = hi
        if "qp0_A" not in first:               # the first sub-problem as the reference's arguments define it (a18)
            first.update(qp0_A=A.copy(), qp0_l=l.copy(), qp0_u=u.copy(), qp0_q=q.copy(), qp0_Pd=Pd.copy(),
                         qp0_X=X_nom[:, :7].copy(), qp0_U=U_nom.copy(), qp0_c=np.stack([c[:7] for c in c_list]),
                         qp0_x0=x0[:7].copy(), qp0_xref=x_ref.copy())
        qp = ompc.OSQPOracle(Pd, q, A, l, u, ompc.OSQPSettings(max_iter=1000))
        if state["rho"] is not None:          # one solver object per controller: its adapted rho and duals persist
            qp.rho = state["rho"]; qp._set_rho_vec(); qp._factor()
        qp.warm_start(x=ompc.traj_to_vec(X_nom[:, :7], U_nom), y=state["y"])
        r = qp.solve()
        state["rho"] = r.rho
        log.append((r.status, r.iters))
        if r.status != ompc.SOLVED:            # gp_mpc.py:478-482: nominal trajectory, cost = inf
            return X_nom, U_nom, np.inf
        state["y"] = r.y
        Xs, Us = ompc.vec_to_traj(r.x, N)
        X_new = X_nom.copy(); X_new[:, :7] = Xs
        cost = float(np.sum(ompc.Q_DIAG * (Xs[:N] - x_ref[:N]) ** 2) + np.sum(ompc.R_DIAG * Us**2)
                     + np.sum(10 * ompc.Q_DIAG * (Xs[N] - x_ref[N]) ** 2))
        return X_new, Us, cost

    ctl = GM.GPMPC(Dyn14(), GP14(), GM.GPMPCConfig(N=N, dt=dt))
    ctl._solve_qp = solve_qp
    ctl.setup()
    x = np.zeros(14); x[:7] = [1.9, 28, 2.0, -1.5, -3.1, 0.2, -0.1]; x[7] = 1.0
    U_ref = np.tile([x[0] * p.g0, 0.0, 0.0], (N, 1))       # hover along the 3-DoF "up" axis for the first call (:268-269)
    out = {k: [] for k in ("x", "xt", "success", "iterations", "X_opt", "U_opt", "cost", "qp_n")}
    for step in range(3):
        xt = x.copy(); xt[4:7] = 0; xt[1] = max(0.5, x[1] - 2.0)         # monte_carlo.py:497-499
        n0 = len(log)
        sol = ctl.solve(x, xt, U_ref=U_ref)
        out["x"].append(x[:7].copy()); out["xt"].append(xt[:7].copy()); out["success"].append(bool(sol.success))
        out["iterations"].append(int(sol.iterations)); out["X_opt"].append(sol.X_opt[:, :7].copy())
        out["U_opt"].append(sol.U_opt.copy()); out["cost"].append(float(sol.cost)); out["qp_n"].append(len(log) - n0)
        print("gpmpc_skeleton", step, sol.success, sol.iterations, sol.status, log[n0:])
        x = Dyn14().step(x, sol.U_opt[0], dt)
    save("gpmpc_skeleton.npz", X=X, U=U, D=Dres, noise=1e-3, U_ref=U_ref, qp_status=np.array([s for s, _ in log]),
         qp_iters=np.array([i for _, i in log]), **first, **{k: np.array(v) for k, v in out.items()})


def safety_fixture(save):
    """13. the reference's own predictive safety filter (src/safety/safety_filter.py:184-361, the gradient path
    `_solve_gradient_qp` that runs when CasADi is absent) around a 14-state embedding of the 7-state Euler model: backup
    gains / Lyapunov matrix as the reference's LQRBackupController produced them, the sampled invariant-set level, and
    (x, u_nominal) -> (u_safe, is_modified, is_safe, safety_margin, backup_value) for 96 seeded cases."""
    import io
    import contextlib
    import types as _t
    for name in ("src.safety",):
        pkg = _t.ModuleType(name)
        pkg.__path__ = [os.path.join(REF, *name.split(".")[1:])]
        sys.modules[name] = pkg
    import src.safety.safety_filter as SF
    from oracle.mpc import EulerDynamics, Rocket3DoFParams, step_3dof, linearize_3dof

    class Embed14:
        """14-state view of the 7-state model: [m, r, v | q = (1,0,0,0), w = 0] with inert attitude states."""
        def __init__(self):
            self.params = Rocket3DoFParams()
            self.params.g_I = self.params.g_vec

        def step(self, x, u, dt=0.1):
            xn = np.asarray(x, float).copy()
            xn[:7] = step_3dof(xn[:7], np.asarray(u, float), dt, self.params)
            return xn

        def linearize(self, x, u, dt=0.1):
            A = np.eye(14); B = np.zeros((14, 3))
            A7, B7 = linearize_3dof(np.asarray(x, float)[:7], np.asarray(u, float), dt, self.params)
            A[:7, :7] = A7; B[:7] = B7
            return A, B

    class CP:
        T_min, T_max, theta_max, gamma_gs = 0.3, 5.0, 60.0, 30.0

    out = {}
    for tag, alpha in (("a25", 25.0), ("a100", 100.0)):
        sf = SF.PredictiveSafetyFilter(Embed14(), SF.SafetyFilterConfig(N=10, dt=0.1), CP())
        x_eq = np.zeros(14); x_eq[0] = 1.8; x_eq[1] = 5.0; x_eq[7] = 1.0
        buf = io.StringIO()
        with contextlib.redirect_stdout(buf):
            sf.initialize(x_eq, backup_alpha=alpha)
        out[f"{tag}_init_log"] = np.array(buf.getvalue())
        rng = np.random.default_rng(17)
        n = 48
        X = np.zeros((n, 14)); X[:, 7] = 1.0
        X[:, 0] = rng.uniform(1.5, 2.0, n); X[:, 1] = rng.uniform(3.0, 8.0, n)
        X[:, 2:4] = rng.normal(0, 0.8, (n, 2)); X[:, 4] = rng.normal(-0.5, 1.5, n); X[:, 5:7] = rng.normal(0, 0.7, (n, 2))
        U = np.stack([rng.uniform(0.1, 5.5, n), rng.normal(0, 1.0, n), rng.normal(0, 1.0, n)], 1)
        res = [sf.filter(X[i], U[i]) for i in range(n)]
        out[f"{tag}_K"] = sf.backup.K[:, :7].copy(); out[f"{tag}_P"] = sf.backup.P[:7, :7].copy()
        out[f"{tag}_Kfull_rest"] = np.abs(sf.backup.K[:, 7:]).max(); out[f"{tag}_alpha"] = sf.invariant_set.alpha
        out[f"{tag}_x_eq"] = sf.backup.x_eq[:7].copy(); out[f"{tag}_u_eq"] = sf.backup.u_eq.copy()
        out[f"{tag}_X"] = X[:, :7].copy(); out[f"{tag}_U"] = U
        out[f"{tag}_u_safe"] = np.array([r.u_safe for r in res])
        out[f"{tag}_modified"] = np.array([r.is_modified for r in res]); out[f"{tag}_safe"] = np.array([r.is_safe for r in res])
        out[f"{tag}_margin"] = np.array([r.safety_margin for r in res]); out[f"{tag}_backup_value"] = np.array([r.backup_value for r in res])
        print(f"   safety {tag}: modified {int(out[f'{tag}_modified'].sum())}/{n}, safe after {int(out[f'{tag}_safe'].sum())}, "
              f"log: {buf.getvalue().strip()[:60]!r}")
    save("safety_filter.npz", **out)


def online_fixture(G, save):
    """11. online data path (reference src/gp/online_update.py): DataBuffer ring semantics, and the accept / update
    sequence of OnlineGPUpdater around a reference SparseGP with injected inducing points (no RNG dependence)."""
    import src.gp.online_update as OU
    rng = np.random.default_rng(21)
    F = rng.normal(size=(40, 3)); F[7] = F[6] + 1e-3; F[20] = F[15] + 2e-2
    T = rng.normal(size=(40, 2)); nov = rng.uniform(0, 1, 40)
    buf = OU.DataBuffer(max_size=12, feature_dim=3, target_dim=2)
    acc = np.array([buf.add_if_novel(F[i], T[i], float(nov[i]), threshold=0.3, min_distance=0.05) for i in range(40)])
    st = buf.get_statistics()
    rZ, rD = buf.get_recent(5)
    out = dict(buf_F=F, buf_T=T, buf_nov=nov, buf_accepted=acc, buf_Z=buf.get_features(), buf_D=buf.get_targets(),
               buf_recent_Z=rZ, buf_recent_D=rD,
               buf_stats=np.array([st["size"], st["total_added"], st["total_rejected"], st["acceptance_rate"],
                                   st["mean_novelty"]]))
    n = 160
    X = rng.normal(size=(n, 4)); U = rng.normal(size=(n, 2))
    d = np.sin(X[:, 0]) + 0.3 * U[:, 1] + 0.05 * rng.normal(size=n)
    ls = np.full(6, 3.0)
    Zind = rng.normal(size=(20, 6))
    gp = G.SparseGP(G.SquaredExponentialARD(6, 1.0, ls), n_inducing=20, noise_variance=1e-2, inducing_points=Zind.copy())
    cfg = OU.OnlineUpdateConfig(buffer_size=60, min_data_for_fit=20, update_interval=10, novelty_threshold=0.05,
                                min_distance=0.01)
    up = OU.OnlineGPUpdater(gp, cfg)
    accepted, novelty, updates = [], [], []
    for i in range(n):
        accepted.append(up.add_observation(X[i], U[i], d[i]))
        if up.should_update():
            r = up.update()
            updates.append([i, r["n_data"], r["total_updates"], float(r["status"] == "success")])
    Zb, Db = up._buffer.get_data()
    Xq = rng.normal(size=(16, 6))
    pred = gp.predict(Xq)
    stt = up.get_statistics()
    out.update(X=X, U=U, d=d, ls=ls, Zind=Zind, accepted=np.array(accepted), updates=np.array(updates), final_Z=Zb,
               final_D=Db, Xq=Xq, pred_mean=pred.mean, pred_var=pred.variance,
               stats=np.array([stt["total_updates"], stt["points_since_update"], stt["size"], stt["total_added"],
                               stt["total_rejected"]]))
    # ResidualCollector with an analytic nominal model
    rc = OU.ResidualCollector(lambda x, u, dt: x + dt * np.concatenate([[0.0], x[4:7], u, 0.5 * x[7:14]]),
                              max_samples=25)
    Xs = rng.normal(size=(30, 14)); Us = rng.normal(size=(30, 3)); Xn = Xs + 0.1 * rng.normal(size=(30, 14))
    for i in range(30):
        rc.record(Xs[i], Us[i], Xn[i], 0.1)
    a, b, c, e = rc.get_training_data()
    out.update(rc_X=Xs, rc_U=Us, rc_Xn=Xn, rc_states=a, rc_controls=b, rc_dv=c, rc_dw=e)
    save("online_update.npz", **out)


def unscented_fixture(G, save):
    """12. UncertaintyPropagator(method="unscented") (uncertainty_prop.py:179-264) on the 14-state embedding of the 7-state
    model, same GP / controls as propagation.npz."""
    import src.mpc.uncertainty_prop as UP
    from oracle.mpc import Rocket3DoFParams, step_3dof
    X, U, Dres = gp_fixture(300, 11)
    gp = G.Simple3DoFGP(use_sparse=False, noise_variance=1e-3); gp.add_data(X, U, Dres); gp.fit()
    p = Rocket3DoFParams()

    class Dyn14:
        params = p

        def step(self, x, u, dt):
            xn = x.copy(); xn[:7] = step_3dof(x[:7], u, dt, p); return xn

    class GP14:
        def predict(self, x, u):
            m, v = gp.predict(x[:7], u); return m, np.zeros(3), v, np.zeros(3)

    x0_14 = np.zeros(14); x0_14[:7] = [1.9, 25, 1, -2, -3.2, 0.3, -0.1]; x0_14[7] = 1.0
    rng = np.random.default_rng(12)
    Useq = np.stack([rng.uniform(1.5, 2.5, 20), rng.uniform(-.3, .3, 20), rng.uniform(-.3, .3, 20)], 1)
    pu = UP.UncertaintyPropagator(Dyn14(), GP14(), method="unscented").propagate(x0_14, Useq, np.eye(14) * 1e-6, dt=0.1)
    save("propagation_unscented.npz", X=X, U=U, D=Dres, noise=1e-3, x0=x0_14[:7], Useq=Useq, means=pu.means[:, :7],
         covs=pu.covariances[:, :7, :7], means_inert=pu.means[:, 7:], cov_cross_max=np.abs(pu.covariances[:, :7, 7:]).max())


def mc_propagation_fixture(G, save):
    """14. UncertaintyPropagator(method="monte_carlo") (uncertainty_prop.py:266-319) on the 14-state embedding, with the
    global numpy RNG seeded and its draw sequence replayed beforehand, so that the particles and the GP-posterior noise the
    reference consumed are part of the fixture (the device path takes them as inputs for bit-level parity; its default
    generator is torch's Philox)."""
    import src.mpc.uncertainty_prop as UP
    from oracle.mpc import Rocket3DoFParams, step_3dof
    X, U, Dres = gp_fixture(300, 11)
    gp = G.Simple3DoFGP(use_sparse=False, noise_variance=1e-3); gp.add_data(X, U, Dres); gp.fit()
    p = Rocket3DoFParams()

    class Dyn14:
        params = p

        def step(self, x, u, dt):
            xn = x.copy(); xn[:7] = step_3dof(x[:7], u, dt, p); return xn

    class GP14:
        def predict(self, x, u):
            m, v = gp.predict(x[:7], u); return m, np.zeros(3), v, np.zeros(3)

    x0_14 = np.zeros(14); x0_14[:7] = [1.9, 25, 1, -2, -3.2, 0.3, -0.1]; x0_14[7] = 1.0
    rng = np.random.default_rng(12)
    N, S = 8, 100
    Useq = np.stack([rng.uniform(1.5, 2.5, N), rng.uniform(-.3, .3, N), rng.uniform(-.3, .3, N)], 1)
    Sigma0 = np.eye(14) * 1e-4
    seed = 2024
    np.random.seed(seed)                       # replay of the reference's draw order (:287, :305-306)
    particles0 = np.random.multivariate_normal(x0_14, Sigma0, S)
    xi = np.zeros((N, S, 3))
    for k in range(N):
        for i in range(S):
            xi[k, i] = np.random.randn(3)
            np.random.randn(3)                 # the rotational draw (unused by the 7-state model)
    np.random.seed(seed)
    pu = UP.UncertaintyPropagator(Dyn14(), GP14(), method="monte_carlo").propagate(x0_14, Useq, Sigma0, dt=0.1)
    save("propagation_mc.npz", X=X, U=U, D=Dres, noise=1e-3, x0=x0_14[:7], Useq=Useq, sigma0=1e-4, particles0=particles0[:, :7],
         xi=xi, means=pu.means[:, :7], covs=pu.covariances[:, :7, :7])


def main():
    Stub = install_shim()
    import src.gp as G
    from src.gp.fast_gp import FastGPPredictor
    os.makedirs(OUT, exist_ok=True)
    save = lambda name, **kw: np.savez_compressed(os.path.join(OUT, name), **kw)
    if "--only-online" in sys.argv:   # add the newest fixture without rewriting the others
        online_fixture(G, save)
        return
    if "--only-rti-loop" in sys.argv:
        rti_loop_fixture(save)
        return
    if "--only-gpmpc-skeleton" in sys.argv:
        gpmpc_skeleton_fixture(G, save)
        return
    if "--only-unscented" in sys.argv:
        unscented_fixture(G, save)
        return

    # ---- 1. kernels ------------------------------------------------------------------
    rng = np.random.default_rng(10)
    D = 11
    X1 = rng.normal(0, 1, (37, D)); X2 = rng.normal(0, 1, (23, D))
    ls = rng.uniform(0.5, 2.0, D); sv = 1.7
    out = dict(X1=X1, X2=X2, ls=ls, sv=sv)
    for name, cls in (("se", G.SquaredExponentialARD), ("m32", G.Matern32), ("m52", G.Matern52)):
        k = cls(D, signal_variance=sv, lengthscales=ls)
        out[f"{name}_self"] = k(X1); out[f"{name}_cross"] = k(X1, X2); out[f"{name}_diag"] = k.diagonal(X1)
    gr = G.SquaredExponentialARD(D, sv, ls).gradients(X1, X2)
    out["se_grad_names"] = np.array(list(gr.keys()))
    out["se_grad"] = np.stack(list(gr.values()))
    out["se_params"] = G.SquaredExponentialARD(D, sv, ls).get_params()
    save("kernels.npz", **out)

    # ---- 2. Simple3DoFGP exact, n=1000, noise 1e-4 (default, cond~9e6) and 1e-2 ----------
    X, U, Dres = gp_fixture()
    Xq, Uq = query_fixture(X, U, 64, 1)
    out = dict(X=X, U=U, D=Dres, Xq=Xq, Uq=Uq)
    for tag, noise in (("n1e-4", 1e-4), ("n1e-2", 1e-2)):
        gp = G.Simple3DoFGP(use_sparse=False, noise_variance=noise)
        gp.add_data(X, U, Dres); gp.fit()
        means = np.zeros((64, 3)); vars_ = np.zeros((64, 3))
        for i in range(64):
            means[i], vars_[i] = gp.predict(Xq[i], Uq[i])
        out[f"{tag}_mean"] = means; out[f"{tag}_var"] = vars_
        out[f"{tag}_alpha"] = np.stack([g._alpha for g in gp.gp.gps])
        out[f"{tag}_ymean"] = np.array([g._y_mean for g in gp.gp.gps])
        out[f"{tag}_ystd"] = np.array([g._y_std for g in gp.gp.gps])
        out[f"{tag}_lml"] = np.array([g.log_marginal_likelihood for g in gp.gp.gps])
        out[f"{tag}_Ldiag"] = np.diag(gp.gp.gps[0]._L).copy()
        out[f"{tag}_Lrow999"] = gp.gp.gps[0]._L[999].copy()
        out[f"{tag}_Lcol0"] = gp.gp.gps[0]._L[:, 0].copy()
    out["Z_train"] = gp.feature_extractor.extract_batch(X, U)
    out["Z_query"] = gp.feature_extractor.extract_batch(Xq, Uq)
    save("simple3dof_exact.npz", **out)

    # ---- 2b. extended-precision arbiter of the same fixture: the posterior in x87 long double
    # (64-bit mantissa) from the reference's feature matrices, and cond(K + noise I).  NOT a reference output: it is
    # the yardstick that tells "differs from the reference" apart from "both are within cond*eps of the truth".
    from oracle import gp as ogp
    arb = {}
    for tag, noise in (("n1e-4", 1e-4), ("n1e-2", 1e-2)):
        K = ogp.gram(ogp.SE, out["Z_train"], out["Z_train"], np.ones(11), 1.0) + noise * np.eye(1000)
        arb[f"{tag}_cond"] = np.linalg.cond(K)
        tm = np.zeros((64, 3)); tv = np.zeros((64, 3))
        for pcol in range(3):
            f = ogp.exact_fit(ogp.SE, out["Z_train"], Dres[:, pcol], np.ones(11), 1.0, noise)
            m, v = ogp.longdouble_exact_predict(f, out["Z_query"])
            tm[:, pcol] = m.astype(np.float64); tv[:, pcol] = v.astype(np.float64)
        arb[f"{tag}_truth_mean"] = tm; arb[f"{tag}_truth_var"] = tv
    save("simple3dof_arbiter.npz", **arb)

    # ---- 3. ExactGP with each kernel, n=200 (+ return_cov, no-normalise) ------------------
    rng = np.random.default_rng(3)
    n = 200
    Xs = rng.normal(0, 1, (n, 5)); ys = np.sin(Xs[:, 0]) + 0.3 * Xs[:, 1] ** 2 + rng.normal(0, 0.05, n)
    Xt = rng.normal(0, 1, (17, 5)); ls5 = rng.uniform(0.7, 1.8, 5)
    out = dict(X=Xs, y=ys, Xt=Xt, ls=ls5, sv=0.8, noise=1e-3)
    for name, cls in (("se", G.SquaredExponentialARD), ("m32", G.Matern32), ("m52", G.Matern52)):
        gp = G.ExactGP(cls(5, 0.8, ls5), noise_variance=1e-3).fit(Xs, ys)
        pr = gp.predict(Xt)
        mc, cov = gp.predict(Xt, return_cov=True)
        out.update({f"{name}_mean": pr.mean, f"{name}_var": pr.variance, f"{name}_std": pr.std, f"{name}_cov": cov,
                    f"{name}_alpha": gp._alpha, f"{name}_L": gp._L, f"{name}_lml": gp.log_marginal_likelihood,
                    f"{name}_ymean": gp._y_mean, f"{name}_ystd": gp._y_std})
    gp = G.ExactGP(G.SquaredExponentialARD(5, 0.8, ls5), noise_variance=1e-3, normalize_y=False).fit(Xs, ys)
    pr = gp.predict(Xt)
    out.update(se_nonorm_mean=pr.mean, se_nonorm_var=pr.variance, se_nonorm_lml=gp.log_marginal_likelihood)
    # jitter escalation: duplicated inputs + tiny noise make K + noise*I numerically singular
    Xd = np.vstack([Xs[:40], Xs[:40]]); yd = np.concatenate([ys[:40], ys[:40]])
    gpj = G.ExactGP(G.SquaredExponentialARD(5, 0.8, ls5 * 50), noise_variance=1e-18).fit(Xd, yd)
    prj = gpj.predict(Xt)
    out.update(jit_X=Xd, jit_y=yd, jit_ls=ls5 * 50, jit_noise=1e-18, jit_mean=prj.mean, jit_var=prj.variance,
               jit_Ldiag=np.diag(gpj._L).copy(), jit_lml=gpj.log_marginal_likelihood)
    save("exact_gp.npz", **out)

    # ---- 4. SparseGP FITC / VFE with injected inducing points -------------------------------
    rng = np.random.default_rng(4)
    n, M = 300, 40
    Xs = rng.normal(0, 1, (n, 6)); ys = np.cos(Xs[:, 0]) * Xs[:, 2] + rng.normal(0, 0.05, n)
    Z = Xs[rng.choice(n, M, replace=False)].copy(); Xt = rng.normal(0, 1, (19, 6)); ls6 = rng.uniform(0.8, 1.6, 6)
    out = dict(X=Xs, y=ys, Z=Z, Xt=Xt, ls=ls6, sv=1.3, noise=1e-3)
    for method in ("fitc", "vfe"):
        gp = G.SparseGP(G.SquaredExponentialARD(6, 1.3, ls6), n_inducing=M, noise_variance=1e-3, method=method,
                        inducing_points=Z.copy()).fit(Xs, ys)
        pr = gp.predict(Xt)
        out.update({f"{method}_mean": pr.mean, f"{method}_var": pr.variance, f"{method}_alpha": gp._alpha,
                    f"{method}_Luu": gp._L_uu, f"{method}_LB": gp._L_B, f"{method}_lml": gp.log_marginal_likelihood})
    out["fitc_Lambda"] = G.SparseGP(G.SquaredExponentialARD(6, 1.3, ls6), M, 1e-3, "fitc", Z.copy()).fit(Xs, ys)._Lambda_diag
    # Simple3DoFGP sparse path with shared injected Z (seed global np.random before kmeans2 is NOT relied on)
    X, U, Dres = gp_fixture(400, 7)
    Xq, Uq = query_fixture(X, U, 16, 8)
    gp = G.Simple3DoFGP(n_inducing=50, use_sparse=True)
    gp.add_data(X, U, Dres)
    Zf = gp.feature_extractor.extract_batch(X, U)
    Zi = Zf[np.random.default_rng(9).choice(400, 50, replace=False)].copy()
    for g in gp.gp.gps:                       # bypass MultiOutputSparseGP.fit's random re-draw (sparse_gp.py:448-451)
        g._Z = Zi.copy(); g.fit(Zf, Dres[:, gp.gp.gps.index(g)])
    gp._is_fitted = True
    ms = np.zeros((16, 3)); vs = np.zeros((16, 3))
    for i in range(16):
        ms[i], vs[i] = gp.predict(Xq[i], Uq[i])
    out.update(s3_X=X, s3_U=U, s3_D=Dres, s3_Z=Zi, s3_Xq=Xq, s3_Uq=Uq, s3_mean=ms, s3_var=vs)
    save("sparse_gp.npz", **out)

    # ---- 5. FastGPPredictor (profiler.benchmark_gp_prediction shapes, reduced) ---------------
    rng = np.random.default_rng(5)
    Xs = rng.normal(0, 1, (200, 17)); Ys = rng.normal(0, 1, (200, 6)); Xt = rng.normal(0, 1, (31, 17))
    ls17 = rng.uniform(2.0, 4.0, 17)
    fg = FastGPPredictor(); fg.fit(Xs, Ys, ls17, 1.2, 0.01)
    mean, var = fg.predict(Xt)
    save("fast_gp.npz", X=Xs, Y=Ys, Xt=Xt, ls=ls17, variance=1.2, noise=0.01, mean=mean, var=var, alpha=fg._alpha,
         Ldiag=np.diag(fg._L).copy())

    # ---- 6. features -------------------------------------------------------------------------
    rng = np.random.default_rng(6)
    X14 = rng.normal(0, 1, (50, 14)); X14[:, 0] = rng.uniform(1, 2, 50); X14[:, 1] = rng.uniform(0, 500, 50)
    X14[:, 4:7] *= 5; q = rng.normal(0, 1, (50, 4)); X14[:, 7:11] = q / np.linalg.norm(q, axis=1, keepdims=True)
    X14[3, 4:7] = 0.0                                   # speed <= 1e-3 branch (features.py:224-231)
    U3 = rng.normal(0, 2, (50, 3))
    f3 = G.Simple3DoFFeatureExtractor(); ft = G.TranslationalFeatureExtractor(); fr = G.RotationalFeatureExtractor()
    save("features.npz", X14=X14, U=U3, f3=f3.extract_batch(X14[:, :7], U3), ft=ft.extract_batch(X14, U3),
         fr=fr.extract_batch(X14, U3))

    # ---- 7. StructuredRocketGP (exact, 6 outputs) ----------------------------------------------
    rng = np.random.default_rng(7)
    n = 150
    Xs = X14[rng.integers(0, 50, n)] + rng.normal(0, 0.1, (n, 14))
    Xs[:, 7:11] /= np.linalg.norm(Xs[:, 7:11], axis=1, keepdims=True)
    Us = rng.normal(0, 2, (n, 3)); Dv = rng.normal(0, 0.1, (n, 3)); Dw = rng.normal(0, 0.05, (n, 3))
    sg = G.StructuredRocketGP(G.StructuredGPConfig(use_sparse=False, noise_variance=1e-3))
    sg.add_data(Xs, Us, Dv, Dw); sg.fit()
    res = [sg.predict(Xs[i] + 0.01, Us[i] - 0.01) for i in range(12)]
    save("structured_gp.npz", X=Xs, U=Us, Dv=Dv, Dw=Dw, dv_mean=np.stack([r[0] for r in res]),
         dw_mean=np.stack([r[1] for r in res]), dv_var=np.stack([r[2] for r in res]), dw_var=np.stack([r[3] for r in res]))

    # ---- 8. QP assembly through the reference's own osqp_rti.py (stub osqp) ----------------------
    from oracle.mpc import EulerDynamics
    import src.mpc.osqp_rti as R
    out = {}
    x0 = np.array([2.0, 30, 0, 0, -3, 0, 0]); xt = np.array([2.0, 0, 0, 0, 0, 0, 0])
    for tag, cls, N in (("fast20", R.FastRTI3DoF, 20), ("fast15", R.FastRTI3DoF, 15), ("fd20", R.OSQPRTIMPC, 20)):
        ctl = cls(EulerDynamics(), R.OSQPRTIConfig(N=N))
        ctl.initialize(x0, xt)
        kw = Stub.last
        out[f"{tag}_P"] = kw["P"].toarray(); out[f"{tag}_q"] = kw["q"]; out[f"{tag}_A"] = kw["A"].toarray()
        out[f"{tag}_l"] = kw["l"]; out[f"{tag}_u"] = kw["u"]
        out[f"{tag}_nnzA"] = kw["A"].nnz; out[f"{tag}_nnzP"] = kw["P"].nnz
        out[f"{tag}_settings"] = np.array([kw["max_iter"], kw["eps_abs"], kw["eps_rel"], kw["scaling"],
                                           float(kw["polish"]), float(kw["warm_start"])])
        # a second build at a generic linearisation point (all thrust components non-zero)
        rng = np.random.default_rng(20 + N)
        Xl = ctl._X_lin + rng.normal(0, 0.3, ctl._X_lin.shape); Xl[:, 0] = np.linspace(2.0, 1.8, N + 1)
        Ul = np.stack([rng.uniform(0.5, 4, N), rng.uniform(-1, 1, N), rng.uniform(-1, 1, N)], 1)
        xc = x0 + rng.normal(0, 0.1, 7)
        A2, l2, u2 = ctl._build_constraint_matrix(Xl, Ul, xc)
        out.update({f"{tag}_Xl": Xl, f"{tag}_Ul": Ul, f"{tag}_xc": xc, f"{tag}_A2": A2.toarray(), f"{tag}_l2": l2,
                    f"{tag}_u2": u2, f"{tag}_nnzA2": A2.nnz})
    out["x0"] = x0; out["xt"] = xt
    save("qp_assembly.npz", **out)

    # ---- 9. UncertaintyPropagator (reference class, 14-state embedding of the 7-state model) ------
    import src.mpc.uncertainty_prop as UP
    from oracle.mpc import Rocket3DoFParams, linearize_3dof, step_3dof
    X, U, Dres = gp_fixture(300, 11)
    gp = G.Simple3DoFGP(use_sparse=False, noise_variance=1e-3); gp.add_data(X, U, Dres); gp.fit()
    p = Rocket3DoFParams()

    class Dyn14:
        params = p

        def step(self, x, u, dt):
            xn = x.copy(); xn[:7] = step_3dof(x[:7], u, dt, p); return xn

        def linearize(self, x, u, dt=None):
            A = np.eye(14); B = np.zeros((14, 3)); A[:7, :7], B[:7] = linearize_3dof(x[:7], u, dt, p); return A, B

    class GP14:
        def predict(self, x, u):
            m, v = gp.predict(x[:7], u); return m, np.zeros(3), v, np.zeros(3)

    x0_14 = np.zeros(14); x0_14[:7] = [1.9, 25, 1, -2, -3.2, 0.3, -0.1]; x0_14[7] = 1.0
    rng = np.random.default_rng(12)
    Useq = np.stack([rng.uniform(1.5, 2.5, 20), rng.uniform(-.3, .3, 20), rng.uniform(-.3, .3, 20)], 1)
    pu = UP.UncertaintyPropagator(Dyn14(), GP14()).propagate(x0_14, Useq, np.eye(14) * 1e-6, dt=0.1)
    save("propagation.npz", X=X, U=U, D=Dres, noise=1e-3, x0=x0_14[:7], Useq=Useq, means=pu.means[:, :7],
         covs=pu.covariances[:, :7, :7])

    # ---- 10. Monte Carlo initial conditions / termination (reference MonteCarloSimulator) ------------
    import src.experiments.monte_carlo as MC
    cfg = MC.SimulationConfig(dt=0.1, max_time=30.0, altitude_mean=30.0, altitude_std=5.0, horizontal_std=3.0,
                              velocity_mean=np.array([-3.0, 0.0, 0.0]), velocity_std=np.array([1.0, 0.5, 0.5]))
    sim = MC.MonteCarloSimulator(EulerDynamics(), controller=None, config=cfg)
    ics = np.stack([sim.sample_initial_condition(seed=42 + i) for i in range(64)])
    lc = MC.LandingConstraints()
    rng = np.random.default_rng(13)
    states = rng.normal(0, 1, (200, 7)) * np.array([0.3, 1.0, 4, 4, 2, 1, 1]) + np.array([1.6, 0.3, 0, 0, -1, 0, 0])
    ok = np.array([lc.check_landing(s, 2.0)[0] for s in states])
    save("monte_carlo.npz", ics=ics, states=states, landing_ok=ok)
    online_fixture(G, save)
    unscented_fixture(G, save)
    mc_loop_fixture(save)
    rti_loop_fixture(save)
    safety_fixture(save)
    mc_propagation_fixture(G, save)
    gpmpc_skeleton_fixture(G, save)
    print("golden fixtures written to", os.path.abspath(OUT))
    for f in sorted(os.listdir(OUT)):
        print(f"  {f}: {os.path.getsize(os.path.join(OUT, f)) / 1024:.1f} KiB")


if __name__ == "__main__":
    main()

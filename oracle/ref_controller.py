"""TEST INFRASTRUCTURE ONLY.  The reference's own `GPMPC` (gp_mpc.py, loaded unmodified from oracle/_ref by oracle/ref_gp.py)
made runnable without its absent third-party parts, for bench.py's CPU legs at the reference's SQP loop (S <= 10):

  * dynamics  -> the in-repo Euler statement (`simdyn` is absent; oracle.mpc.step_3dof / linearize_3dof) inside the 14-state
                 arrays the reference's classes are written for (states 7..13 inert);
  * GP        -> a 4-tuple adapter around a 3-output predictor (the reference's own Simple3DoFGP when oracle/_ref has it);
  * `_solve_qp` (CasADi/IPOPT, absent) -> the OSQP-form QP of `north_star`, built from the A_list / B_list / c_list the
                 reference hands over and solved by oracle.mpc.OSQPOracle (one persistent "solver object": rho and duals kept).

Everything else -- initial controls, rollout with `_predict_with_gp`, `UncertaintyPropagator.propagate` before and after every
SQP iteration, the linearisation loop, the step-size test, warm start -- is the reference's code path (gp_mpc.py:229-369).
tests/golden/gpmpc_skeleton.npz was generated the same way (oracle/gen_golden.py::gpmpc_skeleton_fixture)."""
import numpy as np

from . import mpc as ompc


class _Dyn14:
    def __init__(self, p):
        self.params = p

    def step(self, x, u, dt):
        xn = x.copy()
        xn[:7] = ompc.step_3dof(x[:7], u, dt, self.params)
        return xn

    def linearize(self, x, u, dt=None):
        A = np.eye(14)
        B = np.zeros((14, 3))
        A[:7, :7], B[:7] = ompc.linearize_3dof(x[:7], u, dt, self.params)
        return A, B


class _GP14:
    def __init__(self, predict3):
        self._p = predict3

    def predict(self, x, u):
        m, v = self._p(x[:7], u)
        return m, np.zeros(3), v, np.zeros(3)


class ReferenceGPMPC:
    """`.solve(x0 (7,), x_target (7,)) -> (u0, success, iterations)` through the reference's GPMPC.solve."""

    def __init__(self, gm, predict3, N=20, dt=0.1, settings=None, params=None):
        self.p = params or ompc.Rocket3DoFParams()
        self.N, self.dt = N, dt
        self.settings = settings or ompc.OSQPSettings(max_iter=1000)
        self.ctl = gm.GPMPC(_Dyn14(self.p), _GP14(predict3), gm.GPMPCConfig(N=N, dt=dt))
        self.ctl._solve_qp = self._solve_qp
        self.ctl.setup()
        self._y = None
        self._rho = None
        self.qp_log = []
        lo = np.concatenate([np.tile(np.concatenate([ompc.X_MIN, ompc.U_MIN]), N), ompc.X_MIN])
        hi = np.concatenate([np.tile(np.concatenate([ompc.X_MAX, ompc.U_MAX]), N), ompc.X_MAX])
        self._box = (lo, hi)

    def _solve_qp(self, x0, x_target, X_nom, U_nom, A_list, B_list, c_list, uncertainty, X_ref, U_ref):
        N = self.N
        n, m = ompc.qp_dims(N)
        x_ref = np.tile(x_target[:7], (N + 1, 1)) if X_ref is None else X_ref[:N + 1, :7]
        Pd, q = ompc.build_cost(N, x_ref)
        A = np.zeros((m, n)); l = np.zeros(m); u = np.zeros(m)
        A[:7, :7] = np.eye(7)
        l[:7] = u[:7] = x0[:7]
        row, idx = 7, 0
        for k in range(N):
            Ak, Bk, ck = A_list[k][:7, :7], B_list[k][:7], c_list[k][:7]
            A[row:row + 7, idx:idx + 7] = np.where(np.abs(Ak) > 1e-10, Ak, 0.0)
            A[row:row + 7, idx + 7:idx + 10] = np.where(np.abs(Bk) > 1e-10, Bk, 0.0)
            A[row:row + 7, idx + 10:idx + 17] = -np.eye(7)
            l[row:row + 7] = u[row:row + 7] = -(X_nom[k + 1, :7] - Ak @ X_nom[k, :7] - Bk @ U_nom[k] + ck)
            row += 7; idx += 10
        A[row:, :] = np.eye(n)
        l[row:], u[row:] = self._box
        qp = ompc.OSQPOracle(Pd, q, A, l, u, self.settings, rho0=self._rho)
        qp.warm_start(x=ompc.traj_to_vec(X_nom[:, :7], U_nom), y=self._y)
        r = qp.solve()
        self._rho = r.rho
        self.qp_log.append((r.status, r.iters))
        if r.status != ompc.SOLVED:                       # gp_mpc.py:478-482
            return X_nom, U_nom, np.inf
        self._y = r.y
        Xs, Us = ompc.vec_to_traj(r.x, N)
        X_new = X_nom.copy()
        X_new[:, :7] = Xs
        cost = float(np.sum(ompc.Q_DIAG * (Xs[:N] - x_ref[:N]) ** 2) + np.sum(ompc.R_DIAG * Us**2)
                     + np.sum(10 * ompc.Q_DIAG * (Xs[N] - x_ref[N]) ** 2))
        return X_new, Us, cost

    def solve(self, x0, x_target):
        x = np.zeros(14); x[:7] = x0; x[7] = 1.0
        xt = np.zeros(14); xt[:7] = x_target; xt[7] = 1.0
        U_ref = np.tile([x0[0] * self.p.g0, 0.0, 0.0], (self.N, 1))     # hover along the 3-DoF "up" axis (first call only)
        sol = self.ctl.solve(x, xt, U_ref=U_ref)
        return sol.U_opt[0], bool(sol.success), int(sol.iterations)

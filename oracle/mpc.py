"""Oracle (test infrastructure): numpy restatement of the reference MPC-side numerics.

Follows /root/reference/src/mpc/{osqp_rti,gp_mpc,uncertainty_prop,nominal_mpc}.py and
/root/reference/src/experiments/monte_carlo.py.  The OSQP solve (third-party ``osqp>=0.6.0``,
absent) is restated from the algorithm the reference cites (Stellato et al. 2020, cited at
osqp_rti.py:16-17) -- parity unpinned, cross-checked by ``kkt_active_set_solve``.
The dynamics (third-party ``simdyn``, absent) follow the in-repo Euler statement
nominal_mpc.py:586-605 and the analytic Jacobians osqp_rti.py:682-708 -- parity unpinned.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np
import scipy.linalg as sla

from . import gp as ogp

NX, NU = 7, 3
OSQP_INFTY = 1e30
MIN_SCALING, MAX_SCALING = 1e-4, 1e4
RHO_MIN, RHO_MAX, RHO_TOL, RHO_EQ_OVER_INEQ = 1e-6, 1e6, 1e-4, 1e3

# status codes (OSQP constants)
SOLVED, SOLVED_INACCURATE, MAX_ITER = 1, 2, -2
PRIMAL_INFEASIBLE, PRIMAL_INFEASIBLE_INACCURATE = -3, 3
DUAL_INFEASIBLE, DUAL_INFEASIBLE_INACCURATE = -4, 4
UNSOLVED = -10


# --------------------------------------------------------------------------- dynamics
@dataclass
class Rocket3DoFParams:
    """rocket_3dof.py:33-65 defaults; alpha = 1/(I_sp*g0) (read at nominal_mpc.py:580)."""
    m_dry: float = 1.0
    m_wet: float = 2.0
    I_sp: float = 30.0
    g0: float = 1.0
    T_min: float = 0.0
    T_max: float = 6.5
    g_vec: np.ndarray = field(default_factory=lambda: np.array([-1.0, 0.0, 0.0]))

    @property
    def alpha(self):
        return 1.0 / (self.I_sp * self.g0)


def step_3dof(x, u, dt, p: Rocket3DoFParams):
    """Euler 3-DoF model, nominal_mpc.py:586-605."""
    m, r, v = x[0], x[1:4], x[4:7]
    T = np.sqrt(u[0] ** 2 + u[1] ** 2 + u[2] ** 2)
    xn = np.empty(7)
    xn[0] = m + dt * (-p.alpha * T)
    xn[1:4] = r + dt * v
    xn[4:7] = v + dt * (u / m + p.g_vec)
    return xn


def linearize_3dof(x, u, dt, p: Rocket3DoFParams):
    """Analytic discrete Jacobians, osqp_rti.py:682-708 (T_mag has +1e-10)."""
    m = x[0]
    T_mag = np.sqrt(u[0] ** 2 + u[1] ** 2 + u[2] ** 2) + 1e-10
    A = np.eye(7)
    A[1, 4] = A[2, 5] = A[3, 6] = dt
    A[4, 0] = -u[0] / m**2 * dt
    A[5, 0] = -u[1] / m**2 * dt
    A[6, 0] = -u[2] / m**2 * dt
    B = np.zeros((7, 3))
    B[0, 0] = -p.alpha * u[0] / T_mag * dt
    B[0, 1] = -p.alpha * u[1] / T_mag * dt
    B[0, 2] = -p.alpha * u[2] / T_mag * dt
    B[4, 0] = B[5, 1] = B[6, 2] = dt / m
    return A, B


def linearize_fd(x, u, dt, p, eps=1e-6):
    """Finite-difference Jacobians, osqp_rti.py:374-401."""
    A = np.zeros((7, 7)); B = np.zeros((7, 3))
    f0 = step_3dof(x, u, dt, p)
    for i in range(7):
        xp = x.copy(); xp[i] += eps
        A[:, i] = (step_3dof(xp, u, dt, p) - f0) / eps
    for i in range(3):
        up = u.copy(); up[i] += eps
        B[:, i] = (step_3dof(x, up, dt, p) - f0) / eps
    return A, B


class EulerDynamics:
    """Duck-typed dynamics object (step / linearize / params) used by the oracle and by the
    golden generator to drive the *reference* assembler (osqp_rti.py) and propagator."""

    def __init__(self, params=None):
        self.params = params or Rocket3DoFParams()

    def step(self, x, u, dt=0.1):
        return step_3dof(np.asarray(x, float), np.asarray(u, float), dt, self.params)

    def linearize(self, x, u, dt=None):
        return linearize_3dof(np.asarray(x, float), np.asarray(u, float), 0.1 if dt is None else dt, self.params)


# --------------------------------------------------------------------------- QP assembly
Q_DIAG = np.array([0.0, 10.0, 10.0, 10.0, 1.0, 1.0, 1.0])   # osqp_rti.py:171
R_DIAG = np.ones(3) * 0.01                                   # osqp_rti.py:181
X_MIN = np.array([-np.inf, -100, -100, -100, -50, -50, -50.0])  # osqp_rti.py:198-201
X_MAX = np.array([np.inf, 500, 100, 100, 50, 50, 50.0])
U_MIN = np.array([0.3, -5, -5.0])
U_MAX = np.array([5.0, 5, 5.0])


def qp_dims(N):
    n = (N + 1) * NX + N * NU
    m = NX * (N + 1) + n
    return n, m


def build_cost(N, x_ref):
    """osqp_rti.py:203-258: P = blkdiag(Q,R,...,Q_f) (dense here), q = -Q x_ref."""
    n, _ = qp_dims(N)
    Pd = np.zeros(n); q = np.zeros(n)
    idx = 0
    for k in range(N):
        Pd[idx:idx + NX] = Q_DIAG
        q[idx:idx + NX] = -Q_DIAG * x_ref[k]
        idx += NX
        Pd[idx:idx + NU] = R_DIAG
        idx += NU
    Pd[idx:idx + NX] = Q_DIAG * 10
    q[idx:idx + NX] = -(Q_DIAG * 10) * x_ref[N]
    return Pd, q


def build_constraints(N, X_lin, U_lin, x_init, dt, p, lin="analytic", gp_affine=None, mode="rti"):
    """osqp_rti.py:260-372 -> dense A (m,n), l, u.

    mode="rti": dynamics rows ``A x + B u - x+`` with l=u=+c_k, c_k = f(x_lin,u_lin) - A x_lin - B u_lin
    (sign exactly as the reference writes it).
    mode="gpmpc": deviation-form dynamics of gp_mpc.py:402-411 rewritten in absolute variables:
    x+ = A x + B u + (X_lin[k+1] - A X_lin[k] - B U_lin[k] + c_gp[k]), i.e. l=u=-(...)."""
    n, m = qp_dims(N)
    A = np.zeros((m, n)); l = np.zeros(m); u = np.zeros(m)
    A[:NX, :NX] = np.eye(NX)
    l[:NX] = u[:NX] = x_init
    row, idx = NX, 0
    linf = linearize_3dof if lin == "analytic" else linearize_fd
    for k in range(N):
        Ak, Bk = linf(X_lin[k], U_lin[k], dt, p)
        A[row:row + NX, idx:idx + NX] = np.where(np.abs(Ak) > 1e-10, Ak, 0.0)
        A[row:row + NX, idx + NX:idx + NX + NU] = np.where(np.abs(Bk) > 1e-10, Bk, 0.0)
        A[row:row + NX, idx + NX + NU:idx + 2 * NX + NU] = -np.eye(NX)
        if mode == "rti":
            c = step_3dof(X_lin[k], U_lin[k], dt, p) - Ak @ X_lin[k] - Bk @ U_lin[k]
            if gp_affine is not None:
                c = c + gp_affine[k]
        else:
            c = -(X_lin[k + 1] - Ak @ X_lin[k] - Bk @ U_lin[k] + (0 if gp_affine is None else gp_affine[k]))
        l[row:row + NX] = u[row:row + NX] = c
        row += NX; idx += NX + NU
    A[row:, :] = np.eye(n)
    idx = 0
    for k in range(N):
        l[row + idx:row + idx + NX] = X_MIN; u[row + idx:row + idx + NX] = X_MAX; idx += NX
        l[row + idx:row + idx + NU] = U_MIN; u[row + idx:row + idx + NU] = U_MAX; idx += NU
    l[row + idx:row + idx + NX] = X_MIN; u[row + idx:row + idx + NX] = X_MAX
    return A, l, u


def traj_to_vec(X, U):
    """osqp_rti.py:601-614."""
    N = U.shape[0]
    z = np.zeros(qp_dims(N)[0])
    for k in range(N):
        z[k * 10:k * 10 + 7] = X[k]; z[k * 10 + 7:k * 10 + 10] = U[k]
    z[N * 10:] = X[N]
    return z


def vec_to_traj(z, N):
    """osqp_rti.py:616-631."""
    X = np.zeros((N + 1, NX)); U = np.zeros((N, NU))
    for k in range(N):
        X[k] = z[k * 10:k * 10 + 7]; U[k] = z[k * 10 + 7:k * 10 + 10]
    X[N] = z[N * 10:]
    return X, U


# --------------------------------------------------------------------------- OSQP restatement
@dataclass
class OSQPSettings:
    """Defaults = OSQP 0.6.x defaults overridden by osqp_rti.py:471-477 (max_iter 50, eps 1e-4, scaling 3).
    adaptive_rho_interval: OSQP's default 0 means 'timing based' (non-deterministic); the restatement
    and the GPU solver use a fixed iteration interval instead."""
    rho: float = 0.1
    sigma: float = 1e-6
    alpha: float = 1.6
    max_iter: int = 50
    eps_abs: float = 1e-4
    eps_rel: float = 1e-4
    eps_prim_inf: float = 1e-4
    eps_dual_inf: float = 1e-4
    scaling: int = 3
    check_termination: int = 25
    adaptive_rho: bool = True
    adaptive_rho_interval: int = 25
    adaptive_rho_tolerance: float = 5.0
    warm_start: bool = True
    # KKT solver of the restatement: "dense" (LAPACK LU; what every parity test and golden file uses) or "sparse" (SuperLU on
    # the 0.6 %-dense quasi-definite KKT matrix, closer to the QDLDL factorisation inside the real OSQP; used by bench.py's
    # CPU legs so that the timed CPU baseline does not pay for a dense 561 x 561 factorisation the library would not do)
    linsys: str = "dense"


@dataclass
class OSQPResult:
    x: np.ndarray
    y: np.ndarray
    status: int
    iters: int
    obj_val: float
    pri_res: float
    dua_res: float
    rho: float


def _limit_scaling(v):
    v = np.where(v < MIN_SCALING, 1.0, v)
    return np.where(v > MAX_SCALING, MAX_SCALING, v)


class OSQPOracle:
    """Dense restatement of OSQP (ADMM for min 1/2 x'Px + q'x s.t. l <= Ax <= u).

    Differences from the C library, all documented in DESIGN.md: P is diagonal (vector Pd); the KKT system
    is solved by a dense LU; adaptive rho is iteration-count based; across re-scalings (update with new
    A values) the *unscaled* iterates x, y are kept and re-scaled (the C code keeps the scaled ones)."""

    def __init__(self, Pd, q, A, l, u, settings: OSQPSettings | None = None, rho0: float | None = None):
        """rho0: start from this rho (the adapted value a persistent solver object carries over) -- same result as setting
        `rho` after construction and refactorising, without factorising twice."""
        self.s = settings or OSQPSettings()
        self.n, self.m = A.shape[1], A.shape[0]
        self.Pd0 = np.asarray(Pd, float).copy()
        self.q0 = np.asarray(q, float).copy()
        self.A0 = np.asarray(A, float).copy()
        self.l0 = np.maximum(np.asarray(l, float), -OSQP_INFTY)
        self.u0 = np.minimum(np.asarray(u, float), OSQP_INFTY)
        self.rho = self.s.rho if rho0 is None else float(rho0)
        self.x = np.zeros(self.n); self.y = np.zeros(self.m); self.z = np.zeros(self.m)  # unscaled
        self._scale()
        self._factor()

    # -- scaling (OSQP scaling.c: modified Ruiz + cost normalisation)
    def _scale(self):
        n, m = self.n, self.m
        Pd, q, A = self.Pd0.copy(), self.q0.copy(), self.A0.copy()
        D = np.ones(n); E = np.ones(m); c = 1.0
        for _ in range(self.s.scaling):
            Dt = np.maximum(np.abs(Pd), np.max(np.abs(A), axis=0)) if m else np.abs(Pd)
            Et = np.max(np.abs(A), axis=1)
            Dt = 1.0 / np.sqrt(_limit_scaling(Dt)); Et = 1.0 / np.sqrt(_limit_scaling(Et))
            Pd = Dt * Pd * Dt
            A = (Et[:, None] * A) * Dt[None, :]
            q = Dt * q
            D = D * Dt; E = E * Et
            ct = np.mean(np.abs(Pd))
            qn = float(_limit_scaling(np.array([np.max(np.abs(q))]))[0])
            ct = float(_limit_scaling(np.array([max(ct, qn)]))[0])
            ct = 1.0 / ct
            Pd = Pd * ct; q = q * ct; c = c * ct
        self.D, self.E, self.c = D, E, c
        self.Pd, self.q, self.A = Pd, q, A
        self.l = E * self.l0; self.u = E * self.u0
        self._set_rho_vec()

    def _set_rho_vec(self):
        loose = (self.l < -OSQP_INFTY * MIN_SCALING) & (self.u > OSQP_INFTY * MIN_SCALING)
        eq = (self.u - self.l) < RHO_TOL
        self.rho_vec = np.where(loose, RHO_MIN, np.where(eq, RHO_EQ_OVER_INEQ * self.rho, self.rho))

    def _factor(self):
        n, m = self.n, self.m
        K = np.zeros((n + m, n + m))
        K[:n, :n] = np.diag(self.Pd + self.s.sigma)
        K[:n, n:] = self.A.T
        K[n:, :n] = self.A
        K[n:, n:] = -np.diag(1.0 / self.rho_vec)
        if self.s.linsys == "sparse":
            import scipy.sparse as sp
            import scipy.sparse.linalg as spla
            self._lu = spla.splu(sp.csc_matrix(K))
        else:
            self._lu = sla.lu_factor(K)

    # -- updates (osqp python update(): q, then bounds, then Ax -> rescale + refactor)
    def update(self, q=None, l=None, u=None, A=None):
        if q is not None:
            self.q0 = np.asarray(q, float).copy()
        if l is not None:
            self.l0 = np.maximum(np.asarray(l, float), -OSQP_INFTY)
        if u is not None:
            self.u0 = np.minimum(np.asarray(u, float), OSQP_INFTY)
        if A is not None:
            self.A0 = np.asarray(A, float).copy()
        self._scale()
        self._factor()

    def warm_start(self, x=None, y=None):
        if x is not None:
            self.x = np.asarray(x, float).copy()
            self.z = self.A0 @ self.x
        if y is not None:
            self.y = np.asarray(y, float).copy()

    # -- residuals in unscaled units (auxil.c compute_pri_res / compute_dua_res, scaled_termination=0)
    def _residuals(self, x, z, y):
        Ax = self.A @ x
        Px = self.Pd * x
        Aty = self.A.T @ y
        Ei, Di = 1.0 / self.E, 1.0 / self.D
        pri = np.max(np.abs(Ei * (Ax - z))) if self.m else 0.0
        dua = np.max(np.abs(Di * (Px + self.q + Aty))) / self.c
        eps_pri = self.s.eps_abs + self.s.eps_rel * max(np.max(np.abs(Ei * Ax)), np.max(np.abs(Ei * z)))
        eps_dua = self.s.eps_abs + self.s.eps_rel * max(np.max(np.abs(Di * Px)), np.max(np.abs(Di * Aty)),
                                                         np.max(np.abs(Di * self.q))) / self.c
        return pri, dua, eps_pri, eps_dua, Ax, Px, Aty

    def _rho_estimate(self, x, z, y):
        Ax = self.A @ x; Px = self.Pd * x; Aty = self.A.T @ y
        pri = np.max(np.abs(Ax - z)); dua = np.max(np.abs(Px + self.q + Aty))
        pri /= max(np.max(np.abs(z)), np.max(np.abs(Ax))) + 1e-10
        dua /= max(np.max(np.abs(self.q)), np.max(np.abs(Aty)), np.max(np.abs(Px))) + 1e-10
        return float(min(max(self.rho * np.sqrt(pri / (dua + 1e-10)), RHO_MIN), RHO_MAX))

    def _primal_infeasible(self, dy, eps):
        dy = dy.copy()
        up_inf = self.u > OSQP_INFTY * MIN_SCALING
        lo_inf = self.l < -OSQP_INFTY * MIN_SCALING
        dy = np.where(up_inf & lo_inf, 0.0, np.where(up_inf, np.minimum(dy, 0.0), np.where(lo_inf, np.maximum(dy, 0.0), dy)))
        nrm = np.max(np.abs(self.E * dy))
        if nrm > eps:
            lhs = np.sum(self.u * np.maximum(dy, 0.0) + self.l * np.minimum(dy, 0.0))
            if lhs < -eps * nrm:
                return np.max(np.abs((self.A.T @ dy) / self.D)) < eps * nrm
        return False

    def _dual_infeasible(self, dx, eps):
        nrm = np.max(np.abs(self.D * dx))
        if nrm > eps:
            if np.dot(self.q, dx) < -self.c * eps * nrm:
                if np.max(np.abs((self.Pd * dx) / self.D)) < self.c * eps * nrm:
                    Adx = (self.A @ dx) / self.E
                    up_inf = self.u > OSQP_INFTY * MIN_SCALING
                    lo_inf = self.l < -OSQP_INFTY * MIN_SCALING
                    bad = (~up_inf & (Adx > eps * nrm)) | (~lo_inf & (Adx < -eps * nrm))
                    return not np.any(bad)
        return False

    def solve(self):
        s = self.s
        n, m = self.n, self.m
        if not s.warm_start:
            self.x[:] = 0; self.y[:] = 0; self.z[:] = 0
        # scaled iterates
        x = self.x / self.D
        z = self.E * self.z
        y = self.c * self.y / self.E
        status = UNSOLVED
        it = 0
        dx = np.zeros(n); dy = np.zeros(m)
        for it in range(1, s.max_iter + 1):
            xp, zp = x, z
            rhs = np.concatenate([s.sigma * xp - self.q, zp - y / self.rho_vec])
            sol = self._lu.solve(rhs) if self.s.linsys == "sparse" else sla.lu_solve(self._lu, rhs)
            xt, nu = sol[:n], sol[n:]
            zt = zp + (nu - y) / self.rho_vec
            x = s.alpha * xt + (1 - s.alpha) * xp
            dx = x - xp
            zr = s.alpha * zt + (1 - s.alpha) * zp
            z = np.minimum(np.maximum(zr + y / self.rho_vec, self.l), self.u)
            dy = self.rho_vec * (zr - z)
            y = y + dy
            can_check = s.check_termination and it % s.check_termination == 0
            if can_check:
                status = self._check(x, z, y, dx, dy, approx=False)
                if status != UNSOLVED:
                    break
            if s.adaptive_rho and s.adaptive_rho_interval and it % s.adaptive_rho_interval == 0:
                rn = self._rho_estimate(x, z, y)
                if rn > self.rho * s.adaptive_rho_tolerance or rn < self.rho / s.adaptive_rho_tolerance:
                    self.rho = rn
                    self._set_rho_vec()
                    self._factor()
        if status == UNSOLVED:
            if not (s.check_termination and it % s.check_termination == 0):
                status = self._check(x, z, y, dx, dy, approx=False)
            if status == UNSOLVED:
                status = self._check(x, z, y, dx, dy, approx=True)
            if status == UNSOLVED:
                status = MAX_ITER
        pri, dua, _, _, _, Px, _ = self._residuals(x, z, y)
        # store (unscaled)
        self.x = self.D * x
        self.z = z / self.E
        self.y = self.E * y / self.c
        obj = (0.5 * np.dot(x, Px) + np.dot(self.q, x)) / self.c
        return OSQPResult(self.x.copy(), self.y.copy(), status, it, float(obj), float(pri), float(dua), self.rho)

    def _check(self, x, z, y, dx, dy, approx):
        s = self.s
        f = 10.0 if approx else 1.0
        pri, dua, eps_pri, eps_dua, *_ = self._residuals(x, z, y)
        if approx:
            eps_pri = f * s.eps_abs + f * s.eps_rel * (eps_pri - s.eps_abs) / s.eps_rel
            eps_dua = f * s.eps_abs + f * s.eps_rel * (eps_dua - s.eps_abs) / s.eps_rel
        if self._primal_infeasible(dy, f * s.eps_prim_inf):
            return PRIMAL_INFEASIBLE_INACCURATE if approx else PRIMAL_INFEASIBLE
        if self._dual_infeasible(dx, f * s.eps_dual_inf):
            return DUAL_INFEASIBLE_INACCURATE if approx else DUAL_INFEASIBLE
        if pri <= eps_pri and dua <= eps_dua:
            return SOLVED_INACCURATE if approx else SOLVED
        return UNSOLVED


def kkt_residuals(Pd, q, A, l, u, x, y):
    """Solver-independent optimality check of a QP solution (x, y): the KKT conditions of
    min 1/2 x'Px + q'x s.t. l <= Ax <= u, which are sufficient because the QP is convex.
    Returns (primal infeasibility, stationarity |Px + q + A'y|_inf, complementarity max_i |y_i| * slack_i) with the
    OSQP sign convention (y_i < 0: lower bound active, y_i > 0: upper bound active)."""
    Ax = A @ x
    prim = float(max(np.max(l - Ax), np.max(Ax - u), 0.0))
    stat = float(np.max(np.abs(Pd * x + q + A.T @ y)))
    slack_lo = np.where(np.isfinite(l) & (l > -1e29), Ax - l, 0.0)
    slack_up = np.where(np.isfinite(u) & (u < 1e29), u - Ax, 0.0)
    comp = float(max(np.max(np.maximum(-y, 0.0) * np.abs(slack_lo)), np.max(np.maximum(y, 0.0) * np.abs(slack_up))))
    return prim, stat, comp


def active_set(y, l, u, ytol):
    """(lower-active, upper-active) masks of the inequality rows from the multipliers."""
    ineq = ~((u - l) < 1e-9)
    return (y < -ytol) & ineq, (y > ytol) & ineq


# --------------------------------------------------------------------------- RTI orchestration
@dataclass
class RTISolution:
    u0: np.ndarray
    X_opt: np.ndarray
    U_opt: np.ndarray
    cost: float
    osqp_iterations: int
    success: bool
    status: int


class OSQPRTIOracle:
    """osqp_rti.py:403-599 (initialize / prepare / feedback / step), FastRTI3DoF (analytic) or base (fd)."""

    def __init__(self, params=None, N=15, dt=0.1, lin="analytic", settings=None, g=1.0):
        self.p = params or Rocket3DoFParams()
        self.N, self.dt, self.lin, self.g = N, dt, lin, g
        self.settings = settings or OSQPSettings()

    def initialize(self, x0, x_target):
        N = self.N
        self.x_ref = np.tile(x_target, (N + 1, 1))
        self.u_ref = np.zeros((N, NU)); self.u_ref[:, 0] = x0[0] * self.g       # :422-431
        self.X_lin = np.array([(1 - k / N) * x0 + (k / N) * x_target for k in range(N + 1)])  # :438-441
        self.U_lin = self.u_ref.copy()
        self.X_prev, self.U_prev = self.X_lin.copy(), self.U_lin.copy()
        Pd, q = build_cost(N, self.x_ref)
        A, l, u = build_constraints(N, self.X_lin, self.U_lin, x0, self.dt, self.p, self.lin)
        self.solver = OSQPOracle(Pd, q, A, l, u, self.settings)

    def step(self, x):
        N = self.N
        _, q = build_cost(N, self.x_ref)                                         # prepare :494-496
        A, l, u = build_constraints(N, self.X_lin, self.U_lin, x, self.dt, self.p, self.lin)  # feedback :514
        self.solver.update(q=q, l=l, u=u, A=A)
        if self.settings.warm_start:
            self.solver.warm_start(x=traj_to_vec(self.X_prev, self.U_prev))
        res = self.solver.solve()
        if res.status == SOLVED:                                                 # :532 (see DESIGN quirks)
            X, U = vec_to_traj(res.x, N)
            self.X_lin, self.U_lin = X.copy(), U.copy()
            self.X_prev = np.vstack([X[1:], X[-1:]]); self.U_prev = np.vstack([U[1:], U[-1:]])
            return RTISolution(U[0], X, U, res.obj_val, res.iters, True, res.status)
        return RTISolution(self.U_prev[0], self.X_prev, self.U_prev, np.inf, res.iters, False, res.status)


# --------------------------------------------------------------------------- uncertainty propagation
def propagate_linear(x0, U, Sigma0, dt, p, gp_predict):
    """uncertainty_prop.py:117-177 specialised to the 7-state model (velocity block 4:7 only).
    gp_predict(x,u) -> (mean(3), var(3)).  Returns means (N+1,7), covs (N+1,7,7), dmean (N,3), dvar (N,3)."""
    N = len(U)
    if Sigma0 is None:
        Sigma0 = np.eye(NX) * 1e-6
    means = np.zeros((N + 1, NX)); covs = np.zeros((N + 1, NX, NX))
    dm = np.zeros((N, 3)); dv = np.zeros((N, 3))
    means[0] = x0; covs[0] = Sigma0
    xk, Sk = np.array(x0, float), Sigma0.copy()
    for k in range(N):
        Ad, _ = linearize_3dof(xk, U[k], dt, p)
        d, var = gp_predict(xk, U[k])
        Qgp = np.zeros((NX, NX)); Qgp[4:7, 4:7] = np.diag(var) * dt**2
        xn = step_3dof(xk, U[k], dt, p); xn[4:7] += d * dt
        Sn = Ad @ Sk @ Ad.T + Qgp
        means[k + 1] = xn; covs[k + 1] = Sn; dm[k] = d; dv[k] = var
        xk, Sk = xn, Sn
    return means, covs, dm, dv


def propagate_unscented(x0, U, Sigma0, dt, p, gp_predict, n_aug=14):
    """uncertainty_prop.py:179-264 restricted to the 7 states of the 3-DoF model.  The reference class carries 14 states
    (n = 14 fixes lambda and the weights; 29 sigma points); for the 3-DoF embedding the 7 extra states are inert, their
    block of Sigma never couples with the first 7, and the 14 sigma points that perturb only them propagate exactly like
    the centre point.  So: 7-state Cholesky, 15 distinct points, the centre counted 1 + 14 times, terms added in the
    reference's order (rows 0..28 of its sigma-point array: centre, +real, +inert, -real, -inert).
    alpha = 1e-3 makes w_m[0] ~ -1e6: rounding in the sigma-point images is amplified by 1e6 in the mean."""
    N = len(U)
    n = n_aug
    ni = n_aug - NX
    if Sigma0 is None:
        Sigma0 = np.eye(NX) * 1e-6
    alpha, beta, kappa = 1e-3, 2, 0
    lam = alpha**2 * (n + kappa) - n
    wm0 = lam / (n + lam)
    wc0 = lam / (n + lam) + (1 - alpha**2 + beta)
    wi = 1 / (2 * (n + lam))
    means = np.zeros((N + 1, NX)); covs = np.zeros((N + 1, NX, NX))
    means[0] = x0; covs[0] = Sigma0
    xk, Sk = np.array(x0, float), Sigma0.copy()

    def image(x, u):
        xn = step_3dof(x, u, dt, p)
        d, _ = gp_predict(x, u)
        xn[4:7] += d * dt
        return xn

    for k in range(N):
        Ls = np.linalg.cholesky((n + lam) * Sk + 1e-10 * np.eye(NX))
        c = image(xk, U[k])
        plus = [image(xk + Ls[:, i], U[k]) for i in range(NX)]
        minus = [image(xk - Ls[:, i], U[k]) for i in range(NX)]
        pts = [c] + plus + [c] * ni + minus + [c] * ni
        xn = wm0 * pts[0]
        for q in pts[1:]:
            xn = xn + wi * q
        Sn = np.zeros((NX, NX))
        for i, q in enumerate(pts):
            df = q - xn
            Sn = Sn + (wc0 if i == 0 else wi) * np.outer(df, df)
        _, var = gp_predict(xk, U[k])
        Sn[4:7, 4:7] += np.diag(var) * dt**2
        means[k + 1] = xn; covs[k + 1] = Sn
        xk, Sk = xn, Sn
    return means, covs


def propagate_monte_carlo(x0, U, Sigma0, dt, p, gp_predict, particles0, xi):
    """uncertainty_prop.py:266-319 for the 7-state model with the randomness injected: particles0 (S,7) = the draw of
    :287, xi (N,S,3) = the standard-normal draws of :305 (the reference takes both from the global numpy RNG)."""
    N, S = len(U), particles0.shape[0]
    means = np.zeros((N + 1, NX)); covs = np.zeros((N + 1, NX, NX))
    means[0] = x0; covs[0] = Sigma0
    particles = np.array(particles0, dtype=float)
    for k in range(N):
        nxt = np.zeros_like(particles)
        for i in range(S):
            x_nom = step_3dof(particles[i], U[k], dt, p)
            d, var = gp_predict(particles[i], U[k])
            nxt[i] = x_nom
            nxt[i, 4:7] += (d + np.sqrt(var) * xi[k, i]) * dt
        means[k + 1] = np.mean(nxt, axis=0)
        diff = nxt - means[k + 1]
        covs[k + 1] = (diff.T @ diff) / (S - 1)
        particles = nxt
    return means, covs


# --------------------------------------------------------------------------- GP-MPC (composed 3-DoF path)
@dataclass
class MPCSolutionOracle:
    success: bool
    X_opt: np.ndarray
    U_opt: np.ndarray
    cost: float
    iterations: int
    status: str
    qp_status: list
    qp_iters: list
    covariances: np.ndarray


class GPMPCOracle:
    """gp_mpc.py:229-369 specialised to 7 states, with the inner IPOPT 'QP' replaced by the OSQP-form QP
    of osqp_rti.py: box bounds only; GP mean as affine term c_k=[0,0,0,0,d_v*dt]."""

    def __init__(self, gp_predict, params=None, N=20, dt=0.1, settings=None, max_sqp_iter=10, use_gp_mean=True,
                 keep_rho=True, cone: "ConeRows | None" = None):
        self.gp_predict = gp_predict
        self.cone = cone          # polyhedral thrust / glide-slope rows: the QP is solved in the augmented stage form
        self.p = params or Rocket3DoFParams()
        self.N, self.dt = N, dt
        self.settings = settings or OSQPSettings()
        self.max_sqp_iter, self.use_gp_mean = max_sqp_iter, use_gp_mean
        self.U_warm = None
        self.y_warm = None
        # one OSQP solver object per controller lives across solves (as osqp_rti.py:464-478/517 does): its adapted
        # rho persists; the primal iterate is re-seeded from (X, U) and the duals from the last *solved* QP
        self.keep_rho = keep_rho
        self.rho_warm = None

    def solve(self, x0, x_target):
        N, dt, p = self.N, self.dt, self.p
        x0 = np.asarray(x0, float)
        X = np.zeros((N + 1, NX)); U = np.zeros((N, NU)); X[0] = x0
        if self.U_warm is not None:                                   # gp_mpc.py:266-275
            U_init = self.U_warm
        else:
            # hover thrust along the "up" axis of the 3-DoF frame (axis 0: nominal_mpc.py:651-657,
            # osqp_rti.py:427-431); gp_mpc.py:273 writes [0,0,m*g0] in the 6-DoF body frame
            U_init = np.tile(np.array([x0[0] * p.g0, 0, 0]), (N, 1))
        for k in range(N):                                             # gp_mpc.py:278-281
            U[k] = U_init[k]
            xn = step_3dof(X[k], U[k], dt, p)
            if self.use_gp_mean:
                d, _ = self.gp_predict(X[k], U[k]); xn[4:7] += d * dt
            X[k + 1] = xn
        _, covs, _, _ = propagate_linear(x0, U, np.eye(NX) * 1e-6, dt, p, self.gp_predict)  # :284-289
        x_ref = np.tile(x_target, (N + 1, 1))
        Pd, q = build_cost(N, x_ref)
        converged = False; cost = np.inf; stats = []; its = []
        y = self.y_warm
        for it in range(self.max_sqp_iter):                            # gp_mpc.py:299-353
            c_gp = np.zeros((N, NX))
            if self.use_gp_mean:
                for k in range(N):
                    d, _ = self.gp_predict(X[k], U[k]); c_gp[k, 4:7] = d * dt
            A, l, u = build_constraints(N, X, U, x0, dt, p, "analytic", c_gp, mode="gpmpc")
            rho0 = self.rho_warm if (self.keep_rho and self.rho_warm is not None) else None
            if self.cone is not None:
                AB = np.stack([A[NX + k * NX:NX + (k + 1) * NX, k * 10:(k + 1) * 10] for k in range(N)])
                ABc, beqc, qc, zc = augment_cone_qp(N, AB, l[:NX * (N + 1)], q, X, U, self.cone)
                Pdc, lbc, ubc = cone_shared_data(N, self.cone)
                Ac, lc, uc = dense_stage_constraints(NXC, NU, N, ABc, beqc, lbc, ubc)
                qp = OSQPOracle(Pdc, qc, Ac, lc, uc, self.settings, rho0=rho0)
            else:
                qp = OSQPOracle(Pd, q, A, l, u, self.settings, rho0=rho0)
            qp.warm_start(x=traj_to_vec(X, U) if self.cone is None else zc, y=y)
            res = qp.solve()
            self.rho_warm = res.rho
            stats.append(res.status); its.append(res.iters)
            if res.status == SOLVED:
                zx = res.x
                if self.cone is not None:      # [x; xi; u] stages -> [x; u] stages
                    zx = np.concatenate([np.r_[res.x[k * NSC:k * NSC + NX], res.x[k * NSC + NXC:(k + 1) * NSC]]
                                         for k in range(N)] + [res.x[N * NSC:N * NSC + NX]])
                Xn, Un = vec_to_traj(zx, N); y = res.y
                cost = float(np.sum(Q_DIAG * (Xn[:N] - x_ref[:N]) ** 2) + np.sum(R_DIAG * Un**2)
                             + np.sum(10 * Q_DIAG * (Xn[N] - x_ref[N]) ** 2))
            else:                                                       # gp_mpc.py:478-482: nominal, cost=inf
                Xn, Un = X, U; cost = np.inf
            dX, dU = np.max(np.abs(Xn - X)), np.max(np.abs(Un - U))
            X, U = Xn, Un
            if dX < 1e-4 and dU < 1e-4:
                converged = True
                break
            if it < self.max_sqp_iter - 1:
                _, covs, _, _ = propagate_linear(x0, U, np.eye(NX) * 1e-6, dt, p, self.gp_predict)
        self.U_warm = U.copy(); self.y_warm = y
        return MPCSolutionOracle(converged, X, U, cost, it + 1, "Converged" if converged else "Max iterations",
                                 stats, its, covs)


# --------------------------------------------------------------------------- convex rows
@dataclass
class ConeRows:
    """Polyhedral INNER approximations of the thrust-norm and glide-slope constraints the reference gives IPOPT
    (gp_mpc.py:417-429; defaults of ConstraintParams, constraints.py:35-50), see include/gpmpc_b200.h GpmpcCone3:
    u_x >= T_min (box), u_x +- b u_y <= T_max, u_x +- b u_z <= T_max, and for the next stage's position the pyramid over the
    triangle inscribed in the glide-slope circle with a vertex in the nominal radial direction."""
    T_min: float = 0.5
    T_max: float = 5.0
    gamma_gs_deg: float = 30.0

    @property
    def facet_b(self):
        return float(np.sqrt(2.0 * (self.T_max - self.T_min) / (self.T_max + self.T_min)))

    @property
    def half_tan(self):
        return float(np.tan(np.deg2rad(self.gamma_gs_deg)) / 2.0)


NXC, NSC = 2 * NX, 2 * NX + NU


def glide_normals(y, z):
    rho = np.sqrt(y * y + z * z)
    cy, cz = (y / rho, z / rho) if rho > 1e-9 else (1.0, 0.0)
    c60, s60 = 0.5, 0.8660254037844386
    return np.array([[c60 * cy - s60 * cz, s60 * cy + c60 * cz], [c60 * cy + s60 * cz, -s60 * cy + c60 * cz], [-cy, -cz]])


def cone_values(cone: ConeRows, x, u_prev):
    b = cone.facet_b
    n = glide_normals(x[2], x[3])
    return np.array([u_prev[0] + b * u_prev[1], u_prev[0] - b * u_prev[1], u_prev[0] + b * u_prev[2],
                     u_prev[0] - b * u_prev[2]] + [(n[i, 0] * x[2] + n[i, 1] * x[3]) - cone.half_tan * x[1] for i in range(3)])


def cone_dims(N):
    n = N * NSC + NXC
    return n, NXC * (N + 1) + n


def augment_cone_qp(N, AB, beq, q, X_nom, U_nom, cone: ConeRows):
    """Stage data of the 7-state QP -> stage data of the augmented QP with stage variables [x; xi; u] (the arithmetic of
    qp_augment_cone3_kernel).  AB (N,7,10), beq (7 (N+1),) = [x_init | rhs_0 | ...], q (10 N + 7,)."""
    n, _ = cone_dims(N)
    ABc = np.zeros((N, NXC, NSC)); beqc = np.zeros(NXC * (N + 1)); qc = np.zeros(n); zc = np.zeros(n)
    b, h = cone.facet_b, cone.half_tan
    beqc[:NX] = beq[:NX]
    for k in range(N + 1):
        qc[k * NSC:k * NSC + NX] = q[k * 10:k * 10 + NX]
        zc[k * NSC:k * NSC + NX] = X_nom[k]
        if k > 0:
            zc[k * NSC + NX:k * NSC + NXC] = cone_values(cone, X_nom[k], U_nom[k - 1])
        if k == N:
            break
        qc[k * NSC + NXC:(k + 1) * NSC] = q[k * 10 + NX:(k + 1) * 10]
        zc[k * NSC + NXC:(k + 1) * NSC] = U_nom[k]
        ab, be = AB[k], beq[NX + k * NX:NX + (k + 1) * NX]
        ABc[k, :NX, :NX] = ab[:, :NX]; ABc[k, :NX, NXC:] = ab[:, NX:]
        r = NXC + k * NXC
        beqc[r:r + NX] = be
        for j, (col, sg) in enumerate(((1, b), (1, -b), (2, b), (2, -b))):
            ABc[k, NX + j, NXC + 0] = 1.0; ABc[k, NX + j, NXC + col] = sg
        nrm = glide_normals(X_nom[k + 1][2], X_nom[k + 1][3])
        for i in range(3):
            row = (nrm[i, 0] * ab[2] + nrm[i, 1] * ab[3]) - h * ab[1]
            ABc[k, NX + 4 + i, :NX] = row[:NX]; ABc[k, NX + 4 + i, NXC:] = row[NX:]
            beqc[r + NX + 4 + i] = (nrm[i, 0] * be[2] + nrm[i, 1] * be[3]) - h * be[1]
    return ABc, beqc, qc, zc


def cone_shared_data(N, cone: ConeRows):
    """Pdiag, lb, ub of the augmented QP (shared by all scenarios): zero cost and the facet bounds on xi, u_x >= T_min."""
    n, _ = cone_dims(N)
    Pd = np.zeros(n); lb = np.full(n, -np.inf); ub = np.full(n, np.inf)
    umin = U_MIN.copy(); umin[0] = max(umin[0], cone.T_min)
    xi_hi = np.array([cone.T_max] * 4 + [0.0] * 3)
    for k in range(N + 1):
        o = k * NSC
        Pd[o:o + NX] = Q_DIAG * (10 if k == N else 1)
        lb[o:o + NX] = X_MIN; ub[o:o + NX] = X_MAX
        if k > 0:
            ub[o + NX:o + NX + 4] = xi_hi[:4]
        if k > 1:          # stage 1's position is x_0 + dt v_0 whatever the controls: its glide rows carry no decision
            ub[o + NX + 4:o + NXC] = xi_hi[4:]
        if k < N:
            Pd[o + NXC:o + NSC] = R_DIAG
            lb[o + NXC:o + NSC] = umin; ub[o + NXC:o + NSC] = U_MAX
    return Pd, lb, ub


def dense_stage_constraints(nx, nu, N, AB, beq, lb, ub):
    """Row order of osqp_rti.py:260-372 for any stage size: [x_0 = x_init | A_k x_k + B_k u_k - x_{k+1} = rhs_k | l <= I z <= u]."""
    ns = nx + nu
    n = N * ns + nx; m = nx * (N + 1) + n
    A = np.zeros((m, n)); l = np.zeros(m); u = np.zeros(m)
    A[:nx, :nx] = np.eye(nx)
    for k in range(N):
        r = nx + k * nx
        A[r:r + nx, k * ns:(k + 1) * ns] = AB[k]
        A[r:r + nx, (k + 1) * ns:(k + 1) * ns + nx] = -np.eye(nx)
    neq = nx * (N + 1)
    l[:neq] = u[:neq] = beq
    A[neq:, :] = np.eye(n)
    l[neq:] = lb; u[neq:] = ub
    return A, l, u


def cone_direct_rows(N, X_nom, cone: ConeRows):
    """The same polyhedral constraints written directly as inequality rows on z = [x_0,u_0,...,x_N] of the 7-state QP
    (no auxiliary variables): used by the tests to check that the stage embedding solves the same problem."""
    n, _ = qp_dims(N)
    b, h = cone.facet_b, cone.half_tan
    C = np.zeros((7 * N, n)); hi = np.zeros(7 * N)
    for k in range(N):
        for j, (col, sg) in enumerate(((1, b), (1, -b), (2, b), (2, -b))):
            C[7 * k + j, k * 10 + NX] = 1.0; C[7 * k + j, k * 10 + NX + col] = sg
            hi[7 * k + j] = cone.T_max
        nrm = glide_normals(X_nom[k + 1][2], X_nom[k + 1][3])
        o = (k + 1) * 10
        for i in range(3):
            C[7 * k + 4 + i, o + 1] = -h; C[7 * k + 4 + i, o + 2] = nrm[i, 0]; C[7 * k + 4 + i, o + 3] = nrm[i, 1]
            if k == 0:
                hi[7 * k + 4 + i] = np.inf
    return C, hi


# --------------------------------------------------------------------------- Monte Carlo sampling
@dataclass
class SimConfig:
    """run_experiments.py:359-371 distribution (the SimulationConfig class defaults are infeasible)."""
    dt: float = 0.1
    max_time: float = 30.0
    altitude_mean: float = 30.0
    altitude_std: float = 5.0
    horizontal_std: float = 3.0
    velocity_mean: tuple = (-3.0, 0.0, 0.0)
    velocity_std: tuple = (1.0, 0.5, 0.5)
    mass_mean: float = 2.0
    mass_std: float = 0.1


def sample_initial_condition(seed, cfg: SimConfig):
    """monte_carlo.py:368-399: np.random.seed(seed); 7 randn in fixed order; clips."""
    np.random.seed(seed)
    m = np.clip(cfg.mass_mean + np.random.randn() * cfg.mass_std, 1.5, 2.5)
    alt = np.clip(cfg.altitude_mean + np.random.randn() * cfg.altitude_std, 10, 100)
    hy = np.random.randn() * cfg.horizontal_std
    hz = np.random.randn() * cfg.horizontal_std
    vv = min(cfg.velocity_mean[0] + np.random.randn() * cfg.velocity_std[0], -1)
    vy = cfg.velocity_mean[1] + np.random.randn() * cfg.velocity_std[1]
    vz = cfg.velocity_mean[2] + np.random.randn() * cfg.velocity_std[2]
    return np.array([m, alt, hy, hz, vv, vy, vz])


# outcome codes (monte_carlo.py:24-33 LandingOutcome order)
OUT_RUNNING, OUT_SUCCESS, OUT_CRASH, OUT_FUEL, OUT_VIOLATION, OUT_TIMEOUT, OUT_DIVERGENCE = 0, 1, 2, 3, 4, 5, 6


def check_landing(x, m0):
    """monte_carlo.py:54-104 with default LandingConstraints."""
    if abs(x[1]) > 1.0: return False
    if abs(x[2]) > 5.0 or abs(x[3]) > 5.0: return False
    if abs(x[4]) > 2.0: return False
    if abs(x[5]) > 1.0 or abs(x[6]) > 1.0: return False
    if 1.0 - x[0] / m0 > 1.0 - 0.05: return False
    return True


def termination(x, m0):
    """monte_carlo.py:460-488: returns outcome code or OUT_RUNNING."""
    if x[1] < 0: return OUT_CRASH
    if x[0] <= 1.0 + 0.01: return OUT_FUEL
    if np.any(np.abs(x) > 1e6) or np.any(np.isnan(x)): return OUT_DIVERGENCE
    if x[1] < 1.0 and abs(x[4]) < 5.0:
        return OUT_SUCCESS if check_landing(x, m0) else OUT_VIOLATION
    return OUT_RUNNING


def run_single(solve, x0, dt=0.1, max_time=30.0, params=None, target_mode="per_step", on_solve_failure="divergence",
               drag_cda=0.0, step_fn=None):
    """monte_carlo.py:401-583, the closed loop of one run.  `solve(x, x_target) -> (u0, success)`.
    target_mode "per_step" = the branch for controllers with `.solve` (:495-499), "fixed" = the target of :428-429 that a
    `.step` controller receives through `initialize`; on_solve_failure "divergence" = :507-510, "continue" = the `.step`
    branch (:492-494), which never looks at the flag.  Returns (outcome code, final state, number of plant steps,
    number of failed solves)."""
    p = params or Rocket3DoFParams()
    x = np.asarray(x0, float).copy()
    m0 = x[0]
    x_fixed = np.zeros(NX); x_fixed[0] = m0                       # :428-429
    max_steps = int(max_time / dt)                                # :447
    n_fail = 0
    for step in range(max_steps):                                 # :455
        code = termination(x, m0)                                 # :460-488
        if code != OUT_RUNNING:
            return code, x, step, n_fail
        if target_mode == "per_step":                             # :497-499
            xt = x.copy(); xt[4:7] = 0; xt[1] = max(0.5, x[1] - 2.0)
        else:
            xt = x_fixed
        u, ok = solve(x, xt)
        if not ok:
            n_fail += 1
            if on_solve_failure == "divergence":                  # :507-510
                return OUT_DIVERGENCE, x, step, n_fail
        xn = step_fn(x, u, dt) if step_fn is not None else step_3dof(x, u, dt, p)   # :528
        if drag_cda:
            rho = 1.225 * np.exp(-x[1] / 8500.0)
            sp = np.sqrt(x[4] * x[4] + x[5] * x[5] + x[6] * x[6])
            xn[4:7] = xn[4:7] + dt * (-0.5 * rho * drag_cda * sp * x[4:7] / x[0])
        x = xn
    return OUT_TIMEOUT, x, max_steps, n_fail                      # for/else :539-542


def pd_test_law(x, xt, gains):
    """A deterministic stand-in controller for pinning the LOOP (not a reference function): PD towards the target with
    the reference's control box (osqp_rti.py:200-201).  gains = (kp_alt, kd_alt, kp_lat, kd_lat, bias)."""
    kp, kd, kpl, kdl, bias = gains
    m = x[0]
    u = np.empty(3)
    u[0] = m * (1.0 + bias + kp * (xt[1] - x[1]) + kd * (xt[4] - x[4]))
    u[1] = m * (kpl * (xt[2] - x[2]) + kdl * (xt[5] - x[5]))
    u[2] = m * (kpl * (xt[3] - x[3]) + kdl * (xt[6] - x[6]))
    return np.minimum(np.maximum(u, U_MIN), U_MAX)


def pd_test_gains(i):
    """Gains / failure step of test run i: a spread that visits every outcome class of monte_carlo.py:24-33."""
    table = [
        (0.9, 1.0, 0.3, 0.8, 0.0),     # soft descent at ~1.8 m/s
        (0.5, 1.0, 0.0, 0.0, 0.0),     # no lateral control
        (0.05, 0.1, 0.3, 0.8, -0.6),   # under-thrust: arrives fast
        (0.0, 0.0, 0.0, 0.0, -0.85),   # minimum thrust: crash
        (0.0, 2.0, 0.0, 0.0, 0.0),     # hover: timeout
        (0.5, 1.0, 0.3, 0.8, 0.0),     # solve failure at step 10 (fail_step below)
        (2.0, 0.2, 2.0, 0.1, 1.5),     # aggressive, oscillating: burns fuel
        (0.8, 1.0, 0.3, 0.8, 0.0),
    ]
    return table[i % 8], (10 if i % 8 == 5 else -1)

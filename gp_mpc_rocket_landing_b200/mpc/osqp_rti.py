"""OSQP-based RTI-MPC with the reference's API (src/mpc/osqp_rti.py: `OSQPRTIConfig`, `OSQPRTISolution`, `OSQPRTIMPC`,
`FastRTI3DoF`) plus `initialize_batch` / `step_batch` for B independent controllers.

QP assembly (osqp_rti.py:203-372) and the OSQP solve (:464-478, :496, :517, :524, :527) run in libgpmpc_b200.so: one
assembly thread per (scenario, stage), one ADMM CTA per scenario.  The solver state OSQP keeps inside its object
between solves (primal/dual iterates, adapted rho) lives in per-scenario device buffers.  Status handling follows
:532-552 -- only OSQP status "solved" counts as success (the reference's "solved_inaccurate"
string never matches OSQP 0.6's "solved inaccurate").  No CPU path.
"""
from __future__ import annotations

import time
from dataclasses import dataclass
from typing import Any, Optional, Tuple

import numpy as np

from .. import _lib
from ..gp import engine as ge
from . import engine as me

NX, NU = 7, 3


@dataclass
class OSQPRTIConfig:
    """osqp_rti.py:45-71 (+ the OSQP 0.6 defaults the reference leaves untouched, as explicit fields)."""

    N: int = 15
    dt: float = 0.1
    osqp_verbose: bool = False
    osqp_max_iter: int = 50
    osqp_eps_abs: float = 1e-4
    osqp_eps_rel: float = 1e-4
    osqp_polish: bool = False
    osqp_warm_start: bool = True
    osqp_scaling: int = 3
    max_rti_iter: int = 1
    trust_region: float = 10.0
    n_x: int = 7
    n_u: int = 3
    precompute_matrices: bool = True
    # additions
    osqp_check_termination: int = 25
    osqp_adaptive_rho_interval: int = 25   # iteration based; OSQP's default (0) is timing based / non-deterministic


@dataclass
class OSQPRTISolution:
    """osqp_rti.py:74-86."""

    u0: Any
    X_opt: Any
    U_opt: Any
    cost: float
    prep_time_ms: float
    feedback_time_ms: float
    total_time_ms: float
    osqp_iterations: int
    success: bool


@dataclass
class BatchRTISolution:
    """`step_batch` result: u0 (B,3), X_opt (B,N+1,7), U_opt (B,N,3), cost (B,), osqp_iterations (B,) int,
    success (B,) bool, status (B,) int (OSQP code), times in ms."""

    u0: Any
    X_opt: Any
    U_opt: Any
    cost: Any
    osqp_iterations: Any
    success: Any
    status: Any
    prep_time_ms: float
    feedback_time_ms: float
    total_time_ms: float

    def __getitem__(self, b: int) -> OSQPRTISolution:
        return OSQPRTISolution(u0=self.u0[b], X_opt=self.X_opt[b], U_opt=self.U_opt[b], cost=float(self.cost[b]),
                               prep_time_ms=self.prep_time_ms, feedback_time_ms=self.feedback_time_ms,
                               total_time_ms=self.total_time_ms, osqp_iterations=int(self.osqp_iterations[b]),
                               success=bool(self.success[b]))


class OSQPRTIMPC:
    """osqp_rti.py:89-639.  Base class linearises by forward finite differences (:374-401)."""

    _lin_mode = 1

    def __init__(self, dynamics, config: Optional[OSQPRTIConfig] = None):
        ge.torch_()  # fails loudly without a CUDA device (the reference raises ImportError without osqp, :129-130)
        self.dynamics = dynamics
        self.config = config or OSQPRTIConfig()
        if (self.config.n_x, self.config.n_u) != (NX, NU):
            raise _lib.GpmpcError("the device QP assembly implements the 3-DoF model (n_x=7, n_u=3)")
        self.n_x, self.n_u, self.N = NX, NU, self.config.N
        self.n_vars = (self.N + 1) * self.n_x + self.N * self.n_u
        self._setup_cost_matrices()
        self._setup_constraint_structure()
        self._solver = None
        self._B = 0
        self._batched = False

    # osqp_rti.py:168-201
    def _setup_cost_matrices(self) -> None:
        self.Q = np.diag(me.Q_DIAG)
        self.Q_f = self.Q * 10.0
        self.R = np.eye(NU) * 0.01

    def _setup_constraint_structure(self) -> None:
        N = self.N
        self.n_eq = N * NX + NX
        self.n_ineq = (N + 1) * NX + N * NU
        self.x_min, self.x_max = me.X_MIN.copy(), me.X_MAX.copy()
        self.u_min, self.u_max = me.U_MIN.copy(), me.U_MAX.copy()

    def _dyn(self):
        p = getattr(self.dynamics, "params", None)
        alpha = p.alpha if p is not None else 1.0 / 30.0
        g_vec = getattr(p, "g_vec", None)
        if g_vec is None:
            g_vec = getattr(p, "g_I", np.array([-1.0, 0.0, 0.0]))
        return me.make_dyn(alpha, np.asarray(g_vec, dtype=np.float64), self.config.dt)

    def _settings(self):
        c = self.config
        return _lib.AdmmSettings.reference_defaults(
            max_iter=c.osqp_max_iter, eps_abs=c.osqp_eps_abs, eps_rel=c.osqp_eps_rel, scaling=c.osqp_scaling,
            check_every=c.osqp_check_termination, adaptive_rho_interval=c.osqp_adaptive_rho_interval,
            warm_x=1 if c.osqp_warm_start else 0, warm_y=1 if c.osqp_warm_start else 0, rho_from_input=1)

    # ------------------------------------------------------------------------------------------------ batched core
    def initialize_batch(self, X0, x_target, X_init=None, U_init=None) -> None:
        """osqp_rti.py:403-452 for B scenarios: X0 (B,7); x_target (7,) or (B,7)."""
        torch = ge.torch_()
        N = self.N
        x0 = ge.to_dev(X0)
        B = x0.shape[0]
        xt = ge.to_dev(x_target)
        xt = xt[None, :].expand(B, NX) if xt.dim() == 1 else xt
        f64 = dict(dtype=torch.float64, device=x0.device)
        self._B = B
        self._dynp = self._dyn()
        self._qp = me.QPData.reference(N, Q=np.diag(self.Q), Qf=np.diag(self.Q_f), R=np.diag(self.R),
                                       x_min=self.x_min, x_max=self.x_max, u_min=self.u_min, u_max=self.u_max)
        self._x_ref = xt[:, None, :].expand(B, N + 1, NX).contiguous()
        self._u_ref = torch.zeros((B, N, NU), **f64)
        g = getattr(getattr(self.dynamics, "params", None), "g", 1.0)  # :428 (simdyn exposes g0/g_vec -> 1.0)
        self._u_ref[:, :, 0] = (x0[:, 0] * g)[:, None]
        if X_init is not None:
            self._X_lin = ge.to_dev(X_init).expand(B, N + 1, NX).contiguous().clone()
        else:  # straight line x0 -> x_target, :438-441
            a = (torch.arange(N + 1, **f64) / N)[None, :, None]
            self._X_lin = ((1 - a) * x0[:, None, :] + a * xt[:, None, :]).contiguous()
        self._U_lin = (ge.to_dev(U_init).expand(B, N, NU).contiguous().clone() if U_init is not None
                       else self._u_ref.clone())
        self._X_prev, self._U_prev = self._X_lin.clone(), self._U_lin.clone()
        n, m = me.qp_dims(N)
        self._AB = torch.zeros((B, N, NX, NX + NU), **f64)
        self._beq = torch.zeros((B, NX * (N + 1)), **f64)
        self._q = torch.zeros((B, n), **f64)
        self._cost = torch.zeros((B,), **f64)
        self._ok = torch.zeros((B,), dtype=torch.int32, device=x0.device)
        self._ws = me.AdmmWorkspace(B, N, device=x0.device)   # fresh solver object: x = y = 0, rho = default (:454-478)
        self._solver = self._ws
        self.last_kernel_launches = 0

    def step_batch(self, X, active=None) -> BatchRTISolution:
        """prepare + feedback (osqp_rti.py:480-599) for every scenario; X (B,7) numpy or CUDA tensor."""
        if self._solver is None:
            raise RuntimeError("Must call initialize() before step()")
        torch = ge.torch_()
        t0 = time.perf_counter()
        dev_in = hasattr(X, "is_cuda")
        x = ge.to_dev(X)
        act = None if active is None else ge.to_dev(active, torch.int32)
        # prepare (:494-496: q) + feedback (:514-517: A, l, u) in one assembly launch
        me.qp_assemble(self._dynp, self._qp, self._X_lin, self._U_lin, x, self._x_ref, self._AB, self._beq, self._q,
                       lin_mode=self._lin_mode, rhs_mode=0, active=act)
        if self.config.osqp_warm_start:  # :522-524 primal warm start only; duals / rho persist in the "solver object"
            me.pack_traj(self._X_prev, self._U_prev, self._ws.x, active=act)
        me.admm_solve(self._qp, self._AB, self._beq, self._q, self._settings(), self._ws, active=act)
        me.mpc_update(self._qp, 0, self._ws.x, self._ws.status, self._x_ref, self._X_lin, self._U_lin, self._X_prev,
                      self._U_prev, self._cost, self._ok, active=act, n_iter=None)
        self.last_kernel_launches = 4
        ok = self._ok.to(torch.bool)
        # :532-552: solved -> the new trajectory; otherwise the previous (shifted) solution, cost = inf
        X_opt = torch.where(ok[:, None, None], self._X_lin, self._X_prev)
        U_opt = torch.where(ok[:, None, None], self._U_lin, self._U_prev)
        cost = torch.where(ok, self._ws.obj, torch.full_like(self._ws.obj, float("inf")))
        out = (lambda t: t.clone()) if dev_in else (lambda t: t.cpu().numpy())
        sol = BatchRTISolution(u0=out(U_opt[:, 0]), X_opt=out(X_opt), U_opt=out(U_opt), cost=out(cost),
                               osqp_iterations=out(self._ws.iters), success=out(ok), status=out(self._ws.status),
                               prep_time_ms=0.0, feedback_time_ms=0.0, total_time_ms=0.0)
        sol.total_time_ms = sol.feedback_time_ms = (time.perf_counter() - t0) * 1000
        return sol

    def reset_scenarios(self, mask, X0) -> None:
        """Re-`initialize` the scenarios with mask != 0 from state X0[b] (fresh solver object: zero iterates, default
        rho, straight-line linearisation), leaving the others untouched."""
        torch = ge.torch_()
        mask = ge.to_dev(mask, torch.bool)
        x0 = ge.to_dev(X0)
        N = self.N
        xt = self._x_ref[:, 0, :]
        a = (torch.arange(N + 1, dtype=torch.float64, device=x0.device) / N)[None, :, None]
        Xl = (1 - a) * x0[:, None, :] + a * xt[:, None, :]
        g = getattr(getattr(self.dynamics, "params", None), "g", 1.0)
        Ul = torch.zeros_like(self._U_lin)
        Ul[:, :, 0] = (x0[:, 0] * g)[:, None]
        m3 = mask[:, None, None]
        self._X_lin = torch.where(m3, Xl, self._X_lin).contiguous()
        self._U_lin = torch.where(m3, Ul, self._U_lin).contiguous()
        self._X_prev = torch.where(m3, Xl, self._X_prev).contiguous()
        self._U_prev = torch.where(m3, Ul, self._U_prev).contiguous()
        self._ws.x.masked_fill_(mask[:, None], 0.0)
        self._ws.y.masked_fill_(mask[:, None], 0.0)
        self._ws.rho.masked_fill_(mask, 0.0)

    # ------------------------------------------------------------------------------------------------ reference API
    def initialize(self, x0, x_target, X_init=None, U_init=None) -> None:
        self._batched = False
        self.initialize_batch(np.asarray(x0, dtype=np.float64)[None, :], np.asarray(x_target, dtype=np.float64),
                              None if X_init is None else np.asarray(X_init, dtype=np.float64)[None],
                              None if U_init is None else np.asarray(U_init, dtype=np.float64)[None])

    def prepare(self) -> float:
        """:480-499: the cost vector q is rebuilt inside the fused assembly launch of `feedback`; nothing to do here."""
        return 0.0

    def feedback(self, x_current) -> OSQPRTISolution:
        return self.step_batch(np.asarray(x_current, dtype=np.float64)[None, :])[0]

    def step(self, x_current) -> OSQPRTISolution:
        t0 = time.perf_counter()
        prep = self.prepare()
        sol = self.feedback(x_current)
        sol.prep_time_ms = prep
        sol.total_time_ms = (time.perf_counter() - t0) * 1000
        return sol

    def _solution_to_vector(self, X, U):
        """:601-614."""
        N = self.N
        z = np.zeros(self.n_vars)
        for k in range(N):
            z[k * 10:k * 10 + NX] = X[k]
            z[k * 10 + NX:(k + 1) * 10] = U[k]
        z[N * 10:] = X[N]
        return z

    def _vector_to_solution(self, z) -> Tuple[np.ndarray, np.ndarray]:
        """:616-631."""
        N = self.N
        X, U = np.zeros((N + 1, NX)), np.zeros((N, NU))
        for k in range(N):
            X[k] = z[k * 10:k * 10 + NX]
            U[k] = z[k * 10 + NX:(k + 1) * 10]
        X[N] = z[N * 10:]
        return X, U

    def update_reference(self, x_target) -> None:
        """:633-635."""
        xt = ge.to_dev(x_target)
        xt = xt[None, :].expand(self._B, NX) if xt.dim() == 1 else xt
        self._x_ref = xt[:, None, :].expand(self._B, self.N + 1, NX).contiguous()

    def get_predicted_trajectory(self):
        """:637-639 (B = 1: (N+1,7),(N,3); batched: (B,N+1,7),(B,N,3))."""
        X, U = self._X_lin.cpu().numpy(), self._U_lin.cpu().numpy()
        return (X[0], U[0]) if self._B == 1 else (X, U)


class FastRTI3DoF(OSQPRTIMPC):
    """osqp_rti.py:642-710: analytic Jacobians."""

    _lin_mode = 0

    def __init__(self, dynamics, config: Optional[OSQPRTIConfig] = None):
        config = config or OSQPRTIConfig()
        config.n_x, config.n_u = 7, 3
        super().__init__(dynamics, config)

"""Predictive safety filter on the first control for the 3-DoF path -- the reference's `src.safety` surface that the Monte
Carlo loop touches (src/experiments/monte_carlo.py:519-525 `safety_filter.filter(x, u).u_safe`), batched on the device.

Reference: src/safety/safety_filter.py (`SafetyFilterConfig` :46-70, `SafetyFilterResult` :73-84, `PredictiveSafetyFilter`
`initialize` :146-178, `filter` :184-235, `_check_safety` :237-279, `_solve_gradient_qp` :483-531 -- the branch that runs
without CasADi), src/safety/backup_controller.py (`BackupControllerConfig` :36-53, `LQRBackupController.compute_gains`
:135-172 with its PD fall-back :211-229, `get_control` :231-252) and src/safety/invariant_sets.py (`EllipsoidalInvariantSet`
:52-184).  The reference classes are written for the 14-state model; here: 7 states [m, r, v], identity attitude (its tilt
test always passes).  The gains and the Lyapunov matrix are a handful of host flops (scipy DARE, as in the reference); every
per-scenario evaluation -- predictions with the backup law, the gradient iteration, the re-check -- runs in
`gpmpc_safety_filter_3dof`.  There is no CPU path."""
from __future__ import annotations

import time
from dataclasses import dataclass
from typing import Any, Optional

import numpy as np

from . import _lib
from .gp import engine as ge
from .mpc import engine as me

NX, NU = 7, 3


@dataclass
class SafetyFilterConfig:
    """safety_filter.py:46-70 (fields the 3-DoF filter reads)."""

    N: int = 10
    dt: float = 0.1
    backup_type: str = "ellipsoid"
    backup_margin: float = 0.9
    use_gp_uncertainty: bool = True
    gp_confidence: float = 0.95
    qp_max_iter: int = 50        # iterations of the gradient method (safety_filter.py:487 n_iters)
    qp_tol: float = 1e-6
    filter_mode: str = "soft"
    soft_weight: float = 1e3
    gradient_step: float = 0.1   # safety_filter.py:488 lr


@dataclass
class BackupControllerConfig:
    """backup_controller.py:36-53."""

    Q_diag: Optional[np.ndarray] = None
    R_diag: Optional[np.ndarray] = None
    u_min: Optional[np.ndarray] = None
    u_max: Optional[np.ndarray] = None
    dt: float = 0.1
    controller_type: str = "lqr"


@dataclass
class SafetyFilterResult:
    """safety_filter.py:73-84."""

    u_safe: Any
    u_nominal: Any
    is_modified: Any
    is_safe: Any
    safety_margin: Any
    solve_time: float
    backup_value: Any
    constraint_violations: Any = None


class LQRBackupController:
    """backup_controller.py:56-252 for 7 states: u = u_eq - K (x - x_eq)."""

    def __init__(self, dynamics, config: Optional[BackupControllerConfig] = None):
        self.dynamics = dynamics
        self.config = config or BackupControllerConfig()
        self.n_x, self.n_u = NX, NU
        self.x_eq = self.u_eq = self.K = self.P = None
        q = self.config.Q_diag if self.config.Q_diag is not None else np.array([0.0, 10, 10, 10, 5, 5, 5])  # :109-127
        r = self.config.R_diag if self.config.R_diag is not None else np.array([0.1, 0.1, 0.1])             # :133
        self.Q, self.R = np.diag(np.asarray(q, float)), np.diag(np.asarray(r, float))

    def compute_gains(self, x_eq, u_eq=None) -> None:
        """:135-172 -- discrete Riccati equation at the hover point; PD gains when it has no finite solution (:211-229)."""
        from scipy.linalg import solve_discrete_are
        self.x_eq = np.asarray(x_eq, dtype=np.float64).copy()
        p = self.dynamics.params
        g_vec = getattr(p, "g_I", None)
        if g_vec is None:
            g_vec = p.g_vec
        if u_eq is None:
            self.u_eq = np.array([self.x_eq[0] * abs(g_vec[0] * p.g0), 0.0, 0.0])
        else:
            self.u_eq = np.asarray(u_eq, dtype=np.float64).copy()
        A, B = self.dynamics.linearize(self.x_eq, self.u_eq, self.config.dt)
        try:
            self.P = solve_discrete_are(A, B, self.Q, self.R)
            BtP = B.T @ self.P
            self.K = np.linalg.solve(self.R + BtP @ B, BtP @ A)
        except Exception:  # noqa: BLE001  (the reference prints and falls back)
            self.K = np.zeros((NU, NX))
            for i in range(3):
                self.K[i, 1 + i] = 2.0
                self.K[i, 4 + i] = 1.0
            self.P = self.Q.copy()

    def get_control(self, x):
        """:231-252 (host helper for single states; the batch runs on the device inside the filter)."""
        u = self.u_eq - self.K @ (np.asarray(x, float) - self.x_eq)
        if self.config.u_min is not None:
            u = np.maximum(u, self.config.u_min)
        if self.config.u_max is not None:
            u = np.minimum(u, self.config.u_max)
        return u


class EllipsoidalInvariantSet:
    """invariant_sets.py:52-184: S = {x : (x - x_eq)' P (x - x_eq) <= alpha}."""

    def __init__(self, n_x: int = NX, alpha: float = 1.0):
        self.n_x, self.P, self.x_eq, self.alpha = n_x, None, None, float(alpha)

    def compute_from_lqr(self, P_lqr, x_eq, alpha: Optional[float] = None) -> None:
        self.P, self.x_eq = np.asarray(P_lqr, float).copy(), np.asarray(x_eq, float).copy()
        if alpha is not None:
            self.alpha = float(alpha)

    def get_value(self, x) -> float:
        dx = np.asarray(x, float) - self.x_eq
        return float(dx @ self.P @ dx)

    def contains(self, x) -> bool:
        return self.P is not None and self.get_value(x) <= self.alpha


class PredictiveSafetyFilter:
    """safety_filter.py:87-361 with the reference's constructor / `initialize` / `filter` signatures, plus `filter_batch`.

    constraint_params: any object with T_min, T_max, gamma_gs (degrees) -- e.g. `mpc.ConstraintParams`; None = the
    reference's defaults inside `_check_constraints` / `_solve_gradient_qp` (T in [0.5, 5], no glide-slope test)."""

    def __init__(self, dynamics, config: Optional[SafetyFilterConfig] = None, constraint_params=None,
                 backup_config: Optional[BackupControllerConfig] = None):
        self.dynamics = dynamics
        self.config = config or SafetyFilterConfig()
        self.constraint_params = constraint_params
        self.n_x, self.n_u = NX, NU
        self.backup = LQRBackupController(dynamics, backup_config or BackupControllerConfig(dt=self.config.dt))
        self.invariant_set = EllipsoidalInvariantSet(NX)
        self._initialized = False
        self._setup = None

    def initialize(self, x_eq, u_eq=None, backup_alpha: Optional[float] = None, K=None, P=None) -> None:
        """:146-178.  K, P: inject the gains instead of computing them (tests pin the reference's own)."""
        if K is not None:
            self.backup.x_eq = np.asarray(x_eq, float).copy()
            g = self.dynamics.params
            gv = getattr(g, "g_I", None)
            if gv is None:
                gv = g.g_vec
            self.backup.u_eq = np.asarray(u_eq, float).copy() if u_eq is not None else \
                np.array([self.backup.x_eq[0] * abs(gv[0] * g.g0), 0.0, 0.0])
            self.backup.K, self.backup.P = np.asarray(K, float).copy(), np.asarray(P, float).copy()
        else:
            self.backup.compute_gains(x_eq, u_eq)
        self.invariant_set.compute_from_lqr(self.backup.P, self.backup.x_eq, backup_alpha)
        c, cp = self.config, self.constraint_params
        p = self.dynamics.params
        g_vec = getattr(p, "g_vec", None)
        if g_vec is None:
            g_vec = p.g_I
        s = _lib.Safety3()
        s.dyn = me.make_dyn(p.alpha, np.asarray(g_vec, dtype=np.float64), c.dt)
        s.N, s.n_iters, s.lr = int(c.N), int(c.qp_max_iter), float(c.gradient_step)
        s.alpha, s.backup_margin = float(self.invariant_set.alpha), float(c.backup_margin)
        s.T_min = float(getattr(cp, "T_min", 0.5)) if cp is not None else 0.5
        s.T_max = float(getattr(cp, "T_max", 5.0)) if cp is not None else 5.0
        if cp is not None:
            gam = getattr(cp, "gamma_gs", 30.0)
            s.tan_gamma = float(np.tan(np.deg2rad(gam)))
        else:
            s.tan_gamma = float(np.tan(np.deg2rad(30.0)))
        bc = self.backup.config
        s.has_u_limits = 1 if (bc.u_min is not None or bc.u_max is not None) else 0
        lo = np.full(3, -1e300) if bc.u_min is None else np.asarray(bc.u_min, float)
        hi = np.full(3, 1e300) if bc.u_max is None else np.asarray(bc.u_max, float)
        for i in range(21):
            s.K[i] = float(self.backup.K.ravel()[i])
        for i in range(49):
            s.P[i] = float(self.backup.P.ravel()[i])
        for i in range(7):
            s.x_eq[i] = float(self.backup.x_eq[i])
        for i in range(3):
            s.u_eq[i], s.u_min[i], s.u_max[i] = float(self.backup.u_eq[i]), float(lo[i]), float(hi[i])
        self._setup = s
        self._initialized = True

    def filter_batch(self, X, U_nominal, active=None) -> SafetyFilterResult:
        """B states / nominal controls at once: numpy in -> numpy out, CUDA tensors in -> CUDA tensors out.
        U_nominal may be a strided (B, 3) view of a (B, N, 3) plan (u0 = U_opt[:, 0])."""
        if not self._initialized:
            raise RuntimeError("Must call initialize() first")
        if self.constraint_params is None:
            raise _lib.GpmpcError("the device filter needs constraint_params (T_min, T_max, gamma_gs)")
        torch = ge.torch_()
        t0 = time.perf_counter()
        dev_in = hasattr(X, "is_cuda")
        x = ge.to_dev(X).contiguous()
        u = ge.to_dev(U_nominal)
        B = x.shape[0]
        if u.dim() != 2 or u.stride(1) != 1 or (B > 1 and u.stride(0) < NU):
            u = u.contiguous()
        ldu = u.stride(0) if B > 1 else NU      # (a size-1 leading axis may carry any stride)
        out = torch.empty((B, NU), dtype=torch.float64, device=x.device)
        flags = torch.zeros((B,), dtype=torch.int32, device=x.device)
        margin = torch.empty((B,), dtype=torch.float64, device=x.device)
        bval = torch.empty((B,), dtype=torch.float64, device=x.device)
        act = None if active is None else ge.to_dev(active, torch.int32).contiguous()
        ge.call("gpmpc_safety_filter_3dof", _lib.C.byref(self._setup), B, ge.ptr(x), u.data_ptr(), ldu, ge.ptr(out),
                ge.ptr(flags), ge.ptr(margin), ge.ptr(bval), ge.ptr(act), ge.stream())
        if dev_in:
            return SafetyFilterResult(u_safe=out, u_nominal=u, is_modified=(flags & 1) != 0, is_safe=(flags & 2) != 0,
                                      safety_margin=margin, solve_time=time.perf_counter() - t0, backup_value=bval)
        fl = flags.cpu().numpy()
        return SafetyFilterResult(u_safe=out.cpu().numpy(), u_nominal=np.asarray(U_nominal, float), is_modified=(fl & 1) != 0,
                                  is_safe=(fl & 2) != 0, safety_margin=margin.cpu().numpy(),
                                  solve_time=time.perf_counter() - t0, backup_value=bval.cpu().numpy())

    def filter(self, x, u_nominal) -> SafetyFilterResult:
        """:184-235 for one state (a B = 1 view of the batched kernel)."""
        r = self.filter_batch(np.asarray(x, float)[None], np.asarray(u_nominal, float)[None])
        return SafetyFilterResult(u_safe=r.u_safe[0], u_nominal=np.asarray(u_nominal, float).copy(),
                                  is_modified=bool(r.is_modified[0]), is_safe=bool(r.is_safe[0]),
                                  safety_margin=float(r.safety_margin[0]), solve_time=r.solve_time,
                                  backup_value=float(r.backup_value[0]),
                                  constraint_violations={"terminal": float(r.safety_margin[0])})

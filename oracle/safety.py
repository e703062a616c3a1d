"""Oracle (test infrastructure): numpy restatement of the reference's predictive safety filter for the 7-state model.

Follows /root/reference/src/safety/safety_filter.py -- `filter` :184-235, `_check_safety` :237-279,
`_predict_with_backup` :281-313, `_check_constraints` :315-346 (the tilt test needs the attitude quaternion, which the
7-state model does not carry: identity attitude, always satisfied), `_solve_gradient_qp` :483-531 (the branch taken when
CasADi is absent) -- and backup_controller.py:231-252 (`get_control`), invariant_sets.py:178-184 (`get_value`).
Pinned by tests/golden/safety_filter.npz, written by the reference's own classes around a 14-state embedding."""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np

from .mpc import Rocket3DoFParams, step_3dof


@dataclass
class SafetySetup:
    K: np.ndarray            # (3,7) backup gain
    P: np.ndarray            # (7,7) Lyapunov matrix of the invariant set
    x_eq: np.ndarray
    u_eq: np.ndarray
    alpha: float
    N: int = 10
    dt: float = 0.1
    backup_margin: float = 0.9
    T_min: float = 0.3
    T_max: float = 5.0
    gamma_gs_deg: float = 30.0
    u_min: np.ndarray | None = None
    u_max: np.ndarray | None = None


def backup_control(s: SafetySetup, x):
    """backup_controller.py:231-252."""
    u = s.u_eq - s.K @ (x - s.x_eq)
    if s.u_min is not None:
        u = np.maximum(u, s.u_min)
    if s.u_max is not None:
        u = np.minimum(u, s.u_max)
    return u


def check_constraints(s: SafetySetup, x, u):
    """safety_filter.py:315-346 (thrust magnitude, glide slope; identity attitude)."""
    T = np.linalg.norm(u)
    if T < s.T_min or T > s.T_max:
        return False
    tan_gamma = np.tan(np.deg2rad(s.gamma_gs_deg))
    if x[1] > 0.1:
        if x[1] ** 2 * tan_gamma**2 < x[2] ** 2 + x[3] ** 2:
            return False
    return True


def predict_with_backup(s: SafetySetup, x, u_first, p=None):
    """safety_filter.py:281-313."""
    p = p or Rocket3DoFParams()
    X = np.zeros((s.N + 1, 7))
    X[0] = x
    X[1] = step_3dof(x, u_first, s.dt, p)
    for k in range(1, s.N):
        X[k + 1] = step_3dof(X[k], backup_control(s, X[k]), s.dt, p)
    return X


def value(s: SafetySetup, x):
    dx = x - s.x_eq
    return float(dx @ s.P @ dx)


def check_safety(s: SafetySetup, x, u):
    """safety_filter.py:237-279 -> (is_safe, safety_margin)."""
    ok = check_constraints(s, x, u)
    X = predict_with_backup(s, x, u)
    margin = s.alpha - value(s, X[-1])
    path_ok = True
    for k in range(s.N):
        uk = backup_control(s, X[k]) if k > 0 else u
        if not check_constraints(s, X[k], uk):
            path_ok = False
    return (ok and margin > 0 and path_ok), margin


def solve_gradient_qp(s: SafetySetup, x, u_nominal, n_iters=50, lr=0.1):
    """safety_filter.py:483-531."""
    u = np.array(u_nominal, dtype=float)
    for _ in range(n_iters):
        V = value(s, predict_with_backup(s, x, u)[-1])
        if V <= s.alpha * s.backup_margin:
            break
        eps = 1e-4
        grad = np.zeros(3)
        for i in range(3):
            up = u.copy(); up[i] += eps
            grad[i] = (value(s, predict_with_backup(s, x, up)[-1]) - V) / eps
        u = u - lr * (grad + 2 * (u - u_nominal))
        T = np.linalg.norm(u)
        if T < s.T_min:
            u = u / max(T, 1e-6) * s.T_min
        elif T > s.T_max:
            u = u / T * s.T_max
    return u


def safety_filter(s: SafetySetup, x, u_nominal):
    """safety_filter.py:184-235 -> (u_safe, is_modified, is_safe, safety_margin, backup_value)."""
    safe, margin = check_safety(s, x, u_nominal)
    if safe:
        return np.array(u_nominal, dtype=float), False, True, margin, value(s, predict_with_backup(s, x, u_nominal)[-1])
    u = solve_gradient_qp(s, x, u_nominal)
    safe2, margin2 = check_safety(s, x, u)
    return u, True, safe2, margin2, value(s, predict_with_backup(s, x, u)[-1])

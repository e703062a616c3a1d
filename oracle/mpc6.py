"""TEST INFRASTRUCTURE ONLY (imported by tests/, __graft_entry__.smoke() and bench.py's CPU legs -- never by the product).

CPU restatement (numpy) of the 6-DoF quaternion GP-MPC step of BASELINE configs[3] (14 states
[m, r_I(3), v_I(3), q_BI(4), w_B(3)], 3 controls, N = 30, sparse structured GP):

* continuous model and its discretisation        reference src/mpc/nominal_mpc.py:163-236 (CasADi statement of the model;
                                                 "euler" and the reference's "rk4" = 4 Euler sub-steps)
* quaternion renormalisation after the step      reference src/dynamics/rocket_6dof.py:370-387
* forward-difference linearisation, eps = 1e-6   reference src/mpc/osqp_rti.py:374-401
* cost / bounds for n_x != 7                     reference src/mpc/osqp_rti.py:168-201 (fallback branch: ones, position 10,
                                                 velocity 1; the bound arrays have 7 entries, so states 7..13 are unbounded)
* constraint rows                                reference src/mpc/osqp_rti.py:260-372 (same row order as the 3-DoF QP)
* GP-augmented prediction, 14-state              reference src/mpc/gp_mpc.py:139-175
* affine GP term of the SQP subproblem           reference src/mpc/gp_mpc.py:300-317,402-411

PARITY UNPINNED for the dynamics: `simdyn` (the reference's Rocket6DoFDynamics backend) and CasADi are absent, and the
reference has no test vectors; the in-repo CasADi statement cited above is what both this file and csrc/mpc6.cu
implement.  The GP side (features, sparse FITC) is pinned by tests/golden/{features,sparse_gp,structured_gp}.npz.
"""
from __future__ import annotations

import numpy as np

NX, NU, NS = 14, 3, 17
X_MIN = np.concatenate([[-np.inf, -100, -100, -100, -50, -50, -50], np.full(7, -np.inf)])  # osqp_rti.py:197-198 ([:n_x] of 7)
X_MAX = np.concatenate([[np.inf, 500, 100, 100, 50, 50, 50], np.full(7, np.inf)])
U_MIN = np.array([0.3, -5.0, -5.0])
U_MAX = np.array([5.0, 5.0, 5.0])
Q_DIAG = np.ones(NX); Q_DIAG[1:4] = 10.0; Q_DIAG[4:7] = 1.0    # osqp_rti.py:172-175
QF_DIAG = Q_DIAG * 10.0
R_DIAG = np.ones(NU) * 0.01


class Rocket6DoFParams:
    """rocket_6dof.py:35-89 defaults; alpha = 1 / (I_sp g0)."""

    def __init__(self):
        self.alpha = 1.0 / (30.0 * 1.0)
        self.g = np.array([-1.0, 0.0, 0.0])
        self.J = np.diag([0.02, 1.0, 1.0]) * 0.168
        self.Jinv = np.linalg.inv(self.J)
        self.rT = np.array([-0.25, 0.0, 0.0])


def _cross(a, b):
    return np.array([a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]])


def f6(x, u, p: Rocket6DoFParams):
    """nominal_mpc.py:163-203."""
    m, v, q, w = x[0], x[4:7], x[7:11], x[11:14]
    qw, qx, qy, qz = q
    C = np.array([[1 - 2 * (qy * qy + qz * qz), 2 * (qx * qy - qw * qz), 2 * (qx * qz + qw * qy)],
                  [2 * (qx * qy + qw * qz), 1 - 2 * (qx * qx + qz * qz), 2 * (qy * qz - qw * qx)],
                  [2 * (qx * qz - qw * qy), 2 * (qy * qz + qw * qx), 1 - 2 * (qx * qx + qy * qy)]])
    T_I = np.array([C[r, 0] * u[0] + C[r, 1] * u[1] + C[r, 2] * u[2] for r in range(3)])
    Tm = np.sqrt(u[0] * u[0] + u[1] * u[1] + u[2] * u[2])
    xd = np.zeros(NX)
    xd[0] = -p.alpha * Tm
    xd[1:4] = v
    xd[4:7] = T_I / m + p.g
    qv = q[1:4]
    xd[7] = 0.5 * (-(w[0] * qv[0] + w[1] * qv[1] + w[2] * qv[2]))
    xd[8:11] = 0.5 * (q[0] * w + _cross(w, qv))
    Jw = np.array([p.J[r, 0] * w[0] + p.J[r, 1] * w[1] + p.J[r, 2] * w[2] for r in range(3)])
    rhs = _cross(p.rT, u) - _cross(w, Jw)
    xd[11:14] = np.array([p.Jinv[r, 0] * rhs[0] + p.Jinv[r, 1] * rhs[1] + p.Jinv[r, 2] * rhs[2] for r in range(3)])
    return xd


def step6(x, u, dt, p: Rocket6DoFParams, substeps=1, normalize=True):
    """Euler (substeps = 1) or the reference's "rk4" (= 4 Euler sub-steps, nominal_mpc.py:205-236), then
    quaternion normalisation (rocket_6dof.py:370-387)."""
    x = np.asarray(x, float).copy()
    h = dt / substeps
    for _ in range(substeps):
        x = x + h * f6(x, u, p)
    if normalize:
        q = x[7:11]
        x[7:11] = q / np.sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3])
    return x


def linearize_fd6(x, u, dt, p, eps=1e-6, **kw):
    """osqp_rti.py:374-401."""
    A = np.zeros((NX, NX)); B = np.zeros((NX, NU))
    f0 = step6(x, u, dt, p, **kw)
    for i in range(NX):
        xp = np.array(x, float); xp[i] += eps
        A[:, i] = (step6(xp, u, dt, p, **kw) - f0) / eps
    for i in range(NU):
        up = np.array(u, float); up[i] += eps
        B[:, i] = (step6(x, up, dt, p, **kw) - f0) / eps
    return A, B


def qp_dims(N):
    n = N * NS + NX
    return n, NX * (N + 1) + n


def build_cost(N, x_ref):
    """osqp_rti.py:203-258 with the n_x = 14 weights."""
    n, _ = qp_dims(N)
    Pd = np.zeros(n); q = np.zeros(n)
    for k in range(N):
        Pd[k * NS:k * NS + NX] = Q_DIAG; Pd[k * NS + NX:(k + 1) * NS] = R_DIAG
        q[k * NS:k * NS + NX] = -(Q_DIAG * x_ref[k])
    Pd[N * NS:] = QF_DIAG
    q[N * NS:] = -(QF_DIAG * x_ref[N])
    return Pd, q


def stage_blocks(N, X_lin, U_lin, dt, p, gp_affine=None, **kw):
    """Per-stage [A_k | B_k] (entries <= 1e-10 dropped, osqp_rti.py:301,309) and the right-hand sides of the dynamics
    rows in the GP-MPC convention (gp_mpc.py:402-411): A x + B u - x+ = -(X_lin[k+1] - A X_lin[k] - B U_lin[k] + c_gp)."""
    AB = np.zeros((N, NX, NS)); rhs = np.zeros((N, NX))
    for k in range(N):
        A, B = linearize_fd6(X_lin[k], U_lin[k], dt, p, **kw)
        c = X_lin[k + 1].copy()
        for r in range(NX):
            ax = 0.0
            for j in range(NX):
                ax += A[r, j] * X_lin[k][j]
            bu = 0.0
            for j in range(NU):
                bu += B[r, j] * U_lin[k][j]
            c[r] = (c[r] - ax) - bu
        if gp_affine is not None:
            c = c + gp_affine[k]
        AB[k, :, :NX] = np.where(np.abs(A) > 1e-10, A, 0.0)
        AB[k, :, NX:] = np.where(np.abs(B) > 1e-10, B, 0.0)
        rhs[k] = -c
    return AB, rhs


def dense_constraints(N, AB, x_init, rhs):
    """osqp_rti.py:260-372 row order: [x_0 = x_init | A_k x_k + B_k u_k - x_{k+1} = rhs_k | l <= I z <= u]."""
    n, m = qp_dims(N)
    A = np.zeros((m, n)); l = np.zeros(m); u = np.zeros(m)
    A[:NX, :NX] = np.eye(NX)
    l[:NX] = u[:NX] = x_init
    for k in range(N):
        r = NX + k * NX
        A[r:r + NX, k * NS:(k + 1) * NS] = AB[k]
        A[r:r + NX, (k + 1) * NS:(k + 1) * NS + NX] = -np.eye(NX)
        l[r:r + NX] = u[r:r + NX] = rhs[k]
    neq = NX * (N + 1)
    A[neq:, :] = np.eye(n)
    lo = np.concatenate([np.tile(np.concatenate([X_MIN, U_MIN]), N), X_MIN])
    hi = np.concatenate([np.tile(np.concatenate([X_MAX, U_MAX]), N), X_MAX])
    l[neq:] = lo; u[neq:] = hi
    return A, l, u


def rollout_gp(x0, U, dt, p, predict, **kw):
    """gp_mpc.py:277-281 with _predict_with_gp (:139-175): X[k+1] = step(X[k], U[k]) + dt [0.., d_v, 0.., d_w].
    predict(x, u) -> (d_v (3,), d_w (3,)) or None for the nominal model."""
    N = U.shape[0]
    X = np.zeros((N + 1, NX)); D = np.zeros((N, 6))
    X[0] = x0
    for k in range(N):
        xn = step6(X[k], U[k], dt, p, **kw)
        if predict is not None:
            dv, dw = predict(X[k], U[k])
            D[k, :3], D[k, 3:] = dv, dw
            xn[4:7] = xn[4:7] + dv * dt
            xn[11:14] = xn[11:14] + dw * dt
        X[k + 1] = xn
    return X, D


def propagate_linear6(x0, U, dt, p, predict_full, sigma0=1e-6, **kw):
    """uncertainty_prop.py:117-177 (`_propagate_linear`, the method GPMPC.solve uses: gp_mpc.py:124-127,283-289):
    Sigma_{k+1} = A_k Sigma_k A_k' + Q_k, Q_k = dt^2 diag(var_v) / diag(var_omega) in rows 4:7 / 11:14, along the GP-mean
    chain from x0.  A_k: forward differences of the step (the reference calls dynamics.linearize of the absent simdyn
    backend; the in-repo Jacobian statement is osqp_rti.py:374-401).  predict_full(x, u) -> d_v, d_w, var_v, var_w."""
    N = U.shape[0]
    means = np.zeros((N + 1, NX)); covs = np.zeros((N + 1, NX, NX))
    means[0] = x0
    covs[0] = np.eye(NX) * sigma0
    x = np.array(x0, float); S = covs[0].copy()
    for k in range(N):
        A, _ = linearize_fd6(x, U[k], dt, p, **kw)
        dv, dw, vv, vw = predict_full(x, U[k])
        Q = np.zeros((NX, NX))
        Q[4:7, 4:7] = np.diag(vv) * dt**2
        Q[11:14, 11:14] = np.diag(vw) * dt**2
        xn = step6(x, U[k], dt, p, **kw)
        xn[4:7] += dv * dt
        xn[11:14] += dw * dt
        S = A @ S @ A.T + Q
        means[k + 1] = xn; covs[k + 1] = S
        x = xn
    return means, covs


def hover_guess(x0, N, p):
    """Warm start used by the batched driver: thrust balancing gravity along the body axis."""
    U = np.zeros((N, NU)); U[:, 0] = x0[0] * 1.0
    return U

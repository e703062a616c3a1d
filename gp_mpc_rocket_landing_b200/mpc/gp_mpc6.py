"""6-DoF quaternion GP-MPC (BASELINE configs[3]: 14 states, N = 30, sparse structured GP, 1024 scenarios).

Same per-control-step algorithm as `GPMPC.solve` (reference src/mpc/gp_mpc.py:229-369) on the 14-state model:

1. roll the horizon out with the structured GP means (`_predict_with_gp`, gp_mpc.py:139-175,277-281)  -> gpmpc_rollout_6dof
2. SQP loop (gp_mpc.py:296-345): forward-difference Jacobians of `dynamics.step` (osqp_rti.py:374-401), affine GP term
   (:309-317), QP in the OSQP-RTI layout with the n_x != 7 cost branch (osqp_rti.py:172-175)       -> gpmpc_qp_assemble_6dof
   batched ADMM (replaces the IPOPT call of `_solve_qp`, as in the 3-DoF path)                     -> gpmpc_admm_solve(14, 3)
   unpack / step-size test max|dz| < 1e-4 / cost                                                    -> gpmpc_mpc_update
3. warm start for the next call (gp_mpc.py:355-357).

The reference's 6-DoF controllers need CasADi/IPOPT and `simdyn` (absent): parity is against oracle/mpc6.py, the CPU
restatement of the in-repo model statement; see that file's header for what is pinned and what is not.
There is no CPU fallback: everything below runs through libgpmpc_b200.so on a CUDA device.
"""
from __future__ import annotations

import ctypes as C
import time
from dataclasses import dataclass
from typing import Optional

import numpy as np

from .. import _lib
from ..gp import engine as ge
from . import engine as me
from .nominal_mpc import BatchMPCSolution, MPCSolution

NX6, NU6, NS6 = 14, 3, 17
INF = 1e30
# osqp_rti.py:172-175 (n_x != 7 branch) and :197-200 (the bound arrays have 7 entries: states 7..13 are unbounded)
Q6 = np.ones(NX6); Q6[1:4] = 10.0; Q6[4:7] = 1.0
QF6 = Q6 * 10.0
R6 = np.ones(NU6) * 0.01
X_MIN6 = np.concatenate([[-np.inf, -100.0, -100.0, -100.0, -50.0, -50.0, -50.0], np.full(7, -np.inf)])
X_MAX6 = np.concatenate([[np.inf, 500.0, 100.0, 100.0, 50.0, 50.0, 50.0], np.full(7, np.inf)])
U_MIN6 = np.array([0.3, -5.0, -5.0])
U_MAX6 = np.array([5.0, 5.0, 5.0])


@dataclass
class GPMPC6DoFConfig:
    """gp_mpc.py:48-63 + nominal_mpc.py:42-64 fields that the numeric core reads, then the solver knobs."""

    N: int = 30
    dt: float = 0.1
    use_gp_mean: bool = True
    use_gp_uncertainty: bool = True   # gp_mpc.py:54: propagate the GP variance through the horizon (linear method)
    repropagate_uncertainty: bool = True  # gp_mpc.py:347-353 (computed and discarded by the reference)
    warm_start: bool = True
    max_sqp_iter: int = 10            # gp_mpc.py:296
    qp_max_iter: int = 1000
    qp_eps_abs: float = 1e-4
    qp_eps_rel: float = 1e-4
    qp_check_every: int = 25
    qp_adaptive_rho_interval: int = 25
    qp_scaling: int = 3
    qp_keep_rho: bool = True
    qp_kernel_variant: int = 0        # GpmpcAdmmSettings.kernel_variant (0 = newest kernel that fits)


def _stage_vector(N, xpart, upart):
    v = np.zeros(N * NS6 + NX6)
    for k in range(N):
        v[k * NS6:k * NS6 + NX6] = xpart
        v[k * NS6 + NX6:(k + 1) * NS6] = upart
    v[N * NS6:] = xpart
    return v


class _State6:
    def __init__(self, B, N, device="cuda"):
        torch = ge.torch_()
        f64 = dict(dtype=torch.float64, device=device)
        i32 = dict(dtype=torch.int32, device=device)
        n, m = me.qp_dims(N, NX6, NU6)
        self.B, self.N = B, N
        self.X = torch.zeros((B, N + 1, NX6), **f64)
        self.U = torch.zeros((B, N, NU6), **f64)
        self.U_warm = torch.zeros((B, N, NU6), **f64)
        self.dmean = torch.zeros((B, N, 6), **f64)
        self.AB = torch.zeros((B, N, NX6, NS6), **f64)
        self.beq = torch.zeros((B, NX6 * (N + 1)), **f64)
        self.q = torch.zeros((B, n), **f64)
        self.ws = me.AdmmWorkspace(B, N, nx=NX6, nu=NU6)
        self.cost = torch.zeros((B,), **f64)
        self.active = torch.ones((B,), **i32)
        self.converged = torch.zeros((B,), **i32)
        self.n_iter = torch.zeros((B,), **i32)
        self.qp_iters = torch.zeros((B,), **i32)
        self.has_warm = False
        # uncertainty propagation (uncertainty_prop.py:117-177): its own rollout, Jacobians, process noise and chain
        self.Xr = torch.zeros((B, N + 1, NX6), **f64)
        self.dmean_r = torch.zeros((B, N, 6), **f64)
        self.ABr = torch.zeros((B, N, NX6, NS6), **f64)
        self.qdiag = torch.zeros((B, N, NX6), **f64)
        self.covs = torch.zeros((B, N + 1, NX6, NX6), **f64)
        self.has_cov = False


class GPMPC6DoF:
    """`GPMPC(dynamics, gp_model, config)` for the 14-state model: `.solve(x0, x_target) -> MPCSolution` and the batched
    entry point `.solve_batch(X0 (B,14), x_target (14,) | (B,14)) -> BatchMPCSolution` (B independent controllers, each
    with its own warm start).  `gp_model`: a `StructuredRocketGP` (or None for the nominal controller)."""

    def __init__(self, dynamics, gp_model=None, config: Optional[GPMPC6DoFConfig] = None):
        self.dynamics = dynamics
        self.gp = gp_model
        self.config = config or GPMPC6DoFConfig()
        N = self.config.N
        P = _stage_vector(N, Q6, R6)
        P[N * NS6:] = QF6
        lb = np.maximum(_stage_vector(N, X_MIN6, U_MIN6), -INF)
        ub = np.minimum(_stage_vector(N, X_MAX6, U_MAX6), INF)
        self._qp = me.QPData(N, ge.to_dev(P), ge.to_dev(lb), ge.to_dev(ub), ge.to_dev(Q6), ge.to_dev(QF6), ge.to_dev(R6))
        self._dyn = dynamics.dyn_struct(self.config.dt)
        self._batch = None
        self._single = None
        self._models = (None, None)
        self._is_setup = False
        self.timers: Optional[me.PhaseTimers] = None   # set by the bench tools: CUDA-event phase attribution

    def _span(self, name):
        return self.timers.span(name) if self.timers is not None else me.NULL_SPAN

    def setup(self) -> None:
        """Fetch the device models of the structured GP (fits it lazily, structured_gp.py:245-253)."""
        if self.gp is not None and self.config.use_gp_mean:
            self._models = self.gp.device_mean_models()
        self._is_setup = True

    def reset_warm_start(self) -> None:
        self._batch = None
        self._single = None

    def _settings(self):
        c = self.config
        return _lib.AdmmSettings.reference_defaults(
            max_iter=c.qp_max_iter, eps_abs=c.qp_eps_abs, eps_rel=c.qp_eps_rel, check_every=c.qp_check_every,
            adaptive_rho_interval=c.qp_adaptive_rho_interval, scaling=c.qp_scaling,
            warm_x=1, warm_y=1, rho_from_input=1 if c.qp_keep_rho else 0, commit_y_solved_only=1,
            kernel_variant=int(c.qp_kernel_variant))

    def _rollout(self, st, x0, U, X, dmean):
        mv, mw = self._models
        ge.call("gpmpc_rollout_6dof", C.byref(mv) if mv is not None else None, C.byref(mw) if mw is not None else None,
                C.byref(self._dyn), st.B, st.N, ge.ptr(x0), ge.ptr(U), ge.ptr(X), ge.ptr(dmean), ge.ptr(st.active),
                ge.stream())

    def _propagate(self, st: _State6, Xr):
        """Sigma_{k+1} = A_k Sigma_k A_k' + Q_k along the GP-mean rollout Xr of the current controls
        (uncertainty_prop.py:117-177, Sigma_0 = 1e-6 I as at gp_mpc.py:285-289): Jacobians of `dynamics.step` at
        (Xr[k], U[k]), Q_k = dt^2 diag(var_v) / diag(var_omega) in the velocity / body-rate rows."""
        B, N = st.B, st.N
        dt = float(self.config.dt)
        Xp = Xr[:, :N].reshape(B * N, NX6).contiguous()
        Up = st.U.reshape(B * N, NU6).contiguous()
        _, _, vv, vw = self.gp.predict_batch(Xp, Up)
        st.qdiag[:, :, 4:7] = vv.reshape(B, N, 3) * (dt * dt)
        st.qdiag[:, :, 11:14] = vw.reshape(B, N, 3) * (dt * dt)
        ge.call("gpmpc_linearize_6dof", C.byref(self._dyn), B * N, ge.ptr(Xp), ge.ptr(Up), ge.ptr(st.ABr), ge.stream())
        ge.call("gpmpc_propagate_cov", NX6, B, N, ge.ptr(st.ABr), NS6, NX6 * NS6, ge.ptr(st.qdiag), 1e-6,
                ge.ptr(st.covs), ge.ptr(st.active), ge.stream())
        st.has_cov = True

    def _solve_device(self, st: _State6, x0, x_ref):
        c = self.config
        N = c.N
        use_mean = c.use_gp_mean and (self._models[0] is not None or self._models[1] is not None)
        use_unc = c.use_gp_uncertainty and self.gp is not None and self.gp.n_data > 0
        if st.has_warm and c.warm_start:
            st.U.copy_(st.U_warm)
        else:  # hover along the body axis (gp_mpc.py:266-275)
            st.U.zero_()
            st.U[:, :, 0] = (x0[:, 0] * float(self.dynamics.params.g0))[:, None]
        st.active.fill_(1)
        st.n_iter.zero_()
        st.qp_iters.zero_()
        st.converged.zero_()
        with self._span("rollout"):
            self._rollout(st, x0, st.U, st.X, st.dmean if use_mean else None)
        st.has_cov = False
        if use_unc:
            with self._span("uncertainty"):
                self._propagate(st, st.X)
        S = self._settings()
        ldx = 0 if x_ref.dim() == 2 else x_ref.stride(0)
        for it in range(c.max_sqp_iter):
            if use_mean and it > 0:  # GP means at the new linearisation points (gp_mpc.py:309-314)
                with self._span("gp_mean"):
                    self._gp_points(st)
            sp = self._span("assemble"); sp.__enter__()
            ge.call("gpmpc_qp_assemble_6dof", C.byref(self._dyn), st.B, N, ge.ptr(st.X), ge.ptr(st.U), ge.ptr(x0),
                    ge.ptr(x_ref), ldx, ge.ptr(st.dmean) if use_mean else None, ge.ptr(self._qp.Qd), ge.ptr(self._qp.Qfd),
                    ge.ptr(st.AB), ge.ptr(st.beq), ge.ptr(st.q), ge.ptr(st.active), ge.stream())
            ge.call("gpmpc_pack_traj", NX6, NU6, st.B, N, ge.ptr(st.X), ge.ptr(st.U), ge.ptr(st.ws.x), ge.stream())
            sp.__exit__(None, None, None)
            with self._span("admm"):
                me.admm_solve(self._qp, st.AB, st.beq, st.q, S, st.ws, active=st.active)
            sp = self._span("update"); sp.__enter__()
            st.qp_iters += st.ws.iters * st.active
            ge.call("gpmpc_mpc_update", NX6, NU6, st.B, N, 1, ge.ptr(st.ws.x), ge.ptr(st.ws.status), ge.ptr(x_ref), ldx,
                    ge.ptr(self._qp.Qd), ge.ptr(self._qp.Qfd), ge.ptr(self._qp.Rd), ge.ptr(st.X), ge.ptr(st.U), None, None,
                    ge.ptr(st.cost), ge.ptr(st.converged), ge.ptr(st.active), ge.ptr(st.n_iter), ge.stream())
            sp.__exit__(None, None, None)
            if it == c.max_sqp_iter - 1 or int(st.active.sum().item()) == 0:
                break
            if use_unc and c.repropagate_uncertainty:  # gp_mpc.py:347-353: chain along the rollout of the new controls
                with self._span("uncertainty"):
                    self._rollout(st, x0, st.U, st.Xr, st.dmean_r if use_mean else None)
                    self._propagate(st, st.Xr)
        st.U_warm.copy_(st.U)
        st.has_warm = True

    def _gp_points(self, st):
        """d_v, d_omega at the trajectory points (X[b,k], U[b,k]), k < N, through the structured GP's batched predict."""
        B, N = st.B, st.N
        Xp = st.X[:, :N].reshape(B * N, NX6).contiguous()
        Up = st.U.reshape(B * N, NU6).contiguous()
        dv, dw, _, _ = self.gp.predict_batch(Xp, Up)
        st.dmean[:, :, :3] = dv.reshape(B, N, 3)
        st.dmean[:, :, 3:] = dw.reshape(B, N, 3)

    def last_covariances(self, batch: bool = True):
        """Propagated state covariances (B, N+1, 14, 14) of the last `solve_batch` (or `solve`), numpy; None when the
        controller has no GP or `use_gp_uncertainty` is off (cf. get_uncertainty_at_horizon, gp_mpc.py:486-492)."""
        st = self._batch if batch else self._single
        if st is None or not st.has_cov:
            return None
        return st.covs.cpu().numpy()

    def _state_for(self, B, single):
        attr = "_single" if single else "_batch"
        st = getattr(self, attr)
        if st is None or st.B != B or st.N != self.config.N:
            st = _State6(B, self.config.N)
            setattr(self, attr, st)
        return st

    def solve_batch(self, X0, x_target, _single=False) -> BatchMPCSolution:
        if not self._is_setup:
            self.setup()
        torch = ge.torch_()
        t0 = time.perf_counter()
        dev_in = hasattr(X0, "is_cuda")
        x0 = ge.to_dev(X0).contiguous()
        B, N = x0.shape[0], self.config.N
        xt = ge.to_dev(x_target)
        if xt.dim() == 1:
            x_ref = xt[None, :].expand(N + 1, NX6).contiguous()
        else:
            x_ref = xt[:, None, :].expand(B, N + 1, NX6).contiguous()
        st = self._state_for(B, _single)
        self._solve_device(st, x0, x_ref)
        conv = st.converged.bool()
        sol = dict(u0=st.U[:, 0].clone(), X_opt=st.X.clone(), U_opt=st.U.clone(), cost=st.cost.clone(),
                   iterations=st.n_iter.clone(), success=conv, status=conv.to(torch.int8), qp_status=st.ws.status.clone(),
                   qp_iterations=st.qp_iters.clone())
        if not dev_in:
            sol = {k: v.cpu().numpy() for k, v in sol.items()}
        else:
            torch.cuda.synchronize()
        return BatchMPCSolution(solve_time=time.perf_counter() - t0, **sol)

    def solve(self, x0, x_target, X_ref=None, U_ref=None) -> MPCSolution:
        b = self.solve_batch(np.asarray(x0, dtype=np.float64)[None], np.asarray(x_target, dtype=np.float64), _single=True)
        return b[0]

"""GP-MPC with the reference's API (src/mpc/gp_mpc.py: `GPMPCConfig`, `GPMPC.setup/solve/reset_warm_start`) plus the
new `solve_batch` entry point: B independent controller instances advanced by one fused sequence of device kernels.

Algorithm = `GPMPC.solve` (gp_mpc.py:229-369) specialised to the 3-DoF 7-state model:
  1. GP-augmented rollout of the warm-start controls            (:266-281, `_predict_with_gp` :139-175)
  2. linear covariance propagation with GP variance             (:284-289 -> uncertainty_prop.py:117-177)
  3. up to `max_sqp_iter` (reference: 10) iterations of  linearise + GP-mean affine term + QP  (:296-353),
     convergence max|dX|,max|dU| < 1e-4 (:337-345), failure -> keep the nominal trajectory, cost = inf (:478-482)
  4. warm start = the last iterate, unshifted                   (:355-357)
The inner IPOPT "QP" of the reference (CasADi, absent, non-convex thrust annulus) is replaced by the OSQP-form QP of
osqp_rti.py:203-372 (box bounds, diagonal cost) solved by the batched ADMM kernel -- north_star's design; numerical
parity of u0 with IPOPT is not a target.  Everything numeric runs in libgpmpc_b200.so; there is no CPU path.
"""
from __future__ import annotations

import time
from dataclasses import dataclass
from typing import Optional

import numpy as np

from .. import _lib
from ..gp import engine as ge
from . import engine as me
from .constraints import ConstraintParams, TightenedConstraints
from .cost_functions import CostWeights
from .nominal_mpc import BatchMPCSolution, DeferredHostArray, MPCConfig, MPCSolution

NX, NU = 7, 3


@dataclass
class GPMPCConfig(MPCConfig):
    """gp_mpc.py:48-63 fields, then the knobs the reference hard-codes or leaves to OSQP defaults."""

    use_gp_mean: bool = True
    use_gp_uncertainty: bool = True
    confidence_level: float = 0.95
    max_variance_for_constraint: float = 1.0
    robust_horizon: int = -1
    # --- additions (reference values as defaults) ---
    max_sqp_iter: int = 10            # gp_mpc.py:293 `max_iters = 10`
    sqp_tol: float = 1e-4             # gp_mpc.py:343 (the device kernel implements exactly this constant)
    repropagate_uncertainty: bool = True  # gp_mpc.py:348-353 (computed and discarded by the reference)
    confidence_level: float = 0.95        # gp_mpc.py:57
    max_variance_for_constraint: float = 1.0  # gp_mpc.py:60: no tightening for a stage more uncertain than this
    # Extension (off = reference behaviour, which prepares the tightening and applies it at k = 0 only, gp_mpc.py:414):
    # tighten the state box of every stage k >= 1 of every scenario by kappa * sqrt(diag Sigma_k), kappa = Phi^-1(level)
    tighten_state_bounds: bool = False
    # numpy-in calls: copy u0 / cost / flags to the host at once and X_opt / U_opt on first access (BatchMPCSolution)
    defer_trajectory_copy: bool = True
    qp_max_iter: int = 1000           # config/mpc_params.yaml:48 (osqp_rti.py:53 uses 50, which rarely converges)
    qp_eps_abs: float = 1e-4          # osqp_rti.py:54
    qp_eps_rel: float = 1e-4          # osqp_rti.py:55
    qp_check_every: int = 25          # OSQP default check_termination
    qp_adaptive_rho_interval: int = 25  # iteration based (OSQP's default is timing based, i.e. not reproducible)
    qp_scaling: int = 3               # osqp_rti.py:477
    qp_keep_rho: bool = True          # the adapted rho persists across solves, as inside one OSQP solver object
    variance_via_inverse: bool = True  # horizon variances |L^-1 k*|^2 as one triangular DMMA GEMM against the explicit
                                       # inverse factor (formed once per fit) instead of the blocked in-place solve
                                       # (31 dependent launches); same quantity, rounding differs by O(cond(L) eps)
    overlap_uncertainty: bool = False  # GP variance + covariance chain on a side stream, concurrent with the QP (the QP
                                       # does not read them: gp_mpc.py:414 evaluates the tightening at k = 0 only).  The
                                       # variance GEMM then runs as a persistent one-CTA-per-SM grid so that two ADMM QPs
                                       # fit next to it on every SM.  Measured on B200: the kernels do
                                       # co-reside, but the step stays at 8.83 ms (8.89 ms serial) -- they contend for the
                                       # same shared-memory / issue bandwidth.  Off by default: clean per-kernel timings.


def _is_dev(a):
    return hasattr(a, "is_cuda")


class _BatchState:
    """Device-resident state of B controller instances (buffers are allocated once and reused every step)."""

    def __init__(self, B: int, N: int, model, device="cuda"):
        torch = ge.torch_()
        f64 = dict(dtype=torch.float64, device=device)
        i32 = dict(dtype=torch.int32, device=device)
        n, m = me.qp_dims(N)
        self.B, self.N = B, N
        self.X = torch.zeros((B, N + 1, NX), **f64)
        self.U = torch.zeros((B, N, NU), **f64)
        self.Xs = torch.zeros((B, N + 1, NX), **f64)       # scratch rollout of the re-propagation
        self.X_unc = torch.zeros((B, N + 1, NX), **f64)    # snapshot read by the side-stream covariance chain
        self.U_unc = torch.zeros((B, N, NU), **f64)
        self.active_unc = torch.ones((B,), **i32)
        self.unc_event = None
        self.dmean = torch.zeros((B, N, 3), **f64)
        self.dmean_s = torch.zeros((B, N, 3), **f64)
        self.q_pad = ge.pad_to(B * N, 128)
        self.dvar = torch.zeros((self.q_pad, 3), **f64)
        self.Ks = torch.zeros((self.q_pad, model.f.n_pad), **f64) if model is not None else None
        self.var_work = None  # scratch of the explicit-inverse variance path (allocated on first use)
        self.covs = torch.zeros((B, N + 1, NX, NX), **f64)
        self.lb_t = None   # per-scenario tightened boxes (tighten_state_bounds)
        self._pinned_in = {}
        self.ub_t = None
        self.AB = torch.zeros((B, N, NX, NX + NU), **f64)
        self.beq = torch.zeros((B, NX * (N + 1)), **f64)
        self.q = torch.zeros((B, n), **f64)
        self.cost = torch.zeros((B,), **f64)
        self.converged = torch.zeros((B,), **i32)
        self.active = torch.ones((B,), **i32)
        self.n_iter = torch.zeros((B,), **i32)
        self.qp_iters = torch.zeros((B,), **i32)
        self.ws = me.AdmmWorkspace(B, N, device=device)
        self.U_warm = torch.zeros((B, N, NU), **f64)
        self.X_warm = torch.zeros((B, N + 1, NX), **f64)
        self.has_warm = False
        self._pinned = [{}, {}]   # two alternating sets of pinned staging buffers for the host results
        self._flip = 0

    def from_host(self, key: str, arr):
        """numpy -> device through a persistent pinned staging buffer (async H2D on the current stream).  Safe to reuse
        on the next call: every numpy-in call ends with the stream synchronisation of `to_host`."""
        torch = ge.torch_()
        a = np.ascontiguousarray(np.asarray(arr, dtype=np.float64))
        b = self._pinned_in.get(key)
        if b is None or tuple(b.shape) != a.shape:
            b = torch.empty(a.shape, dtype=torch.float64, device="cpu", pin_memory=True)
            self._pinned_in[key] = b
        b.numpy()[...] = a
        return b.to("cuda", non_blocking=True)

    def to_host(self, tensors: dict) -> dict:
        """Device -> pinned host copies of the result tensors (async on the current stream, one synchronise).
        The returned numpy arrays are views of pinned buffers that stay valid until the next-but-one call."""
        torch = ge.torch_()
        self._flip ^= 1
        bufs = self._pinned[self._flip]
        out = {}
        for k, t in tensors.items():
            b = bufs.get(k)
            if b is None or b.shape != t.shape or b.dtype != t.dtype:
                b = torch.empty(t.shape, dtype=t.dtype, device="cpu", pin_memory=True)
                bufs[k] = b
            b.copy_(t, non_blocking=True)
            out[k] = b
        torch.cuda.current_stream().synchronize()
        return {k: v.numpy() for k, v in out.items()}


class GPMPC:
    """gp_mpc.py:66-502 (3-DoF specialisation; `gp_model` = `Simple3DoFGP` or None for nominal MPC)."""

    def __init__(self, dynamics, gp_model, config: Optional[GPMPCConfig] = None,
                 constraint_params: Optional[ConstraintParams] = None, cost_weights: Optional[CostWeights] = None):
        self.dynamics = dynamics
        self.gp = gp_model
        self.config = config or GPMPCConfig()
        self.constraint_params = constraint_params or ConstraintParams()
        self.cost_weights = cost_weights or CostWeights()
        self.n_x, self.n_u = NX, NU
        self._tightened_constraints = TightenedConstraints(base_params=self.constraint_params,
                                                           confidence_level=self.config.confidence_level)
        self._X_warm = None
        self._U_warm = None
        self._single: Optional[_BatchState] = None
        self._batch: Optional[_BatchState] = None
        self._model = None
        self._model_struct = None
        self._qp = None
        self._dyn = None
        self._is_setup = False
        self._side = None
        self.last_kernel_launches = 0
        self.timers: Optional[me.PhaseTimers] = None   # set to me.PhaseTimers() to time the phases (bench.py)

    def _span(self, name):
        return self.timers.span(name) if self.timers is not None else me.NULL_SPAN

    # ------------------------------------------------------------------------------------------------ setup
    def setup(self) -> None:
        """Upload the scenario-independent data: dynamics parameters, cost / bound vectors, the fitted GP factor."""
        ge.torch_()
        p = self.dynamics.params
        g_vec = getattr(p, "g_vec", None)
        if g_vec is None:
            g_vec = getattr(p, "g_I")
        self._dyn = me.make_dyn(p.alpha, np.asarray(g_vec, dtype=np.float64), self.config.dt)
        Qd, Qfd, Rd = self.cost_weights.diag_3dof()
        self._qp = me.QPData.reference(self.config.N, Q=Qd, Qf=Qfd, R=Rd)
        self.refresh_gp()
        self._is_setup = True

    def refresh_gp(self) -> None:
        """(Re)read the GP's fitted device state -- call after the GP was refit (e.g. new hyper-parameters)."""
        self._model = None
        self._model_struct = None
        if self.gp is not None and (self.config.use_gp_mean or self.config.use_gp_uncertainty):
            dm = self.gp.device_model() if hasattr(self.gp, "device_model") else None
            if dm is None and getattr(self.gp, "n_data", 0) > 0:
                raise _lib.GpmpcError("GPMPC needs an exact GP whose outputs share one factor "
                                      "(Simple3DoFGP(use_sparse=False)); sparse models are served by predict_batch")
            if dm is not None:
                self._model = dm
                self._model_struct = dm.struct()

    def reset_warm_start(self) -> None:
        """gp_mpc.py:494-497."""
        self._X_warm = None
        self._U_warm = None
        for st in (self._single, self._batch):
            if st is not None:
                st.has_warm = False
                st.ws.y.zero_()
                st.ws.rho.zero_()

    def reset_scenarios(self, mask, X0) -> None:
        """Per-scenario `reset_warm_start` for the batched state: scenarios with mask != 0 restart from the hover
        guess at mass X0[b,0] with zero duals (a fresh controller instance for a respawned Monte Carlo run)."""
        st = self._batch
        if st is None or not st.has_warm:
            return
        torch = ge.torch_()
        g0 = float(self.dynamics.params.g0)
        if _is_dev(mask):
            mask_d = mask.to(torch.int32).contiguous()
            hover0 = (ge.to_dev(X0)[:, 0] * g0).contiguous()
        else:   # host arrays: two small pinned H2D copies, one launch
            mask_d = st.from_host("reset_mask", np.asarray(mask) != 0).to(torch.int32)
            hover0 = st.from_host("reset_hover", np.asarray(X0, dtype=np.float64)[:, 0] * g0)
        if not st.U_warm.is_contiguous():
            st.U_warm = st.U_warm.contiguous()
        ge.call("gpmpc_reset_scenarios", st.B, st.N, NU, st.ws.y.shape[1], ge.ptr(mask_d), ge.ptr(hover0),
                ge.ptr(st.U_warm), ge.ptr(st.ws.y), ge.ptr(st.ws.rho), ge.stream())

    def _settings(self):
        c = self.config
        return _lib.AdmmSettings.reference_defaults(
            max_iter=c.qp_max_iter, eps_abs=c.qp_eps_abs, eps_rel=c.qp_eps_rel, check_every=c.qp_check_every,
            adaptive_rho_interval=c.qp_adaptive_rho_interval, scaling=c.qp_scaling,
            warm_x=1, warm_y=1, rho_from_input=1 if c.qp_keep_rho else 0, commit_y_solved_only=1)

    # ------------------------------------------------------------------------------------------------ core
    def _variance_and_cov(self, st: _BatchState, X, U, active):
        """GP variances of the rows of st.Ks (blocked DMMA TRSM with fused sum of squares) -> covariance chain."""
        f = self._model.f
        with self._span("gp_variance"):
            if self.config.variance_via_inverse:
                if st.var_work is None:
                    st.var_work = ge.torch_().empty(((f.n_pad // 64) * st.Ks.shape[0],), dtype=st.Ks.dtype,
                                                    device=st.Ks.device)
                sumsq = f.sumsq_linv(st.Ks, st.var_work, one_cta_per_sm=self.config.overlap_uncertainty)
            else:
                sumsq = f.sumsq_inplace(st.Ks)
        with self._span("cov"):
            ge.call("gpmpc_predict_var", ge.ptr(sumsq), None, st.B * st.N, 3, f.sv, 1e-10, ge.ptr(self._model.ystd),
                    ge.ptr(st.dvar), 3, 0, ge.stream())
            me.propagate_cov(self._dyn, X, U, st.dvar, st.covs, sigma0=1e-6, active=active)
        if self.config.variance_via_inverse:
            return 4  # triangular GEMM + partial-sum reduction + var + cov
        return 2 * (f.n_pad // 64) + 2  # launches: memset is not a kernel; (n/64 - 1) GEMM + n/64 W-GEMM + var + cov

    def _fork_uncertainty(self, st: _BatchState, X, U, active):
        """`_variance_and_cov` on the side stream: the blocked DMMA TRSM (tensor pipe) runs concurrently with the
        latency-bound ADMM kernel of the main stream.  Inputs that the main stream later overwrites are snapshotted."""
        torch = ge.torch_()
        if not self.config.overlap_uncertainty:
            return self._variance_and_cov(st, X, U, active)
        if self._side is None:
            self._side = torch.cuda.Stream(priority=-1)
        self._join_uncertainty(st)
        st.X_unc.copy_(X)
        st.U_unc.copy_(U)
        st.active_unc.copy_(active)
        main = torch.cuda.current_stream()
        self._side.wait_stream(main)
        with torch.cuda.stream(self._side):
            n = self._variance_and_cov(st, st.X_unc, st.U_unc, st.active_unc)
            st.unc_event = torch.cuda.Event()
            st.unc_event.record(self._side)
        return n + 3

    def _join_uncertainty(self, st: _BatchState):
        if st.unc_event is not None:
            ge.torch_().cuda.current_stream().wait_event(st.unc_event)
            st.unc_event = None

    def _solve_device(self, st: _BatchState, x0, x_ref, U_ref=None, active_in=None):
        """One GP-MPC solve for every scenario of `st` (all tensors on the device).  Returns #kernel launches."""
        c = self.config
        N = c.N
        use_model = self._model is not None
        use_mean = c.use_gp_mean and use_model
        use_unc = c.use_gp_uncertainty and use_model
        launches = 0
        # --- initial controls: warm start | U_ref | hover along the up axis (gp_mpc.py:266-275; the 3-DoF frame of
        #     nominal_mpc.py:651-657 / osqp_rti.py:427-431 has "up" = axis 0)
        if st.has_warm and c.warm_start:
            st.U.copy_(st.U_warm)
        elif U_ref is not None:
            st.U.copy_(U_ref[..., :N, :].expand_as(st.U))
        else:
            st.U.zero_()
            st.U[:, :, 0] = (x0[:, 0] * float(self.dynamics.params.g0))[:, None]
        if active_in is None:
            st.active.fill_(1)
        else:
            st.active.copy_(active_in)
        st.n_iter.zero_()
        st.qp_iters.zero_()
        st.converged.zero_()
        # --- 1. rollout with the GP mean
        need_ks = use_unc
        with self._span("rollout"):
            me.rollout(self._model_struct if (use_mean or use_unc) else None, self._dyn, x0, st.U, st.X,
                       st.dmean if use_model else None, st.Ks if need_ks else None, active=st.active)
        launches += 1
        if use_model and not use_mean:
            # uncertainty only: the GP mean must not enter the prediction (use_gp_mean=False)
            me.rollout(None, self._dyn, x0, st.U, st.X, None, None, active=st.active)
            launches += 1
        # --- 2. uncertainty propagation
        if use_unc:
            launches += self._fork_uncertainty(st, st.X, st.U, st.active)
        # --- 3. SQP iterations
        S = self._settings()
        tighten = use_unc and c.tighten_state_bounds
        if tighten:
            from scipy.stats import norm
            kappa = float(norm.ppf(c.confidence_level))
            if st.lb_t is None:
                n_var = self._qp.lb.shape[0]
                st.lb_t = ge.torch_().empty((st.B, n_var), dtype=self._qp.lb.dtype, device=self._qp.lb.device)
                st.ub_t = ge.torch_().empty_like(st.lb_t)
        for it in range(c.max_sqp_iter):
            if tighten:   # boxes from the latest covariance chain (initial propagation or the re-propagation below)
                self._join_uncertainty(st)
                me.tighten_bounds(self._qp, st.covs, kappa, c.max_variance_for_constraint, st.lb_t, st.ub_t)
                launches += 1
            if use_mean and it > 0:
                with self._span("gp_mean"):
                    me.gp_mean_points(self._model_struct, st.X, st.U, st.dmean, None, active=st.active)
                launches += 1
            with self._span("assemble"):
                me.qp_assemble(self._dyn, self._qp, st.X, st.U, x0, x_ref, st.AB, st.beq, st.q,
                               c_gp=st.dmean if use_mean else None, lin_mode=0, rhs_mode=1, active=st.active)
                me.pack_traj(st.X, st.U, st.ws.x, active=st.active)
            with self._span("admm"):
                me.admm_solve(self._qp, st.AB, st.beq, st.q, S, st.ws, active=st.active,
                              lb=st.lb_t if tighten else None, ub=st.ub_t if tighten else None)
            with self._span("update"):
                st.qp_iters += st.ws.iters * st.active
                me.mpc_update(self._qp, 1, st.ws.x, st.ws.status, x_ref, st.X, st.U, None, None, st.cost,
                              st.converged, active=st.active, n_iter=st.n_iter)
            launches += 5
            if it == c.max_sqp_iter - 1:
                break
            if int(st.active.sum().item()) == 0:
                break
            if use_unc and c.repropagate_uncertainty:  # gp_mpc.py:348-353
                self._join_uncertainty(st)   # the rollout below overwrites Ks
                with self._span("rollout"):
                    me.rollout(self._model_struct, self._dyn, x0, st.U, st.Xs, st.dmean_s, st.Ks, active=st.active)
                launches += 1 + self._fork_uncertainty(st, st.Xs, st.U, st.active)
        self._join_uncertainty(st)
        # --- 4. warm start for the next call (unshifted, gp_mpc.py:355-357)
        st.U_warm.copy_(st.U)
        st.X_warm.copy_(st.X)
        st.has_warm = True
        return launches

    def _state_for(self, B: int, single: bool) -> _BatchState:
        attr = "_single" if single else "_batch"
        st = getattr(self, attr)
        if st is None or st.B != B or st.N != self.config.N:
            st = _BatchState(B, self.config.N, self._model)
            setattr(self, attr, st)
        return st

    def _x_ref(self, x_target, X_ref, B):
        torch = ge.torch_()
        N = self.config.N
        if X_ref is not None:
            r = ge.to_dev(X_ref)
            return r[..., : N + 1, :].contiguous()
        t = ge.to_dev(x_target)
        if t.dim() == 1:
            return t[None, :].expand(N + 1, NX).contiguous()
        return t[:, None, :].expand(B, N + 1, NX).contiguous()

    # ------------------------------------------------------------------------------------------------ public API
    def solve_batch(self, X0, x_target, X_ref=None, U_ref=None, active=None) -> BatchMPCSolution:
        """B independent `solve` calls (each scenario keeps its own warm start between calls).

        X0 (B,7); x_target (7,) or (B,7); optional X_ref (N+1,7)|(B,N+1,7), U_ref (N,3)|(B,N,3); optional `active`
        (B,) mask -- scenarios with 0 are skipped (their outputs repeat the previous call).  numpy in -> numpy out,
        CUDA float64 tensors in -> CUDA tensors out."""
        if not self._is_setup:
            self.setup()
        torch = ge.torch_()
        t0 = time.perf_counter()
        dev_in = _is_dev(X0)
        if dev_in:
            x0 = ge.to_dev(X0)
            if x0.dim() == 1:
                x0 = x0[None, :]
            B = x0.shape[0]
            st = self._state_for(B, single=False)
        else:   # host buffers: H2D from pinned staging memory
            X0h = np.atleast_2d(np.asarray(X0, dtype=np.float64))
            B = X0h.shape[0]
            st = self._state_for(B, single=False)
            x0 = st.from_host("x0", X0h)
            if X_ref is None and not _is_dev(x_target):
                x_target = st.from_host("xt", x_target)
        x_ref = self._x_ref(x_target, X_ref, B)
        act = None if active is None else ge.to_dev(active, torch.int32)
        uref = None if U_ref is None else ge.to_dev(U_ref)
        self.last_kernel_launches = self._solve_device(st, x0, x_ref, uref, act)
        return self._collect(st, dev_in, t0)

    def _collect(self, st: _BatchState, dev_in: bool, t0: float, with_cov: bool = False) -> BatchMPCSolution:
        torch = ge.torch_()
        success = st.converged.to(torch.bool)
        status = st.converged.to(torch.int8)
        has_cov = self._model is not None and self.config.use_gp_uncertainty
        if dev_in:  # stream-ordered results, no host synchronisation (solve_time is then the launch time only)
            sol = BatchMPCSolution(
                u0=st.U[:, 0].clone(), X_opt=st.X.clone(), U_opt=st.U.clone(), cost=st.cost.clone(),
                iterations=st.n_iter.clone(), success=success, status=status, solve_time=0.0,
                qp_status=st.ws.status.clone(), qp_iterations=st.qp_iters.clone(),
                covariances=st.covs.clone() if has_cov else None)
        else:
            defer = self.config.defer_trajectory_copy and not with_cov
            t = dict(cost=st.cost, it=st.n_iter, ok=success, status=status, qs=st.ws.status, qi=st.qp_iters)
            if defer:
                t["u0"] = st.U[:, 0].contiguous()
            else:
                t["X"], t["U"] = st.X, st.U
            if has_cov and with_cov:
                t["cov"] = st.covs
            h = st.to_host(t)
            if defer:
                h["X"], h["U"] = DeferredHostArray(st.X.clone()), DeferredHostArray(st.U.clone())
            else:
                h["u0"] = h["U"][:, 0]
            sol = BatchMPCSolution(
                u0=h["u0"], X_opt=h["X"], U_opt=h["U"], cost=h["cost"], iterations=h["it"], success=h["ok"],
                status=h["status"], solve_time=0.0, qp_status=h["qs"], qp_iterations=h["qi"],
                covariances=h.get("cov"))
        sol.solve_time = time.perf_counter() - t0
        return sol

    def last_covariances(self, batch: bool = True):
        """Propagated state covariances (B,N+1,7,7) of the last `solve_batch` (or `solve`) as a numpy array.  The
        reference computes them and discards them (gp_mpc.py:486-492 returns None); they are kept on the device and
        copied to the host only on request (34 MB at B=4096)."""
        st = self._batch if batch else self._single
        if st is None or self._model is None or not self.config.use_gp_uncertainty:
            return None
        return st.covs.cpu().numpy()

    def solve(self, x0, x_target, X_ref=None, U_ref=None) -> MPCSolution:
        """gp_mpc.py:229-369 signature; a B = 1 view of the batched path with its own warm start."""
        if not self._is_setup:
            self.setup()
        t0 = time.perf_counter()
        x0d = ge.to_dev(np.asarray(x0, dtype=np.float64)[None, :])
        st = self._state_for(1, single=True)
        x_ref = self._x_ref(np.asarray(x_target, dtype=np.float64), X_ref, 1)
        uref = None if U_ref is None else ge.to_dev(np.asarray(U_ref, dtype=np.float64))
        self.last_kernel_launches = self._solve_device(st, x0d, x_ref, uref, None)
        b = self._collect(st, False, t0, with_cov=True)
        b.X_opt, b.U_opt, b.u0 = b.X_opt.copy(), b.U_opt.copy(), b.u0.copy()   # detach from the pinned staging
        if b.covariances is not None:
            b.covariances = b.covariances.copy()
        self._X_warm, self._U_warm = b.X_opt[0].copy(), b.U_opt[0].copy()
        self._last_batch = b
        return b[0]

    def get_uncertainty_at_horizon(self):
        """gp_mpc.py:486-492 returns None in the reference; here: the propagated covariances (N+1,7,7) of the last
        `solve`, when uncertainty propagation is on."""
        b = getattr(self, "_last_batch", None)
        return None if b is None or b.covariances is None else b.covariances[0]

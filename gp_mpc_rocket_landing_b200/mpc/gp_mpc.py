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
    # Extension (off = the OSQP-form box QP): polyhedral inner approximations of the thrust-norm and
    # glide-slope constraints the reference's sub-problem hands to IPOPT (gp_mpc.py:417-429) as rows of the ADMM QP,
    # parameters from `constraint_params` (T_min, T_max, gamma_gs); include/gpmpc_b200.h GpmpcCone3
    convex_constraints: bool = False
    qp_max_iter: int = 1000           # config/mpc_params.yaml:48 (osqp_rti.py:53 uses 50, which rarely converges)
    qp_eps_abs: float = 1e-4          # osqp_rti.py:54
    qp_eps_rel: float = 1e-4          # osqp_rti.py:55
    qp_check_every: int = 25          # OSQP default check_termination
    qp_adaptive_rho_interval: int = 25  # iteration based (OSQP's default is timing based, i.e. not reproducible)
    qp_scaling: int = 3               # osqp_rti.py:477
    qp_keep_rho: bool = True          # the adapted rho persists across solves, as inside one OSQP solver object
    qp_kernel_variant: int = 0        # GpmpcAdmmSettings.kernel_variant (0 = newest kernel that fits; 2 / 1 = older generations)
    variance_via_inverse: bool = True  # horizon variances |L^-1 k*|^2 as one triangular DMMA GEMM against the explicit
                                       # inverse factor (formed once per fit) instead of the blocked in-place solve
                                       # (31 dependent launches); same quantity, rounding differs by O(cond(L) eps)


def _is_dev(a):
    return hasattr(a, "is_cuda")


class _BatchState:
    """Device-resident state of B controller instances (buffers are allocated once and reused every step) and the
    `GpmpcSolve3` plan that hands their addresses to the fused C entry point."""

    def __init__(self, B: int, N: int, model, device="cuda"):
        torch = ge.torch_()
        f64 = dict(dtype=torch.float64, device=device)
        i32 = dict(dtype=torch.int32, device=device)
        n, m = me.qp_dims(N)
        self.B, self.N = B, N
        self.x_ref = torch.zeros((B, N + 1, NX), **f64)
        self.X = torch.zeros((B, N + 1, NX), **f64)
        self.U = torch.zeros((B, N, NU), **f64)
        self.Xs = torch.zeros((B, N + 1, NX), **f64)       # scratch rollout of the re-propagation
        self.dmean = torch.zeros((B, N, 3), **f64)
        self.dmean_s = torch.zeros((B, N, 3), **f64)
        self.q_pad = ge.pad_to(B * N, 128)
        self.dvar = torch.zeros((self.q_pad, 3), **f64)
        self.sumsq = torch.zeros((self.q_pad,), **f64)
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
        self.has_warm = torch.zeros((B,), **i32)     # per scenario: 0 -> hover / U_ref initial guess
        self.cone = None      # buffers of the augmented (14, 3) stage QP (GPMPCConfig.convex_constraints)
        self.plan = None
        self.generation = 0   # number of solves run on this state (DeferredHostArray checks it)
        self.keep = None      # tensors referenced by the plan for the duration of a call
        self._pinned = [{}, {}]   # two alternating sets of pinned staging buffers for the host results
        self._flip = 0

    @property
    def duals(self):
        """The dual state that persists between control steps (y of the QP that is actually solved)."""
        return self.cone["y"] if self.cone is not None else self.ws.y

    def from_host(self, key: str, arr):
        """numpy -> device through a persistent pinned staging buffer (async H2D on the current stream).  Safe to reuse
        on the next call: every numpy-in call ends with the stream synchronisation of `to_host`."""
        torch = ge.torch_()
        a = np.ascontiguousarray(np.asarray(arr, dtype=np.float64))
        b = self._pinned_in.get(key)
        if b is None or tuple(b[0].shape) != a.shape:
            b = (torch.empty(a.shape, dtype=torch.float64, device="cpu", pin_memory=True),
                 torch.empty(a.shape, dtype=torch.float64, device="cuda"))
            self._pinned_in[key] = b
        b[0].numpy()[...] = a
        b[1].copy_(b[0], non_blocking=True)
        return b[1]

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
        self.last_kernel_launches = 0
        self.timers: Optional[me.EventTimers] = None   # set to me.EventTimers(n) to time the phases (bench.py)

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
        for st in (self._single, self._batch):
            if st is not None:
                st.plan = None
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
                st.has_warm.zero_()
                st.duals.zero_()
                st.ws.rho.zero_()

    def reset_scenarios(self, mask, X0=None) -> None:
        """Per-scenario `reset_warm_start` for the batched state: scenarios with mask != 0 restart from the hover
        guess with zero duals and the default rho (a fresh controller instance for a respawned Monte Carlo run)."""
        st = self._batch
        if st is None:
            return
        torch = ge.torch_()
        if _is_dev(mask):
            mask_d = mask.to(torch.int32).contiguous()
        else:   # host mask: one small pinned H2D copy, one launch
            mask_d = st.from_host("reset_mask", (np.asarray(mask) != 0).astype(np.float64)).to(torch.int32)
        ge.call("gpmpc_reset_scenarios", st.B, st.N, NU, st.duals.shape[1], ge.ptr(mask_d), None, None, ge.ptr(st.duals),
                ge.ptr(st.ws.rho), ge.ptr(st.has_warm), ge.stream())

    def _settings(self):
        c = self.config
        return _lib.AdmmSettings.reference_defaults(
            max_iter=c.qp_max_iter, eps_abs=c.qp_eps_abs, eps_rel=c.qp_eps_rel, check_every=c.qp_check_every,
            adaptive_rho_interval=c.qp_adaptive_rho_interval, scaling=c.qp_scaling,
            warm_x=1, warm_y=1, rho_from_input=1 if c.qp_keep_rho else 0, commit_y_solved_only=1,
            kernel_variant=int(c.qp_kernel_variant))

    # ------------------------------------------------------------------------------------------------ core
    def _make_plan(self, st: _BatchState) -> "_lib.Solve3":
        """Fill the GpmpcSolve3 plan with the addresses of the state buffers (once per state / GP refresh)."""
        torch = ge.torch_()
        c = self.config
        use_model = self._model is not None
        use_unc = c.use_gp_uncertainty and use_model
        p = _lib.Solve3()
        p.B, p.N = st.B, st.N
        p.use_gp_mean = 1 if (c.use_gp_mean and use_model) else 0
        p.use_gp_uncertainty = 1 if use_unc else 0
        p.repropagate = 1 if c.repropagate_uncertainty else 0
        p.has_model = 1 if use_model else 0
        p.dyn = self._dyn
        p.g0 = float(self.dynamics.params.g0)
        if use_model:
            p.model = self._model_struct
            f = self._model.f
            if use_unc and c.variance_via_inverse:
                W = f.ensure_linv()
                p.Linv, p.ldi = ge.ptr(W), W.stride(0)
                if st.var_work is None:
                    st.var_work = torch.empty(((f.n_pad // 64) * st.q_pad,), dtype=torch.float64, device=st.Ks.device)
                p.var_work = ge.ptr(st.var_work)
            p.Ks, p.ldk = ge.ptr(st.Ks), st.Ks.stride(0)
        p.q_pad = st.q_pad
        if use_unc and c.tighten_state_bounds:
            from scipy.stats import norm
            p.tighten_kappa = float(norm.ppf(c.confidence_level))
            p.tighten_max_variance = float(c.max_variance_for_constraint)
            if st.lb_t is None:
                n_var = self._qp.lb.shape[0]
                st.lb_t = torch.empty((st.B, n_var), dtype=self._qp.lb.dtype, device=self._qp.lb.device)
                st.ub_t = torch.empty_like(st.lb_t)
            p.lb_t, p.ub_t = ge.ptr(st.lb_t), ge.ptr(st.ub_t)
        qp, ws = self._qp, st.ws
        for name, t in (("Pdiag", qp.Pdiag), ("lb", qp.lb), ("ub", qp.ub), ("Qd", qp.Qd), ("Qfd", qp.Qfd), ("Rd", qp.Rd),
                        ("x_ref", st.x_ref), ("X", st.X), ("U", st.U), ("Xs", st.Xs), ("dmean", st.dmean),
                        ("dmean_s", st.dmean_s), ("sumsq", st.sumsq), ("dvar", st.dvar), ("covs", st.covs),
                        ("AB", st.AB), ("beq", st.beq), ("q", st.q), ("cost", st.cost), ("converged", st.converged),
                        ("active", st.active), ("n_iter", st.n_iter), ("qp_iters", st.qp_iters), ("z", ws.x),
                        ("y", ws.y), ("rho", ws.rho), ("obj", ws.obj), ("pri", ws.pri), ("dua", ws.dua),
                        ("qp_status", ws.status), ("qp_it", ws.iters), ("U_warm", st.U_warm), ("X_warm", st.X_warm)):
            setattr(p, name, ge.ptr(t))
        if c.convex_constraints:
            self._plan_cone(st, p)
        return p

    def _plan_cone(self, st: _BatchState, p) -> None:
        """Buffers and shared data of the augmented stage QP [x (7); xi (7); u (3)] (GpmpcCone3)."""
        torch = ge.torch_()
        if self.config.tighten_state_bounds:
            raise _lib.GpmpcError("convex_constraints and tighten_state_bounds are exclusive")
        cp, N, B = self.constraint_params, st.N, st.B
        NXC, NSC = 2 * NX, 2 * NX + NU
        n_c = N * NSC + NXC
        m_c = NXC * (N + 1) + n_c
        qp = self._qp
        if st.cone is None:
            f64 = dict(dtype=torch.float64, device=st.X.device)
            Qd, Qfd, Rd = (np.asarray(t.cpu().numpy(), dtype=np.float64) for t in (qp.Qd, qp.Qfd, qp.Rd))
            lb7, ub7 = qp.lb.cpu().numpy(), qp.ub.cpu().numpy()
            Pd = np.zeros(n_c); lb = np.full(n_c, -me.INF); ub = np.full(n_c, me.INF)
            xi_hi = np.array([cp.T_max] * 4 + [0.0] * 3)
            for k in range(N + 1):
                o, o7 = k * NSC, k * (NX + NU)
                Pd[o:o + NX] = Qfd if k == N else Qd
                lb[o:o + NX], ub[o:o + NX] = lb7[o7:o7 + NX], ub7[o7:o7 + NX]
                if k > 0:
                    ub[o + NX:o + NX + 4] = xi_hi[:4]
                if k > 1:      # the position of stage 1 is x_0 + dt v_0: no decision enters its rows (redundant or infeasible)
                    ub[o + NX + 4:o + NXC] = xi_hi[4:]
                if k < N:
                    Pd[o + NXC:o + NSC] = Rd
                    lb[o + NXC:o + NSC], ub[o + NXC:o + NSC] = lb7[o7 + NX:o7 + NX + NU], ub7[o7 + NX:o7 + NX + NU]
                    lb[o + NXC] = max(lb[o + NXC], cp.T_min)      # |u| >= T_min (non-convex) -> u_x >= T_min
            st.cone = dict(AB=torch.zeros((B, N, NXC, NSC), **f64), beq=torch.zeros((B, NXC * (N + 1)), **f64),
                           q=torch.zeros((B, n_c), **f64), z=torch.zeros((B, n_c), **f64), y=torch.zeros((B, m_c), **f64),
                           Pdiag=ge.to_dev(Pd), lb=ge.to_dev(lb), ub=ge.to_dev(ub))
        p.cone.enabled = 1
        p.cone.T_max = float(cp.T_max)
        p.cone.facet_b = float(np.sqrt(2.0 * (cp.T_max - cp.T_min) / (cp.T_max + cp.T_min)))
        p.cone.half_tan_gs = float(np.tan(cp.gamma_gs_rad) / 2.0)
        for name, key in (("AB_c", "AB"), ("beq_c", "beq"), ("q_c", "q"), ("z_c", "z"), ("y_c", "y"), ("Pdiag_c", "Pdiag"),
                          ("lb_c", "lb"), ("ub_c", "ub")):
            setattr(p, name, ge.ptr(st.cone[key]))

    def _solve_device(self, st: _BatchState, x0, x_target=None, X_ref=None, U_ref=None, active_in=None,
                      per_step_target=False):
        """One GP-MPC solve for every scenario of `st` (all tensors on the device): ONE call of the fused C entry point
        `gpmpc_solve_batch_3dof` (8 kernel launches for S = 1, no host synchronisation).  Returns #kernel launches."""
        c = self.config
        if st.plan is None:
            st.plan = self._make_plan(st)
        p = st.plan
        p.max_sqp_iter = int(c.max_sqp_iter)
        p.admm = self._settings()
        p.target_mode = 1 if per_step_target else 0
        p.x0 = ge.ptr(x0)
        p.x_target = ge.ptr(x_target) if x_target is not None else None
        if X_ref is not None:
            p.X_ref, p.ld_xref_in = ge.ptr(X_ref), (0 if X_ref.dim() == 2 else X_ref.stride(0))
        else:
            p.X_ref, p.ld_xref_in = None, 0
        if U_ref is not None:
            p.U_ref, p.ld_uref = ge.ptr(U_ref), (0 if U_ref.dim() == 2 else U_ref.stride(0))
        else:
            p.U_ref, p.ld_uref = None, 0
        p.active_in = ge.ptr(active_in) if active_in is not None else None
        p.has_warm = ge.ptr(st.has_warm) if c.warm_start else None
        if self.timers is not None:
            self.timers.attach(p)
        else:
            p.events, p.event_phase, p.event_cap, p.event_count = None, None, 0, None
        st.keep = (x0, x_target, X_ref, U_ref, active_in)
        st.generation += 1
        n = _lib.c_int(0)
        ge.call("gpmpc_solve_batch_3dof", _lib.C.byref(p), _lib.C.byref(n), ge.stream())
        return n.value

    def _state_for(self, B: int, single: bool) -> _BatchState:
        attr = "_single" if single else "_batch"
        st = getattr(self, attr)
        if st is None or st.B != B or st.N != self.config.N:
            st = _BatchState(B, self.config.N, self._model)
            setattr(self, attr, st)
        return st

    def _targets(self, st, x_target, X_ref, U_ref, B):
        """Device views of the target / reference arguments in the layouts GpmpcSolve3 takes."""
        N = self.config.N
        xr = ur = xt = None
        if X_ref is not None:
            xr = ge.to_dev(X_ref)[..., : N + 1, :].contiguous()
        elif x_target is not None:
            xt = ge.to_dev(x_target)
            if xt.dim() == 1:
                xt = xt[None, :].expand(B, NX)
            xt = xt.contiguous()
        if U_ref is not None:
            ur = ge.to_dev(U_ref)[..., :N, :].contiguous()
        return xt, xr, ur

    # ------------------------------------------------------------------------------------------------ public API
    def solve_batch(self, X0, x_target=None, X_ref=None, U_ref=None, active=None,
                    per_step_target: bool = False) -> BatchMPCSolution:
        """B independent `solve` calls (each scenario keeps its own warm start between calls).

        X0 (B,7); x_target (7,) or (B,7); optional X_ref (N+1,7)|(B,N+1,7), U_ref (N,3)|(B,N,3); optional `active`
        (B,) mask -- scenarios with 0 are skipped (their outputs repeat the previous call).  per_step_target=True derives
        every scenario's target from its current state on the device, as the reference's Monte Carlo loop does for
        controllers with `.solve` (monte_carlo.py:497-499: x_target = x, v = 0, altitude = max(0.5, altitude - 2));
        x_target is then ignored.  numpy in -> numpy out; CUDA float64 tensors in -> CUDA tensors out, which are VIEWS of
        the controller state, valid until the next call (clone what must survive it)."""
        if not self._is_setup:
            self.setup()
        torch = ge.torch_()
        t0 = time.perf_counter()
        dev_in = _is_dev(X0)
        if dev_in:
            x0 = ge.to_dev(X0)
            if x0.dim() == 1:
                x0 = x0[None, :]
            x0 = x0.contiguous()
            B = x0.shape[0]
            st = self._state_for(B, single=False)
        else:   # host buffers: H2D from pinned staging memory
            X0h = np.atleast_2d(np.asarray(X0, dtype=np.float64))
            B = X0h.shape[0]
            st = self._state_for(B, single=False)
            x0 = st.from_host("x0", X0h)
            if X_ref is None and x_target is not None and not per_step_target and not _is_dev(x_target):
                xth = np.asarray(x_target, dtype=np.float64)
                x_target = st.from_host("xt", np.broadcast_to(xth, (B, NX)))
        if per_step_target:
            x_target = None
        elif x_target is None and X_ref is None:
            raise ValueError("solve_batch needs x_target, X_ref or per_step_target=True")
        xt, xr, ur = self._targets(st, x_target, X_ref, U_ref, B)
        act = None if active is None else ge.to_dev(active, torch.int32).contiguous()
        self.last_kernel_launches = self._solve_device(st, x0, xt, xr, ur, act, per_step_target)
        return self._collect(st, dev_in, t0)

    def _collect(self, st: _BatchState, dev_in: bool, t0: float, with_cov: bool = False) -> BatchMPCSolution:
        torch = ge.torch_()
        has_cov = self._model is not None and self.config.use_gp_uncertainty
        if dev_in:  # stream-ordered views of the controller state: no copies, no host synchronisation
            sol = BatchMPCSolution(
                u0=st.U[:, 0], X_opt=st.X, U_opt=st.U, cost=st.cost, iterations=st.n_iter, success=st.converged,
                status=st.converged, solve_time=0.0, qp_status=st.ws.status, qp_iterations=st.qp_iters,
                covariances=st.covs if has_cov else None)
        else:
            defer = self.config.defer_trajectory_copy and not with_cov
            t = dict(cost=st.cost, it=st.n_iter, ok=st.converged, qs=st.ws.status, qi=st.qp_iters)
            if defer:
                t["u0"] = st.U[:, 0].contiguous()
            else:
                t["X"], t["U"] = st.X, st.U
            if has_cov and with_cov:
                t["cov"] = st.covs
            h = st.to_host(t)
            if defer:
                h["X"], h["U"] = DeferredHostArray(st.X, st), DeferredHostArray(st.U, st)
            else:
                h["u0"] = h["U"][:, 0]
            ok = h["ok"] != 0
            sol = BatchMPCSolution(
                u0=h["u0"], X_opt=h["X"], U_opt=h["U"], cost=h["cost"], iterations=h["it"], success=ok,
                status=ok.astype(np.int8), solve_time=0.0, qp_status=h["qs"], qp_iterations=h["qi"],
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
        xt, xr, ur = self._targets(st, np.asarray(x_target, dtype=np.float64), X_ref, U_ref, 1)
        self.last_kernel_launches = self._solve_device(st, x0d, xt, xr, ur, None, False)
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

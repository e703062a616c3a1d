"""Device-side MPC engine: thin Python over the MPC entry points of the C-ABI (include/gpmpc_b200.h).

QP layout (reference osqp_rti.py): z = [x_0,u_0,...,x_{N-1},u_{N-1},x_N], n = 10N+7,
rows = [x_0 = x_init (7) | A_k x_k + B_k u_k - x_{k+1} = b_k (7N) | l <= I z <= u (n)], m = 7(N+1)+n.
The dynamics blocks are stored dense per stage (AB: B x N x 7 x 10, explicit zeros = fixed superset pattern).
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import numpy as np

from .. import _lib
from .._lib import AdmmSettings, Dyn3, call, ptr, stream
from ..gp.engine import to_dev, torch_

NX, NU, NS = 7, 3, 10
INF = 1e30

# osqp_rti.py:168-201 cost weights and box bounds
Q_DIAG = np.array([0.0, 10.0, 10.0, 10.0, 1.0, 1.0, 1.0])
QF_DIAG = Q_DIAG * 10.0
R_DIAG = np.ones(3) * 0.01
X_MIN = np.array([-np.inf, -100.0, -100.0, -100.0, -50.0, -50.0, -50.0])
X_MAX = np.array([np.inf, 500.0, 100.0, 100.0, 50.0, 50.0, 50.0])
U_MIN = np.array([0.3, -5.0, -5.0])
U_MAX = np.array([5.0, 5.0, 5.0])


class PhaseTimers:
    """CUDA-event timers around the phases of a batched solve (events on the launching stream).  Used by bench.py to
    attribute the step time to kernels live, inside the timed region; disabled (None) in normal operation."""

    def __init__(self):
        self.pairs = {}

    class _Span:
        def __init__(self, owner, name):
            self.o, self.n = owner, name

        def __enter__(self):
            torch = torch_()
            self.a = torch.cuda.Event(enable_timing=True)
            self.b = torch.cuda.Event(enable_timing=True)
            self.a.record()

        def __exit__(self, *exc):
            self.b.record()
            self.o.pairs.setdefault(self.n, []).append((self.a, self.b))

    def span(self, name):
        return PhaseTimers._Span(self, name)

    def totals(self):
        """name -> (total ms, number of spans); synchronises."""
        torch_().cuda.synchronize()
        return {k: (sum(a.elapsed_time(b) for a, b in v), len(v)) for k, v in self.pairs.items()}

    def reset(self):
        self.pairs = {}


class EventTimers:
    """CUDA-event phase attribution of the fused solve (`gpmpc_solve_batch_3dof` records one event per phase boundary
    on the launching stream when a pool is attached).  Used by bench.py to time every kernel live, inside the timed region."""

    def __init__(self, capacity: int = 4096):
        torch_()
        self.cap = int(capacity)
        self.events = (C.c_void_p * self.cap)()
        self.phase = (C.c_int * self.cap)()
        self.count = C.c_int(0)
        call("gpmpc_events_create", self.cap, self.events)

    def attach(self, plan):
        plan.events = C.cast(self.events, C.POINTER(C.c_void_p))
        plan.event_phase = C.cast(self.phase, C.POINTER(C.c_int))
        plan.event_cap = self.cap
        plan.event_count = C.pointer(self.count)

    def reset(self):
        self.count.value = 0

    def totals(self):
        """phase name -> (total ms, number of spans); the time between the last event of one solve and the `begin` of
        the next (plant step, respawn, host gaps) is reported as "between_solves".  Synchronises."""
        torch_().cuda.synchronize()
        out = {}
        ms = C.c_float(0.0)
        n = self.count.value
        for i in range(1, n):
            ph = self.phase[i]
            name = "between_solves" if ph == 0 else _lib.PHASE_NAMES[ph]
            call("gpmpc_events_elapsed_ms", self.events[i - 1], self.events[i], C.byref(ms))
            t, c = out.get(name, (0.0, 0))
            out[name] = (t + ms.value, c + 1)
        return out

    def step_times_ms(self):
        """Per-solve wall time on the device: begin(i) -> begin(i+1) for consecutive solves (includes what runs between
        them), for the p50 / p99 step latency."""
        torch_().cuda.synchronize()
        idx = [i for i in range(self.count.value) if self.phase[i] == 0]
        ms = C.c_float(0.0)
        out = []
        for a, b in zip(idx[:-1], idx[1:]):
            call("gpmpc_events_elapsed_ms", self.events[a], self.events[b], C.byref(ms))
            out.append(ms.value)
        return out

    def __del__(self):
        try:
            _lib.lib.gpmpc_events_destroy(self.cap, self.events)
        except Exception:  # noqa: BLE001
            pass


class _NullSpan:
    def __enter__(self):
        return None

    def __exit__(self, *exc):
        return False


NULL_SPAN = _NullSpan()


def qp_dims(N, nx=NX, nu=NU):
    n = N * (nx + nu) + nx
    return n, nx * (N + 1) + n


def make_dyn(alpha, g_vec, dt):
    d = Dyn3()
    d.alpha = float(alpha)
    for i in range(3):
        d.g[i] = float(g_vec[i])
    d.dt = float(dt)
    return d


def stage_vector(N, xpart, upart):
    """Tile per-stage (7,) and (3,) vectors into the z ordering (n,)."""
    n, _ = qp_dims(N)
    v = np.zeros(n)
    for k in range(N):
        v[k * NS:k * NS + NX] = xpart
        v[k * NS + NX:(k + 1) * NS] = upart
    v[N * NS:] = xpart
    return v


@dataclass
class QPData:
    """Shared (scenario-independent) QP data on the device."""
    N: int
    Pdiag: object
    lb: object
    ub: object
    Qd: object
    Qfd: object
    Rd: object

    @classmethod
    def reference(cls, N, Q=Q_DIAG, Qf=QF_DIAG, R=R_DIAG, x_min=X_MIN, x_max=X_MAX, u_min=U_MIN, u_max=U_MAX):
        n, _ = qp_dims(N)
        P = stage_vector(N, Q, R)
        P[N * NS:] = Qf
        lb = np.maximum(stage_vector(N, x_min, u_min), -INF)
        ub = np.minimum(stage_vector(N, x_max, u_max), INF)
        return cls(N, to_dev(P), to_dev(lb), to_dev(ub), to_dev(Q), to_dev(Qf), to_dev(R))


class AdmmWorkspace:
    """Per-scenario solver state that OSQP keeps between solves (x, y, rho) + per-solve outputs."""

    def __init__(self, B, N, device=None, nx=NX, nu=NU):
        torch = torch_()
        n, m = qp_dims(N, nx, nu)
        kw = dict(dtype=torch.float64, device=device or "cuda")
        self.B, self.N, self.n, self.m, self.nx, self.nu = B, N, n, m, nx, nu
        self.x = torch.zeros((B, n), **kw)
        self.y = torch.zeros((B, m), **kw)
        self.rho = torch.zeros((B,), **kw)
        self.obj = torch.zeros((B,), **kw)
        self.pri = torch.zeros((B,), **kw)
        self.dua = torch.zeros((B,), **kw)
        self.status = torch.zeros((B,), dtype=torch.int32, device=kw["device"])
        self.iters = torch.zeros((B,), dtype=torch.int32, device=kw["device"])


def admm_solve(qp: QPData, AB, beq, q, settings: AdmmSettings, ws: AdmmWorkspace, active=None, lb=None, ub=None):
    """lb, ub: optional per-scenario boxes (B, n) replacing the shared qp.lb / qp.ub (settings.bound_stride)."""
    settings.bound_stride = 0 if lb is None else int(lb.stride(0))
    lb, ub = (qp.lb, qp.ub) if lb is None else (lb, ub)
    call("gpmpc_admm_solve", ws.nx, ws.nu, qp.N, ws.B, ptr(AB), ptr(beq), ptr(q), ptr(qp.Pdiag), ptr(lb), ptr(ub),
         C.byref(settings), ptr(active), ptr(ws.x), ptr(ws.y), ptr(ws.rho), ptr(ws.status), ptr(ws.iters),
         ptr(ws.obj), ptr(ws.pri), ptr(ws.dua), stream())
    return ws


def rollout(model_struct, dyn, x0, U, X, dmean=None, Ks=None, active=None):
    B, N = U.shape[0], U.shape[1]
    call("gpmpc_rollout_3dof", C.byref(model_struct) if model_struct is not None else None, C.byref(dyn), B, N,
         ptr(x0), ptr(U), ptr(X), ptr(dmean), ptr(Ks), Ks.stride(0) if Ks is not None else 0, ptr(active), stream())


def gp_mean_points(model_struct, X, U, dmean, Ks=None, active=None):
    """GP means at the trajectory points (X[b,k], U[b,k]), k < N (gp_mpc.py:309-317)."""
    B, N = U.shape[0], U.shape[1]
    call("gpmpc_gp_mean_points_3dof", C.byref(model_struct), B, N, ptr(X), ptr(U), ptr(dmean), ptr(Ks),
         Ks.stride(0) if Ks is not None else 0, ptr(active), stream())


def pack_traj(X, U, z, active=None):
    """z = [x_0,u_0,...,x_N] (osqp_rti.py:601-614)."""
    B, N = U.shape[0], U.shape[1]
    call("gpmpc_pack_traj_3dof", B, N, ptr(X), ptr(U), ptr(z), ptr(active), stream())


def linearize(dyn, X, U, AB, lin_mode=0):
    """[A_d | B_d] at the flattened points of X (...,7), U (...,3) -> AB (...,7,10)."""
    npts = X.numel() // NX
    call("gpmpc_linearize_3dof", C.byref(dyn), npts, lin_mode, ptr(X), ptr(U), ptr(AB), stream())


def tighten_bounds(qp: QPData, covs, kappa, max_variance, lb_out, ub_out, nx=NX, nu=NU, active=None):
    """Per-scenario boxes lb + kappa sigma_k, ub - kappa sigma_k for the states of stages k >= 1 (gpmpc_tighten_bounds)."""
    B = covs.shape[0]
    call("gpmpc_tighten_bounds", nx, nu, B, qp.N, ptr(qp.lb), ptr(qp.ub), ptr(covs), float(kappa), float(max_variance),
         ptr(lb_out), ptr(ub_out), ptr(active), stream())


def propagate_cov(dyn, X, U, dvar, covs, sigma0=1e-6, active=None):
    B, N = U.shape[0], U.shape[1]
    call("gpmpc_propagate_cov_3dof", C.byref(dyn), B, N, ptr(X), ptr(U), ptr(dvar), float(sigma0), ptr(covs),
         ptr(active), stream())


def qp_assemble(dyn, qp: QPData, X_lin, U_lin, x_init, x_ref, AB, beq, q, c_gp=None, lin_mode=0, rhs_mode=0,
                active=None):
    """x_ref: (N+1,7) shared by all scenarios or (B,N+1,7)."""
    B, N = U_lin.shape[0], U_lin.shape[1]
    ldxref = 0 if x_ref.dim() == 2 else x_ref.stride(0)
    call("gpmpc_qp_assemble_3dof", C.byref(dyn), B, N, lin_mode, rhs_mode, ptr(X_lin), ptr(U_lin), ptr(x_init),
         ptr(x_ref), ldxref, ptr(c_gp), ptr(qp.Qd), ptr(qp.Qfd), ptr(AB), ptr(beq), ptr(q), ptr(active), stream())


def mpc_update(qp: QPData, mode, z, qp_status, x_ref, X, U, X_prev, U_prev, cost, converged, active=None, n_iter=None):
    B, N = U.shape[0], U.shape[1]
    ldxref = 0 if x_ref.dim() == 2 else x_ref.stride(0)
    call("gpmpc_mpc_update_3dof", B, N, mode, ptr(z), ptr(qp_status), ptr(x_ref), ldxref, ptr(qp.Qd), ptr(qp.Qfd),
         ptr(qp.Rd), ptr(X), ptr(U), ptr(X_prev), ptr(U_prev), ptr(cost), ptr(converged), ptr(active), ptr(n_iter),
         stream())


def dense_constraints(N, AB, beq, lb, ub):
    """Host helper (tests / debugging): dense (A, l, u) of one scenario from the stage-block storage."""
    n, m = qp_dims(N)
    A = np.zeros((m, n)); l = np.zeros(m); u = np.zeros(m)
    A[:NX, :NX] = np.eye(NX)
    for k in range(N):
        r = NX + k * NX
        A[r:r + NX, k * NS:(k + 1) * NS] = AB[k]
        A[r:r + NX, (k + 1) * NS:(k + 1) * NS + NX] = -np.eye(NX)
    A[NX * (N + 1):] = np.eye(n)
    l[:NX * (N + 1)] = u[:NX * (N + 1)] = beq
    l[NX * (N + 1):] = lb; u[NX * (N + 1):] = ub
    return A, l, u

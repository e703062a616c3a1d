"""`MPCConfig`, `MPCSolution` and `NominalMPC3DoF` with the reference's API (src/mpc/nominal_mpc.py:42-82, :537-679).

The reference's `NominalMPC3DoF` is a CasADi/IPOPT NLP (third-party, absent).  Here it is the same 3-DoF model and
cost (nominal_mpc.py:579-631: Euler dynamics, Q = diag(0,10,10,10,1,1,1), R = 0.01 I, terminal 10 Q) solved by the
batched SQP-on-OSQP-form-QP path of `GPMPC` with no GP -- see gp_mpc.py in this package and DESIGN.md "Deviations".
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Any, Optional


@dataclass
class MPCConfig:
    """nominal_mpc.py:42-64."""

    N: int = 20
    dt: float = 0.1
    integration_method: str = "rk4"
    max_iter: int = 100
    tol: float = 1e-6
    warm_start: bool = True
    verbose: bool = False
    soft_constraints: bool = False
    constraint_slack_weight: float = 1e4
    use_reference: bool = False


@dataclass
class MPCSolution:
    """nominal_mpc.py:67-82."""

    success: bool
    X_opt: Any
    U_opt: Any
    cost: float
    solve_time: float
    iterations: int
    status: str

    @property
    def u0(self):
        return self.U_opt[0]


class DeferredHostArray:
    """A device array that becomes a numpy array the first time the owning solution's attribute is read.  `owner` (the
    controller's batch state) counts its solves: reading after a later solve overwrote the array is an error, not a
    silent read of newer data."""

    def __init__(self, tensor, owner=None):
        self.tensor = tensor
        self.owner = owner
        self.generation = getattr(owner, "generation", None)

    def fetch(self):
        if self.owner is not None and self.owner.generation != self.generation:
            raise RuntimeError("X_opt / U_opt of an earlier solve_batch were overwritten by a later call: read them before "
                               "the next call, or set GPMPCConfig.defer_trajectory_copy=False")
        return self.tensor.cpu().numpy()


@dataclass
class BatchMPCSolution:
    """Result of `solve_batch` (new entry point): B independent `MPCSolution`s as arrays.

    u0 (B,3), X_opt (B,N+1,7), U_opt (B,N,3), cost (B,), iterations (B,) int, success (B,) bool (numpy calls; 0/1 int32
    for CUDA-tensor calls, whose results are views of the controller state valid until the next call),
    status (B,) int8 (1 = "Converged", 0 = "Max iterations"), qp_status (B,) int (OSQP code of the last QP),
    qp_iterations (B,) int (ADMM iterations summed over the SQP loop), covariances (B,N+1,7,7) or None.
    numpy arrays when the inputs were numpy, CUDA tensors when they were CUDA tensors.  For numpy calls u0, cost and the
    flags are copied to the host at once; X_opt / U_opt (7 MB at B = 4096) are device snapshots that turn into numpy arrays
    on first access (`GPMPCConfig.defer_trajectory_copy`), since a closed loop reads u0 only."""

    u0: Any
    X_opt: Any
    U_opt: Any
    cost: Any
    iterations: Any
    success: Any
    status: Any
    solve_time: float
    qp_status: Any = None
    qp_iterations: Any = None
    covariances: Any = None

    _DEFERRABLE = ("X_opt", "U_opt")

    def __getattribute__(self, name):
        v = object.__getattribute__(self, name)
        if name in BatchMPCSolution._DEFERRABLE and isinstance(v, DeferredHostArray):
            v = v.fetch()                       # host-buffer calls: the trajectories cross PCIe only when they are read
            object.__setattr__(self, name, v)
        return v

    def __len__(self):
        return self.u0.shape[0]

    def __getitem__(self, b: int) -> MPCSolution:
        ok = bool(self.success[b])
        return MPCSolution(success=ok, X_opt=self.X_opt[b], U_opt=self.U_opt[b], cost=float(self.cost[b]),
                           solve_time=self.solve_time, iterations=int(self.iterations[b]),
                           status="Converged" if ok else "Max iterations")


class NominalMPC3DoF:
    """nominal_mpc.py:537-679 signature: `NominalMPC3DoF(dynamics, config=None)`, `.setup()`, `.solve(x0, x_target)`."""

    def __init__(self, dynamics, config: Optional[MPCConfig] = None):
        from .gp_mpc import GPMPC, GPMPCConfig
        self.dynamics = dynamics
        self.config = config or MPCConfig()
        c = self.config
        self._impl = GPMPC(dynamics, None, GPMPCConfig(N=c.N, dt=c.dt, warm_start=c.warm_start, verbose=c.verbose,
                                                       use_gp_mean=False, use_gp_uncertainty=False))
        self.n_x, self.n_u = 7, 3
        self._is_setup = False

    def setup(self) -> None:
        self._impl.setup()
        self._is_setup = True

    def solve(self, x0, x_target) -> MPCSolution:
        if not self._is_setup:
            self.setup()
        return self._impl.solve(x0, x_target)

    def solve_batch(self, X0, x_target=None, **kw) -> BatchMPCSolution:
        if not self._is_setup:
            self.setup()
        return self._impl.solve_batch(X0, x_target, **kw)

    @property
    def _batch(self):
        return self._impl._batch

    def reset_warm_start(self) -> None:
        self._impl.reset_warm_start()

    def reset_scenarios(self, mask, X0=None) -> None:
        self._impl.reset_scenarios(mask, X0)

"""3-DoF rocket dynamics object with the duck-typed protocol the reference's controllers consume
(`step(x,u,dt)`, `linearize(x,u,dt)`, `.params.{g0, alpha, g_vec, ...}`; src/dynamics/rocket_3dof.py:33-157).

The reference wraps the third-party `simdyn` package (absent, undeclared -- SURVEY R4); the arithmetic here is the
in-repo Euler statement (nominal_mpc.py:586-605) and the analytic discrete Jacobians (osqp_rti.py:682-708), evaluated
by the same device kernels the batched hot path uses (B = 1 views).  Any object with this protocol can be passed to
the controllers instead; only `.params.alpha / g_vec` and `dt` reach the device.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Optional

import numpy as np


@dataclass
class Rocket3DoFConfig:
    """rocket_3dof.py:33-82."""

    m_dry: float = 1.0
    m_wet: float = 2.0
    I_sp: float = 30.0
    g0: float = 1.0
    T_min: float = 0.0
    T_max: float = 6.5
    g_I: Optional[np.ndarray] = None  # noqa: N815
    enable_drag: bool = False
    rho: float = 1.0
    C_D: float = 0.5
    A_ref: float = 0.5
    gamma_gs: float = np.deg2rad(30.0)
    v_max: float = np.inf
    default_dt: float = 0.1

    def __post_init__(self):
        if self.g_I is None:
            self.g_I = np.array([-1.0, 0.0, 0.0])

    @classmethod
    def normalized_defaults(cls) -> "Rocket3DoFConfig":
        return cls()

    @classmethod
    def fuel_optimal_defaults(cls) -> "Rocket3DoFConfig":
        return cls(m_wet=2.0, m_dry=1.0, T_min=0.3, T_max=5.0, I_sp=300.0, g_I=np.array([-9.81, 0.0, 0.0]))


@dataclass
class Rocket3DoFParams:
    """The `simdyn` parameter record the reference reads (`params.g0`, `.alpha`, `.g_vec`, `.m_dry` ...)."""

    m_dry: float = 1.0
    m_wet: float = 2.0
    I_sp: float = 30.0
    g0: float = 1.0
    T_min: float = 0.0
    T_max: float = 6.5
    g_vec: np.ndarray = field(default_factory=lambda: np.array([-1.0, 0.0, 0.0]))
    gamma_gs: float = np.deg2rad(30.0)

    @property
    def alpha(self) -> float:
        return 1.0 / (self.I_sp * self.g0)


class Rocket3DoFDynamics:
    """rocket_3dof.py:85-157 interface on the Euler model; state [m, r(3), v(3)], control T(3)."""

    IDX_MASS = 0
    IDX_POS = slice(1, 4)
    IDX_VEL = slice(4, 7)
    N_STATE = 7
    N_CONTROL = 3

    def __init__(self, config: Optional[Rocket3DoFConfig] = None):
        self.config = config or Rocket3DoFConfig.normalized_defaults()
        c = self.config
        self._params = Rocket3DoFParams(m_dry=c.m_dry, m_wet=c.m_wet, I_sp=c.I_sp, g0=c.g0, T_min=c.T_min,
                                        T_max=c.T_max, g_vec=np.array(c.g_I, dtype=np.float64), gamma_gs=c.gamma_gs)

    @property
    def params(self) -> Rocket3DoFParams:
        return self._params

    @property
    def n_state(self) -> int:
        return self.N_STATE

    @property
    def n_control(self) -> int:
        return self.N_CONTROL

    def create_initial_state(self, altitude: float = 10.0, mass: Optional[float] = None, velocity=None, lateral=None):
        x = np.zeros(7)
        x[0] = self.config.m_wet if mass is None else mass
        x[1] = altitude
        if lateral is not None:
            x[2:4] = lateral
        if velocity is not None:
            x[4:7] = velocity
        return x

    def hover_thrust(self, x):
        """Thrust that cancels gravity at mass x[0]."""
        return -np.asarray(x)[0] * self._params.g_vec

    # ---- numerics on the device (B = 1 views of the batched kernels) ---------------------------------------------
    def _dyn(self, dt):
        from .mpc import engine as me
        return me.make_dyn(self._params.alpha, self._params.g_vec, self.config.default_dt if dt is None else dt)

    def step_batch(self, X, U, dt=None):
        """(B,7),(B,3) -> (B,7): one Euler step per scenario (nominal_mpc.py:586-605)."""
        from .gp import engine as ge
        from .mpc import engine as me
        torch = ge.torch_()
        dev_in = hasattr(X, "is_cuda")
        Xd, Ud = ge.to_dev(X), ge.to_dev(U)
        B = Xd.shape[0]
        out = torch.empty((B, 2, 7), dtype=torch.float64, device=Xd.device)
        me.rollout(None, self._dyn(dt), Xd, Ud.reshape(B, 1, 3), out)
        Xn = out[:, 1]
        return Xn if dev_in else Xn.cpu().numpy()

    def step(self, x, u, dt=None):
        return self.step_batch(np.asarray(x, dtype=np.float64)[None], np.asarray(u, dtype=np.float64)[None], dt)[0]

    def linearize(self, x, u, dt=None):
        """Discrete Jacobians (A_d (7,7), B_d (7,3)), osqp_rti.py:682-708, without the sparsity filter."""
        from .gp import engine as ge
        from .mpc import engine as me
        torch = ge.torch_()
        x = np.asarray(x, dtype=np.float64)
        u = np.asarray(u, dtype=np.float64)
        AB = torch.zeros((1, 1, 7, 10), dtype=torch.float64, device="cuda")
        me.linearize(self._dyn(dt), ge.to_dev(x[None]), ge.to_dev(u[None, None]), AB)
        ab = AB[0, 0].cpu().numpy()
        return ab[:, :7].copy(), ab[:, 7:].copy()


# ====================================================================================================== 6-DoF
@dataclass
class Rocket6DoFConfig:
    """rocket_6dof.py:35-89 (the fields the per-step numeric core reads)."""

    m_dry: float = 1.0
    m_wet: float = 2.0
    J_B: Optional[np.ndarray] = None  # noqa: N815
    I_sp: float = 30.0
    g0: float = 1.0
    T_min: float = 1.5
    T_max: float = 6.5
    r_T_B: Optional[np.ndarray] = None  # noqa: N815
    g_I: Optional[np.ndarray] = None  # noqa: N815
    default_dt: float = 0.1
    use_rk4: bool = False  # True: the reference's "rk4" = 4 Euler sub-steps (nominal_mpc.py:205-236)

    def __post_init__(self):
        if self.J_B is None:
            self.J_B = np.diag([0.02, 1.0, 1.0]) * 0.168
        if self.r_T_B is None:
            self.r_T_B = np.array([-0.25, 0.0, 0.0])
        if self.g_I is None:
            self.g_I = np.array([-1.0, 0.0, 0.0])

    @classmethod
    def szmuk_defaults(cls) -> "Rocket6DoFConfig":
        return cls()


@dataclass
class Rocket6DoFParams:
    """The parameter record the reference's 6-DoF prediction model reads (nominal_mpc.py:170-174)."""

    g_I: np.ndarray  # noqa: N815
    J_B: np.ndarray  # noqa: N815
    r_T_B: np.ndarray  # noqa: N815
    I_sp: float = 30.0
    g0: float = 1.0
    m_dry: float = 1.0
    m_wet: float = 2.0

    @property
    def alpha(self) -> float:
        return 1.0 / (self.I_sp * self.g0)


class Rocket6DoFDynamics:
    """rocket_6dof.py:92-420 interface on the in-repo statement of the model (nominal_mpc.py:163-236): state
    [m, r_I(3), v_I(3), q_BI(4), w_B(3)], control T_B(3); `step` renormalises the quaternion (rocket_6dof.py:370-387);
    `linearize` is the forward-difference rule the OSQP controller applies to `step` (osqp_rti.py:374-401)."""

    IDX_MASS = 0
    IDX_POS = slice(1, 4)
    IDX_VEL = slice(4, 7)
    IDX_QUAT = slice(7, 11)
    IDX_OMEGA = slice(11, 14)
    N_STATE = 14
    N_CONTROL = 3

    def __init__(self, config: Optional[Rocket6DoFConfig] = None):
        self.config = config or Rocket6DoFConfig.szmuk_defaults()
        c = self.config
        self._params = Rocket6DoFParams(g_I=np.array(c.g_I, dtype=np.float64), J_B=np.array(c.J_B, dtype=np.float64),
                                        r_T_B=np.array(c.r_T_B, dtype=np.float64), I_sp=c.I_sp, g0=c.g0, m_dry=c.m_dry,
                                        m_wet=c.m_wet)

    @property
    def params(self) -> Rocket6DoFParams:
        return self._params

    @property
    def n_state(self) -> int:
        return self.N_STATE

    @property
    def n_control(self) -> int:
        return self.N_CONTROL

    def create_initial_state(self, altitude: float = 10.0, mass: Optional[float] = None, velocity=None, lateral=None):
        x = np.zeros(14)
        x[0] = self.config.m_wet if mass is None else mass
        x[1] = altitude
        if lateral is not None:
            x[2:4] = lateral
        if velocity is not None:
            x[4:7] = velocity
        x[7] = 1.0
        return x

    def dyn_struct(self, dt=None):
        from . import _lib
        p = self._params
        d = _lib.Dyn6()
        d.alpha = float(p.alpha)
        Jinv = np.linalg.inv(p.J_B)
        for i in range(3):
            d.g[i] = float(p.g_I[i])
            d.rT[i] = float(p.r_T_B[i])
        for i in range(9):
            d.J[i] = float(p.J_B.reshape(-1)[i])
            d.Jinv[i] = float(Jinv.reshape(-1)[i])
        d.dt = float(self.config.default_dt if dt is None else dt)
        d.substeps = 4 if self.config.use_rk4 else 1
        d.normalize_quat = 1
        return d

    def step_batch(self, X, U, dt=None):
        """(B,14),(B,3) -> (B,14)."""
        import ctypes as C
        from .gp import engine as ge
        torch = ge.torch_()
        dev_in = hasattr(X, "is_cuda")
        Xd, Ud = ge.to_dev(X).contiguous(), ge.to_dev(U).contiguous()
        Xn = torch.empty_like(Xd)
        ge.call("gpmpc_step_6dof", C.byref(self.dyn_struct(dt)), Xd.shape[0], ge.ptr(Xd), ge.ptr(Ud), ge.ptr(Xn),
                ge.stream())
        return Xn if dev_in else Xn.cpu().numpy()

    def step(self, x, u, dt=None):
        return self.step_batch(np.asarray(x, dtype=np.float64)[None], np.asarray(u, dtype=np.float64)[None], dt)[0]

    def f_discrete(self, x, u, dt=None):
        return self.step(x, u, dt)

    def linearize_batch(self, X, U, dt=None):
        """(B,14),(B,3) -> AB (B,14,17) = [A_d | B_d] on the device."""
        import ctypes as C
        from .gp import engine as ge
        torch = ge.torch_()
        Xd, Ud = ge.to_dev(X).contiguous(), ge.to_dev(U).contiguous()
        AB = torch.empty((Xd.shape[0], 14, 17), dtype=torch.float64, device=Xd.device)
        ge.call("gpmpc_linearize_6dof", C.byref(self.dyn_struct(dt)), Xd.shape[0], ge.ptr(Xd), ge.ptr(Ud), ge.ptr(AB),
                ge.stream())
        return AB

    def linearize(self, x, u, dt=None):
        ab = self.linearize_batch(np.asarray(x, dtype=np.float64)[None], np.asarray(u, dtype=np.float64)[None], dt)
        ab = ab[0].cpu().numpy()
        return ab[:, :14].copy(), ab[:, 14:].copy()

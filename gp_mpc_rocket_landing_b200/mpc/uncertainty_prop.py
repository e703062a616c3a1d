"""Horizon uncertainty propagation with the reference's API (src/mpc/uncertainty_prop.py:34-177), 7-state model.

`propagate` = GP-augmented mean rollout + GP variance (blocked DMMA TRSM) + Sigma_{k+1} = A Sigma A' + Q_gp, all on
the device (`gpmpc_rollout_3dof`, `gpmpc_trsm_sumsq`, `gpmpc_propagate_cov_3dof`).  Only the "linear" method is on
the hot path; "unscented" / "monte_carlo" (uncertainty_prop.py:179-319) are SURVEY 8(f) "next" rows and raise.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Any, Optional, Tuple

import numpy as np

from .. import _lib
from ..gp import engine as ge
from . import engine as me


@dataclass
class PropagatedUncertainty:
    """uncertainty_prop.py:34-55."""

    means: Any
    covariances: Any

    def get_std(self, k: int):
        return np.sqrt(np.diag(self.covariances[k]))

    def get_confidence_bounds(self, k: int, confidence: float = 0.95) -> Tuple[Any, Any]:
        from scipy.stats import norm
        kappa = norm.ppf((1 + confidence) / 2)
        std = self.get_std(k)
        return self.means[k] - kappa * std, self.means[k] + kappa * std


class UncertaintyPropagator:
    """uncertainty_prop.py:58-177 (n_x = 7 specialisation; gp_model = Simple3DoFGP with an exact GP)."""

    def __init__(self, dynamics, gp_model, method: str = "linear"):
        self.dynamics = dynamics
        self.gp = gp_model
        self.method = method
        self.n_x, self.n_u = 7, 3

    def _dyn(self, dt):
        p = self.dynamics.params
        g_vec = getattr(p, "g_vec", None)
        if g_vec is None:
            g_vec = p.g_I
        return me.make_dyn(p.alpha, np.asarray(g_vec, dtype=np.float64), dt)

    def propagate_batch(self, X0, U, Sigma_0: Optional[float] = None, dt: float = 0.1):
        """X0 (B,7), U (B,N,3) -> means (B,N+1,7), covs (B,N+1,7,7), dmean (B,N,3), dvar (B,N,3) as CUDA tensors.
        Sigma_0: scalar s for Sigma_0 = s*I (default 1e-6, uncertainty_prop.py:133-134)."""
        if self.method != "linear":
            raise _lib.GpmpcError(f"propagation method {self.method!r} is not on the device hot path (linear only)")
        torch = ge.torch_()
        x0, Ud = ge.to_dev(X0), ge.to_dev(U)
        B, N = Ud.shape[0], Ud.shape[1]
        model = self.gp.device_model() if self.gp is not None else None
        dyn = self._dyn(dt)
        f64 = dict(dtype=torch.float64, device=x0.device)
        X = torch.zeros((B, N + 1, 7), **f64)
        covs = torch.zeros((B, N + 1, 7, 7), **f64)
        dmean = torch.zeros((B, N, 3), **f64)
        q_pad = ge.pad_to(B * N, 128)
        dvar = torch.zeros((q_pad, 3), **f64)
        s0 = 1e-6 if Sigma_0 is None else float(Sigma_0)
        if model is None:
            me.rollout(None, dyn, x0, Ud, X)
            me.propagate_cov(dyn, X, Ud, None, covs, sigma0=s0)
            return X, covs, dmean, dvar[: B * N].reshape(B, N, 3)
        f = model.f
        Ks = torch.zeros((q_pad, f.n_pad), **f64)
        me.rollout(model.struct(), dyn, x0, Ud, X, dmean, Ks)
        sumsq = f.sumsq_inplace(Ks)
        ge.call("gpmpc_predict_var", ge.ptr(sumsq), None, B * N, 3, f.sv, 1e-10, ge.ptr(model.ystd), ge.ptr(dvar), 3, 0,
                ge.stream())
        me.propagate_cov(dyn, X, Ud, dvar, covs, sigma0=s0)
        return X, covs, dmean, dvar[: B * N].reshape(B, N, 3)

    def propagate(self, x0, U, Sigma_0=None, dt: float = 0.1) -> PropagatedUncertainty:
        s0 = None
        if Sigma_0 is not None:
            S = np.asarray(Sigma_0, dtype=np.float64)
            if S.ndim == 2 and not np.allclose(S, S[0, 0] * np.eye(S.shape[0])):
                raise _lib.GpmpcError("the device covariance chain starts from Sigma_0 = s*I (the reference's only use)")
            s0 = float(S[0, 0]) if S.ndim == 2 else float(S)
        X, covs, _, _ = self.propagate_batch(np.asarray(x0, dtype=np.float64)[None], np.asarray(U, dtype=np.float64)[None],
                                             s0, dt)
        return PropagatedUncertainty(means=X[0].cpu().numpy(), covariances=covs[0].cpu().numpy())

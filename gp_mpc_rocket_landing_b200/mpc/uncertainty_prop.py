"""Horizon uncertainty propagation with the reference's API (src/mpc/uncertainty_prop.py:34-177), 7-state model.

`propagate` = GP-augmented mean rollout + GP variance (blocked DMMA TRSM) + Sigma_{k+1} = A Sigma A' + Q_gp, all on
the device (`gpmpc_rollout_3dof`, `gpmpc_trsm_sumsq`, `gpmpc_propagate_cov_3dof`).  "unscented"
(uncertainty_prop.py:179-264) runs stage by stage: sigma points, their images through the N = 1 rollout kernel (step +
GP mean), the GP variance at the centre, recombination (`gpmpc_ut_sigma_3dof`, `gpmpc_ut_combine_3dof`).
"monte_carlo" (:266-319): S particles per scenario, one stage at a time -- the N = 1 rollout kernel images all B*S particles
(step + GP mean) and writes their cross-covariance rows, one triangular GEMM gives the B*S variances, `gpmpc_mc_combine_3dof`
draws from the GP posterior and forms the sample mean / covariance.  The reference draws from the global numpy RNG; here the
initial particles and the standard-normal draws are inputs (for parity tests) or come from torch's Philox generator.
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
        if self.method == "unscented":
            return self._propagate_unscented_batch(X0, U, Sigma_0, dt)
        if self.method == "monte_carlo":
            return self._propagate_monte_carlo_batch(X0, U, Sigma_0, dt)
        if self.method != "linear":
            raise _lib.GpmpcError(f"unknown propagation method {self.method!r} (linear, unscented, monte_carlo)")
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

    N_REFERENCE_STATES = 14   # uncertainty_prop.py:86: the reference class is written for 14 states; this fixes lambda

    def _propagate_unscented_batch(self, X0, U, Sigma_0, dt):
        """uncertainty_prop.py:179-264 for B scenarios at once; same return as the linear method."""
        torch = ge.torch_()
        x0, Ud = ge.to_dev(X0), ge.to_dev(U)
        B, N = Ud.shape[0], Ud.shape[1]
        model = self.gp.device_model() if self.gp is not None else None
        dyn = self._dyn(dt)
        f64 = dict(dtype=torch.float64, device=x0.device)
        n = self.N_REFERENCE_STATES
        a, beta, kappa = 1e-3, 2, 0
        lam = a**2 * (n + kappa) - n
        wm0 = lam / (n + lam)
        wc0 = lam / (n + lam) + (1 - a**2 + beta)
        wi = 1 / (2 * (n + lam))
        s0 = 1e-6 if Sigma_0 is None else float(Sigma_0)
        means = torch.zeros((B, N + 1, 7), **f64)
        covs = torch.zeros((B, N + 1, 7, 7), **f64)
        dmean = torch.zeros((B, N, 3), **f64)
        dvar = torch.zeros((B, N, 3), **f64)
        means[:, 0] = x0
        covs[:, 0] = torch.eye(7, **f64) * s0
        SP = torch.empty((B * 15, 7), **f64)
        Usp = torch.empty((B * 15, 1, 3), **f64)
        Xn = torch.empty((B * 15, 2, 7), **f64)
        dm = torch.zeros((B * 15, 1, 3), **f64)
        info = torch.zeros((1,), dtype=torch.int32, device=x0.device)
        xk, Sk = x0.clone(), covs[:, 0].clone()
        xn, Sn = torch.empty_like(xk), torch.empty_like(Sk)
        if model is not None:
            f = model.f
            Ks = torch.zeros((ge.pad_to(B * 15, 128), f.n_pad), **f64)
            Kc = torch.zeros((ge.pad_to(B, 128), f.n_pad), **f64)
            var = torch.zeros((Kc.shape[0], 3), **f64)
            var_work = torch.empty(((f.n_pad // 64) * Kc.shape[0],), **f64)
            ms = model.struct()
        for k in range(N):
            Uk = Ud[:, k].contiguous()
            ge.call("gpmpc_ut_sigma_3dof", B, n + lam, ge.ptr(xk), ge.ptr(Sk), ge.ptr(Uk), Uk.stride(0), ge.ptr(SP),
                    ge.ptr(Usp), ge.ptr(info), ge.stream())
            if model is None:
                me.rollout(None, dyn, SP, Usp, Xn)
                vk = None
            else:
                me.rollout(ms, dyn, SP, Usp, Xn, dm, Ks)
                Kc[:B] = Ks[0:B * 15:15]                     # cross-covariance rows of the centre points
                sumsq = f.sumsq_linv(Kc, var_work)            # one triangular GEMM against the explicit inverse factor
                ge.call("gpmpc_predict_var", ge.ptr(sumsq), None, B, 3, f.sv, 1e-10, ge.ptr(model.ystd), ge.ptr(var), 3, 0,
                        ge.stream())
                vk = var
                dmean[:, k] = dm[0:B * 15:15, 0]
                dvar[:, k] = var[:B]
            ge.call("gpmpc_ut_combine_3dof", B, n - 7, wm0, wc0, wi, ge.ptr(Xn), ge.ptr(vk), float(dt), ge.ptr(xn),
                    ge.ptr(Sn), ge.stream())
            means[:, k + 1] = xn
            covs[:, k + 1] = Sn
            xk, xn = xn, xk
            Sk, Sn = Sn, Sk
        bad = int(info.item())
        if bad:
            raise np.linalg.LinAlgError(f"unscented propagation: (n + lambda) Sigma + 1e-10 I of scenario {bad - 1} is not "
                                        "positive definite")
        return means, covs, dmean, dvar

    def _propagate_monte_carlo_batch(self, X0, U, Sigma_0, dt, n_samples: int = 100, particles0=None, xi=None, seed=None):
        """uncertainty_prop.py:266-319 for B scenarios at once; same return as the linear method (dmean / dvar: the GP mean
        and variance averaged over the particles).  particles0 (B,S,7), xi (B,N,S,3): injected randomness (the reference's
        draws at :287 and :305); default: x0 + sqrt(Sigma_0) z and standard normals from a Philox generator seeded with
        `seed` -- a documented deviation from the reference's global numpy RNG, the moments agree within Monte Carlo error."""
        torch = ge.torch_()
        x0, Ud = ge.to_dev(X0), ge.to_dev(U)
        B, N = Ud.shape[0], Ud.shape[1]
        model = self.gp.device_model() if self.gp is not None else None
        dyn = self._dyn(dt)
        f64 = dict(dtype=torch.float64, device=x0.device)
        s0 = 1e-6 if Sigma_0 is None else float(Sigma_0)
        gen = None
        if particles0 is None or xi is None:
            gen = torch.Generator(device=x0.device)
            gen.manual_seed(0 if seed is None else int(seed))
        if particles0 is None:
            S = int(n_samples)
            part = x0[:, None, :] + np.sqrt(s0) * torch.randn((B, S, 7), generator=gen, **f64)
        else:
            part = ge.to_dev(particles0)
            S = part.shape[1]
        part = part.reshape(B * S, 7).contiguous()
        xi_d = ge.to_dev(xi) if xi is not None else None
        means = torch.zeros((B, N + 1, 7), **f64)
        covs = torch.zeros((B, N + 1, 7, 7), **f64)
        dmean = torch.zeros((B, N, 3), **f64)
        dvar = torch.zeros((B, N, 3), **f64)
        means[:, 0] = x0
        covs[:, 0] = torch.eye(7, **f64) * s0
        Q = B * S
        q_pad = ge.pad_to(Q, 128)
        Xn = torch.empty((Q, 2, 7), **f64)
        Usp = torch.empty((Q, 1, 3), **f64)
        dm = torch.zeros((Q, 1, 3), **f64)
        var = torch.zeros((q_pad, 3), **f64)
        nxt = torch.empty((Q, 7), **f64)
        mk = torch.empty((B, 7), **f64)
        Sk = torch.empty((B, 7, 7), **f64)
        if model is not None:
            f = model.f
            Ks = torch.zeros((q_pad, f.n_pad), **f64)
            var_work = torch.empty(((f.n_pad // 64) * q_pad,), **f64)
            ms = model.struct()
        for k in range(N):
            Usp[:, 0] = Ud[:, k].repeat_interleave(S, dim=0)
            if model is None:
                me.rollout(None, dyn, part, Usp, Xn)
            else:
                me.rollout(ms, dyn, part, Usp, Xn, dm, Ks)
                sumsq = f.sumsq_linv(Ks, var_work)
                ge.call("gpmpc_predict_var", ge.ptr(sumsq), None, Q, 3, f.sv, 1e-10, ge.ptr(model.ystd), ge.ptr(var), 3, 0,
                        ge.stream())
                dmean[:, k] = dm[:, 0].reshape(B, S, 3).mean(dim=1)
                dvar[:, k] = var[:Q].reshape(B, S, 3).mean(dim=1)
            z = xi_d[:, k].contiguous() if xi_d is not None else torch.randn((B, S, 3), generator=gen, **f64)
            ge.call("gpmpc_mc_combine_3dof", B, S, ge.ptr(Xn), ge.ptr(var), ge.ptr(z), float(dt), ge.ptr(nxt), ge.ptr(mk),
                    ge.ptr(Sk), ge.stream())
            means[:, k + 1] = mk
            covs[:, k + 1] = Sk
            part, nxt = nxt, part
        return means, covs, dmean, dvar

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


class ConstraintTightening:
    """uncertainty_prop.py:322-416: Gaussian chance-constraint back-offs kappa * sqrt(g' Sigma g), kappa = Phi^{-1}(p)."""

    def __init__(self, confidence: float = 0.95):
        from scipy.stats import norm
        self.confidence = confidence
        self.kappa = norm.ppf(confidence)

    def tighten_linear_constraint(self, a, b: float, mu, Sigma) -> float:  # noqa: ARG002 (mu unused, as in the reference)
        a = np.asarray(a, dtype=np.float64)
        return b + self.kappa * np.sqrt(a @ np.asarray(Sigma) @ a)

    def tighten_quadratic_constraint(self, x_mu, x_Sigma, constraint_func, n_samples: int = 100) -> float:
        """Sampling estimate from the global numpy RNG (uncertainty_prop.py:367-387)."""
        samples = np.random.multivariate_normal(x_mu, x_Sigma, n_samples)
        g = np.array([constraint_func(s) for s in samples])
        return np.percentile(g, (1 - self.confidence) * 100)

    def compute_back_offs(self, uncertainty: PropagatedUncertainty, constraint_gradients) -> np.ndarray:
        N = len(uncertainty.means) - 1
        nc = len(constraint_gradients[0]) if constraint_gradients else 0
        out = np.zeros((N, nc))
        for k in range(min(N, len(constraint_gradients))):
            S = np.asarray(uncertainty.covariances[k])
            for j, g in enumerate(constraint_gradients[k]):
                g = np.asarray(g, dtype=np.float64)
                out[k, j] = self.kappa * np.sqrt(g @ S @ g)
        return out

    def compute_back_offs_batch(self, covs, gradients):
        """All scenarios and stages at once: covs (B, N+1, n, n) as returned by `propagate_batch` (CUDA tensor or numpy),
        gradients (J, n) shared by all stages -> (B, N, J) back-offs kappa * sqrt(g_j' Sigma_{b,k} g_j)."""
        if hasattr(covs, "is_cuda"):
            torch = ge.torch_()
            G = ge.to_dev(np.asarray(gradients, dtype=np.float64))
            q = torch.einsum("ji,bkil,jl->bkj", G, covs[:, :-1], G)
            return self.kappa * q.clamp_min(0.0).sqrt()
        G = np.asarray(gradients, dtype=np.float64)
        q = np.einsum("ji,bkil,jl->bkj", G, np.asarray(covs)[:, :-1], G)
        return self.kappa * np.sqrt(np.maximum(q, 0.0))


class TubeBasedRobustness:
    """uncertainty_prop.py:419-470: w_{k+1} = |A_k| w_k + d_max dt in the velocity (and, for 14 states, body-rate) rows."""

    def __init__(self, dynamics, d_max: float = 0.1):
        self.dynamics = dynamics
        self.d_max = d_max

    def compute_tube(self, X_nom, U_nom, dt: float) -> np.ndarray:
        X_nom, U_nom = np.asarray(X_nom, dtype=np.float64), np.asarray(U_nom, dtype=np.float64)
        N, n_x = len(U_nom), X_nom.shape[1]
        widths = np.zeros((N + 1, n_x))
        w = np.zeros(n_x)
        for k in range(N):
            A_d, _ = self.dynamics.linearize(X_nom[k], U_nom[k], dt=dt)
            w = np.abs(np.asarray(A_d)) @ w
            w[4:7] += self.d_max * dt
            w[11:14] += self.d_max * dt     # empty slice for the 7-state model
            widths[k + 1] = w
        return widths

    # ---- batched device forms (7-state model) -----------------------------------------------------------------------
    def _device_tube(self, X_nom, U_nom, dt, W):
        torch = ge.torch_()
        X, U = ge.to_dev(X_nom), ge.to_dev(U_nom)
        B, N = U.shape[0], U.shape[1]
        p = self.dynamics.params
        g_vec = getattr(p, "g_vec", None)
        dyn = me.make_dyn(p.alpha, np.asarray(g_vec if g_vec is not None else p.g_I, dtype=np.float64), dt)
        Xp = X[:, :N].reshape(B * N, 7).contiguous()
        Up = U.reshape(B * N, 3).contiguous()
        AB = torch.empty((B * N, 7, 10), dtype=torch.float64, device=X.device)
        me.linearize(dyn, Xp, Up, AB)
        tubes = torch.empty((B, N + 1, 7), dtype=torch.float64, device=X.device)
        ge.call("gpmpc_propagate_tube", 7, B, N, ge.ptr(AB), 10, 70, ge.ptr(W), ge.ptr(tubes), None, ge.stream())
        return tubes

    def compute_tube_batch(self, X_nom, U_nom, dt: float):
        """`compute_tube` for B trajectories at once: X_nom (B,N+1,7), U_nom (B,N,3) -> (B,N+1,7) CUDA tensor
        (Jacobians by `gpmpc_linearize_3dof`, recursion by `gpmpc_propagate_tube`)."""
        torch = ge.torch_()
        U = ge.to_dev(U_nom)
        W = torch.zeros((U.shape[0], U.shape[1], 7), dtype=torch.float64, device=U.device)
        W[:, :, 4:7] = self.d_max * dt
        return self._device_tube(X_nom, U, dt, W)

    def propagate_with_gp_batch(self, X_nom, U_nom, dt: float, gp_model, n_sigma: float = 2.0):
        """GP-variance tubes (safety/tube_mpc.py:118-183) for B trajectories: w_k = n_sigma * sqrt(var_k) in the velocity rows,
        variances from one batched `gp_model.predict_batch` over all (scenario, stage) points.
        -> (tubes (B,N+1,7), gp_vars (B,N,3)) CUDA tensors."""
        torch = ge.torch_()
        X, U = ge.to_dev(X_nom), ge.to_dev(U_nom)
        B, N = U.shape[0], U.shape[1]
        _, var = gp_model.predict_batch(X[:, :N].reshape(B * N, 7).contiguous(), U.reshape(B * N, 3).contiguous())
        var = ge.to_dev(var).reshape(B, N, 3)
        W = torch.zeros((B, N, 7), dtype=torch.float64, device=U.device)
        W[:, :, 4:7] = n_sigma * var.sqrt()
        return self._device_tube(X, U, dt, W), var

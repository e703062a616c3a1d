"""User-facing residual models with the reference's API (src/gp/structured_gp.py): `Simple3DoFGP` (3 outputs) and
`StructuredRocketGP` (2 x 3 outputs).  Storage is the reference's Python lists; features, fit and predict run on the
B200.  New: `predict_batch` on `Simple3DoFGP` (name taken from StructuredRocketGP.predict_batch, :270) and
`device_model()` which hands the fitted factor to the fused GP-MPC horizon kernels.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Any, Dict, Optional, Tuple

import numpy as np

from .exact_gp import MultiOutputExactGP
from .features import AtmosphereModel, CombinedFeatureExtractor, Simple3DoFFeatureExtractor
from .kernels import SquaredExponentialARD, _is_dev


@dataclass
class StructuredGPConfig:
    """structured_gp.py:42-63."""

    n_inducing: int = 50
    noise_variance: float = 1e-4
    use_sparse: bool = True
    reference_velocity: float = 10.0
    include_altitude: bool = True
    include_density: bool = True
    signal_variance: float = 0.1
    lengthscales_translational: Optional[Any] = None
    lengthscales_rotational: Optional[Any] = None
    max_data_points: int = 1000
    novelty_threshold: float = 0.1


def _make_multi(input_dim, use_sparse, n_inducing, noise_variance):
    if use_sparse:
        from .sparse_gp import MultiOutputSparseGP
        return MultiOutputSparseGP(input_dim=input_dim, output_dim=3, n_inducing=n_inducing,
                                   noise_variance=noise_variance)
    return MultiOutputExactGP(input_dim=input_dim, output_dim=3, noise_variance=noise_variance)


class StructuredRocketGP:
    """structured_gp.py:66-411."""

    def __init__(self, config: Optional[StructuredGPConfig] = None):
        self.config = config or StructuredGPConfig()
        self.feature_extractor = CombinedFeatureExtractor(atmosphere=AtmosphereModel(),
                                                          reference_velocity=self.config.reference_velocity)
        n_feat_v = self.feature_extractor.n_features_translational
        n_feat_omega = self.feature_extractor.n_features_rotational
        # structured_gp.py:120-130 builds two kernels it never uses (noqa: F841); kept for attribute parity only
        ls_v = np.ones(n_feat_v) if self.config.lengthscales_translational is None else self.config.lengthscales_translational
        ls_o = np.ones(n_feat_omega) if self.config.lengthscales_rotational is None else self.config.lengthscales_rotational
        self._unused_kernels = (
            SquaredExponentialARD(n_feat_v, self.config.signal_variance, ls_v),
            SquaredExponentialARD(n_feat_omega, self.config.signal_variance, ls_o),
        )
        self.gp_v = _make_multi(n_feat_v, self.config.use_sparse, self.config.n_inducing, self.config.noise_variance)
        self.gp_omega = _make_multi(n_feat_omega, self.config.use_sparse, self.config.n_inducing,
                                    self.config.noise_variance)
        self.X_data: list = []
        self.U_data: list = []
        self.D_v_data: list = []
        self.D_omega_data: list = []
        self._is_fitted = False

    @property
    def n_data(self) -> int:
        return len(self.X_data)

    def add_data(self, X, U, D_v, D_omega) -> None:
        X, U, D_v, D_omega = (np.atleast_2d(a) for a in (X, U, D_v, D_omega))
        for i in range(X.shape[0]):
            self.X_data.append(X[i])
            self.U_data.append(U[i])
            self.D_v_data.append(D_v[i])
            self.D_omega_data.append(D_omega[i])
        if len(self.X_data) > self.config.max_data_points:  # FIFO cap, :197-204
            excess = len(self.X_data) - self.config.max_data_points
            self.X_data = self.X_data[excess:]
            self.U_data = self.U_data[excess:]
            self.D_v_data = self.D_v_data[excess:]
            self.D_omega_data = self.D_omega_data[excess:]
        self._is_fitted = False

    def fit(self) -> None:
        if self.n_data == 0:
            raise RuntimeError("No data to fit")
        X, U = np.array(self.X_data), np.array(self.U_data)
        Z_v = self.feature_extractor.extract_batch_translational(X, U)
        Z_omega = self.feature_extractor.extract_batch_rotational(X, U)
        self.gp_v.fit(Z_v, np.array(self.D_v_data))
        self.gp_omega.fit(Z_omega, np.array(self.D_omega_data))
        self._is_fitted = True

    def _ensure(self, n):
        """:245-253: lazy fit; with no data return the prior (zeros, signal_variance)."""
        if not self._is_fitted:
            if self.n_data > 0:
                self.fit()
            else:
                zero = np.zeros(3) if n is None else np.zeros((n, 3))
                prior = np.full(zero.shape, self.config.signal_variance)
                return zero, zero, prior, prior
        return None

    def predict(self, x, u) -> Tuple[Any, Any, Any, Any]:
        d = self._ensure(None)
        if d is not None:
            return d
        z_v = self.feature_extractor.extract_translational(x, u)
        z_omega = self.feature_extractor.extract_rotational(x, u)
        d_v_mean, d_v_var = self.gp_v.predict(z_v.reshape(1, -1))
        d_omega_mean, d_omega_var = self.gp_omega.predict(z_omega.reshape(1, -1))
        return d_v_mean.flatten(), d_omega_mean.flatten(), d_v_var.flatten(), d_omega_var.flatten()

    def predict_batch(self, X, U) -> Tuple[Any, Any, Any, Any]:
        d = self._ensure(X.shape[0])
        if d is not None:
            return d
        Z_v = self.feature_extractor.extract_batch_translational(X, U)
        Z_omega = self.feature_extractor.extract_batch_rotational(X, U)
        d_v_mean, d_v_var = self.gp_v.predict(Z_v)
        d_omega_mean, d_omega_var = self.gp_omega.predict(Z_omega)
        return d_v_mean, d_omega_mean, d_v_var, d_omega_var

    def device_mean_models(self):
        """(gp_v, gp_omega) posterior-mean model structs for `gpmpc_rollout_6dof` (None, None before any data).  The
        returned objects keep the device tensors alive; both GPs must have outputs sharing one factor (the default)."""
        if not self._is_fitted:
            if self.n_data == 0:
                return None, None
            self.fit()
        out = []
        for gp in (self.gp_v, self.gp_omega):
            shared = getattr(gp, "_shared", None)
            if shared is not None:
                out.append(shared.mean_struct())
            else:
                model = gp.device_model()
                if model is None:
                    raise RuntimeError("StructuredRocketGP: the fused horizon kernels need outputs sharing one factor")
                out.append(model.struct())
        return tuple(out)

    def get_full_residual(self, x, u):
        d_v_mean, d_omega_mean, d_v_var, d_omega_var = self.predict(x, u)
        d_mean, d_var = np.zeros(14), np.zeros(14)
        d_mean[4:7], d_mean[11:14] = d_v_mean, d_omega_mean
        d_var[4:7], d_var[11:14] = d_v_var, d_omega_var
        return d_mean, d_var

    def is_novel(self, x, u) -> bool:
        _, _, d_v_var, d_omega_var = self.predict(x, u)
        max_var = max(np.max(d_v_var), np.max(d_omega_var))
        return bool(max_var > self.config.novelty_threshold * self.config.signal_variance)

    def optimize_hyperparameters(self) -> Dict[str, Any]:
        return {"n_data": self.n_data, "is_fitted": self._is_fitted}

    def save(self, path: str) -> None:
        data = {"config": self.config, "X_data": self.X_data, "U_data": self.U_data, "D_v_data": self.D_v_data,
                "D_omega_data": self.D_omega_data, "is_fitted": self._is_fitted}
        np.save(path, data, allow_pickle=True)

    def load(self, path: str) -> None:
        data = np.load(path, allow_pickle=True).item()
        self.X_data, self.U_data = data["X_data"], data["U_data"]
        self.D_v_data, self.D_omega_data = data["D_v_data"], data["D_omega_data"]
        if data["is_fitted"]:
            self.fit()

    def __repr__(self) -> str:
        return (f"StructuredRocketGP(n_data={self.n_data}, n_inducing={self.config.n_inducing}, "
                f"fitted={self._is_fitted})")


class Simple3DoFGP:
    """structured_gp.py:414-496."""

    def __init__(self, n_inducing: int = 50, noise_variance: float = 1e-4, use_sparse: bool = True):
        self.feature_extractor = Simple3DoFFeatureExtractor()
        self.gp = _make_multi(self.feature_extractor.n_features, use_sparse, n_inducing, noise_variance)
        self.X_data: list = []
        self.U_data: list = []
        self.D_data: list = []
        self._is_fitted = False

    @property
    def n_data(self) -> int:
        return len(self.X_data)

    def add_data(self, X, U, D) -> None:
        X, U, D = np.atleast_2d(X), np.atleast_2d(U), np.atleast_2d(D)
        for i in range(X.shape[0]):
            self.X_data.append(X[i])
            self.U_data.append(U[i])
            self.D_data.append(D[i])
        self._is_fitted = False

    def fit(self) -> None:
        if self.n_data == 0:
            raise RuntimeError("No data")
        Z = self.feature_extractor.extract_batch(np.array(self.X_data), np.array(self.U_data))
        self.gp.fit(Z, np.array(self.D_data))
        self._is_fitted = True

    def predict(self, x, u):
        """-> (mean (3,), var (3,)); no data -> (zeros(3), 0.1 * ones(3)) (:485-489)."""
        if not self._is_fitted:
            if self.n_data > 0:
                self.fit()
            else:
                return np.zeros(3), np.ones(3) * 0.1
        z = self.feature_extractor.extract(x, u)
        mean, var = self.gp.predict(z.reshape(1, -1))
        return mean.flatten(), var.flatten()

    def predict_batch(self, X, U):
        """(B,7),(B,3) -> mean (B,3), var (B,3): batch = loop of `predict` (numpy or CUDA tensors)."""
        if not self._is_fitted:
            if self.n_data > 0:
                self.fit()
            else:
                B = X.shape[0]
                if _is_dev(X):
                    import torch
                    return (torch.zeros((B, 3), dtype=torch.float64, device=X.device),
                            torch.full((B, 3), 0.1, dtype=torch.float64, device=X.device))
                return np.zeros((B, 3)), np.ones((B, 3)) * 0.1
        Z = self.feature_extractor.extract_batch(X, U)
        return self.gp.predict(Z)

    def device_model(self):
        """Fitted exact-GP device state (shared factor, alpha) for the fused horizon kernels, or None."""
        if not self._is_fitted:
            if self.n_data == 0:
                return None
            self.fit()
        return getattr(self.gp, "device_model", None)

    def __repr__(self) -> str:
        return f"Simple3DoFGP(n_data={self.n_data}, fitted={self._is_fitted})"

"""Sparse (inducing-point) GP regression with the reference's API (src/gp/sparse_gp.py), fit and predict on the B200.

FITC (default) and VFE exactly as the reference states them, quirks included: the predictive mean
is ``K*u @ alpha`` with the whitened alpha (no ``L_uu^-T``, sparse_gp.py:281) and the VFE variance mixes whitened and
unwhitened factors (:293-299).  Inducing-point initialisation stays on the host and calls the same
``scipy.cluster.vq.kmeans2(minit="points")`` / ``np.random.choice`` as the reference (:122-148), so a seeded run picks
the same Z as the reference does.  ``MultiOutputSparseGP`` factorises once when its outputs share Z and
hyper-parameters (L_uu and L_B do not depend on the targets); the reference refits per output (:448-454).
"""
from __future__ import annotations

from typing import Optional, Tuple

import numpy as np
from scipy.cluster.vq import kmeans2

from . import engine as E
from .exact_gp import GPPrediction, _host
from .kernels import Kernel, SquaredExponentialARD, _is_dev


def _np2d(X):
    return np.atleast_2d(np.asarray(X, dtype=np.float64))


class SparseGP:
    """sparse_gp.py:36-389."""

    def __init__(self, kernel: Kernel, n_inducing: int = 50, noise_variance: float = 1e-4, method: str = "fitc",
                 inducing_points=None, jitter: float = 1e-6):
        self.kernel = kernel
        self.n_inducing = n_inducing
        self._noise_variance = noise_variance
        self.method = method
        self.jitter = jitter
        self._Z = inducing_points
        self.X_train = None
        self.y_train = None
        self.n_train = 0
        self._y_mean = 0.0
        self._y_std = 1.0
        self._model: Optional[E.SparseModelDev] = None
        self._col = 0
        self._log_marginal_likelihood = None

    # ---- reference attribute surface -------------------------------------------------------------------------------
    @property
    def inducing_points(self):
        return self._Z

    @property
    def noise_variance(self) -> float:
        return self._noise_variance

    @noise_variance.setter
    def noise_variance(self, value: float) -> None:
        self._noise_variance = value
        self._invalidate_cache()

    def _invalidate_cache(self) -> None:
        self._model = None
        self._log_marginal_likelihood = None

    @property
    def _L_uu(self):
        if self._model is None:
            return None
        M = self._model.f.M
        return _host(self._model.f.L_uu[:M, :M])

    @property
    def _L_B(self):
        if self._model is None:
            return None
        M = self._model.f.M
        return _host(self._model.f.L_B[:M, :M])

    @property
    def _alpha(self):
        if self._model is None:
            return None
        return _host(self._model.alpha[self._col, : self._model.f.M])

    @property
    def _Lambda_diag(self):
        if self._model is None or self._model.f.Lam is None:
            return None
        return _host(self._model.f.Lam[: self.n_train])

    def _initialize_inducing_points(self, X):
        """sparse_gp.py:122-148 (host side, same library calls and RNG stream as the reference)."""
        X = _np2d(_host(X) if _is_dev(X) else X)
        if X.shape[0] <= self.n_inducing:
            return X.copy()
        try:
            Z, _ = kmeans2(X, self.n_inducing, minit="points")
        except Exception:  # noqa: BLE001
            idx = np.random.choice(X.shape[0], self.n_inducing, replace=False)
            Z = X[idx].copy()
        return Z

    # ---- fit -------------------------------------------------------------------------------------------------------
    def _make_factor(self) -> E.SparseFactor:
        kind = getattr(self.kernel, "kind", None)
        if kind is None:
            raise NotImplementedError("SparseGP on the B200 needs an SE-ARD / Matern kernel (fused Gram kernel)")
        return E.SparseFactor(kind, self.kernel.lengthscales, self.kernel.signal_variance, self._noise_variance,
                              method=self.method, jitter=self.jitter)

    def _attach(self, model: E.SparseModelDev, col: int, X):
        self._model, self._col = model, col
        self.X_train = X
        self.n_train = model.f.n
        self.y_train = _host(model.Yn[col, : model.f.n])
        self._y_mean = float(model.ymean[col].item())
        self._y_std = float(model.ystd[col].item())
        self._log_marginal_likelihood = float(model.lml[col].item())

    def fit(self, X, y) -> "SparseGP":
        """sparse_gp.py:150-253."""
        Xk = X if _is_dev(X) else _np2d(X)
        yk = y.flatten() if _is_dev(y) else np.atleast_1d(np.asarray(y, dtype=np.float64)).flatten()
        if self._Z is None:
            self._Z = self._initialize_inducing_points(Xk)
        factor = self._make_factor().fit(Xk, self._Z)
        model = E.SparseModelDev(factor).fit_targets(yk)
        self._attach(model, 0, Xk)
        return self

    # ---- predict ---------------------------------------------------------------------------------------------------
    def predict(self, X, return_std: bool = True) -> GPPrediction:
        """sparse_gp.py:255-305."""
        if self._model is None:
            raise RuntimeError("Must call fit() before predict()")
        dev_in = _is_dev(X)
        Xq = X if dev_in else _np2d(X)
        if dev_in and Xq.dim() == 1:
            Xq = Xq.reshape(1, -1)
        mean, var = self._model.predict(E.to_dev(Xq), want_var=return_std, col=self._col)
        mean = mean[:, 0]
        if return_std:
            var = var[:, 0]
            std = var.sqrt()
        else:
            var = mean.new_zeros(mean.shape)
            std = mean.new_zeros(mean.shape)
        if dev_in:
            return GPPrediction(mean=mean, variance=var, std=std)
        return GPPrediction(mean=_host(mean), variance=_host(var), std=_host(std))

    def predict_f(self, X) -> Tuple[object, object]:
        pred = self.predict(X, return_std=True)
        return pred.mean, pred.variance

    @property
    def log_marginal_likelihood(self) -> float:
        if self._log_marginal_likelihood is None:
            raise RuntimeError("Must call fit() first")
        return self._log_marginal_likelihood

    def update(self, X_new, y_new) -> "SparseGP":
        """sparse_gp.py:328-353: refit on the augmented data (Z is kept)."""
        if self.X_train is None:
            return self.fit(X_new, y_new)
        X_old = _host(self.X_train) if _is_dev(self.X_train) else self.X_train
        y_denorm = self.y_train * self._y_std + self._y_mean
        X_all = np.vstack([X_old, _np2d(X_new)])
        y_all = np.concatenate([y_denorm, np.atleast_1d(np.asarray(y_new, dtype=np.float64))])
        return self.fit(X_all, y_all)

    def optimize_inducing_locations(self, n_iterations: int = 100, learning_rate: float = 0.01,  # noqa: ARG002
                                    verbose: bool = False) -> dict:  # noqa: ARG002
        """sparse_gp.py:355-380: re-run k-means on the current data and refit."""
        if self.X_train is not None:
            self._Z = self._initialize_inducing_points(self.X_train)
            y_denorm = self.y_train * self._y_std + self._y_mean
            self.fit(self.X_train, y_denorm)
        return {"n_iterations": 0, "method": "kmeans"}

    def __repr__(self) -> str:
        return (f"SparseGP(n_inducing={self.n_inducing}, method='{self.method}', "
                f"noise_variance={self._noise_variance:.6f}, n_train={self.n_train})")


def _same_sparse_hypers(gps) -> bool:
    g0 = gps[0]
    k0 = g0.kernel
    for g in gps[1:]:
        k = g.kernel
        if (type(k) is not type(k0) or g._noise_variance != g0._noise_variance or g.method != g0.method
                or g.jitter != g0.jitter or k.signal_variance != k0.signal_variance
                or not np.array_equal(k.lengthscales, k0.lengthscales)):
            return False
        if g._Z is None or g0._Z is None or not np.array_equal(g._Z, g0._Z):
            return False
    return True


class MultiOutputSparseGP:
    """sparse_gp.py:391-511."""

    def __init__(self, input_dim: int, output_dim: int, n_inducing: int = 50, noise_variance: float = 1e-4,
                 share_inducing: bool = True):
        self.input_dim = input_dim
        self.output_dim = output_dim
        self.n_inducing = n_inducing
        self.share_inducing = share_inducing
        self.gps = [SparseGP(SquaredExponentialARD(input_dim), n_inducing, noise_variance) for _ in range(output_dim)]
        self._shared: Optional[E.SparseModelDev] = None

    def fit(self, X, Y) -> "MultiOutputSparseGP":
        """sparse_gp.py:430-456 (Z is re-drawn on every fit when share_inducing, as in the reference)."""
        Xk = X if _is_dev(X) else _np2d(X)
        Yk = Y if _is_dev(Y) else _np2d(Y)
        if Yk.shape[1] != self.output_dim:
            Yk = Yk.T
        if self.share_inducing:
            Z = self.gps[0]._initialize_inducing_points(Xk)
            for gp in self.gps:
                gp._Z = Z.copy()
        self._shared = None
        if _same_sparse_hypers(self.gps):
            factor = self.gps[0]._make_factor().fit(Xk, self.gps[0]._Z)
            model = E.SparseModelDev(factor).fit_targets(Yk)
            for i, gp in enumerate(self.gps):
                gp._attach(model, i, Xk)
            self._shared = model
        else:
            for i, gp in enumerate(self.gps):
                gp.fit(Xk, Yk[:, i])
        return self

    def predict_device(self, Xq):
        """-> (mean (Q,P), var (Q,P)) CUDA tensors; one fused pass when the outputs share a factor."""
        if self._shared is not None and all(g._model is self._shared for g in self.gps):
            return self._shared.predict(Xq)
        torch = E.torch_()
        ms, vs = [], []
        for gp in self.gps:
            if gp._model is None:
                raise RuntimeError("Must call fit() before predict()")
            m, v = gp._model.predict(Xq, col=gp._col)
            ms.append(m)
            vs.append(v)
        return torch.cat(ms, dim=1), torch.cat(vs, dim=1)

    def predict(self, X) -> Tuple[object, object]:
        """sparse_gp.py:458-482."""
        dev_in = _is_dev(X)
        Xq = X if dev_in else _np2d(X)
        if dev_in and Xq.dim() == 1:
            Xq = Xq.reshape(1, -1)
        for gp in self.gps:
            if gp._model is None:
                raise RuntimeError("Must call fit() before predict()")
        mean, var = self.predict_device(E.to_dev(Xq))
        if dev_in:
            return mean, var
        return _host(mean), _host(var)

    def predict_f(self, X):
        return self.predict(X)

    def update(self, X_new, Y_new) -> "MultiOutputSparseGP":
        """sparse_gp.py:484-504: per-output refit on the augmented data (Z kept)."""
        Y_new = _np2d(Y_new)
        if Y_new.shape[1] != self.output_dim:
            Y_new = Y_new.T
        for i, gp in enumerate(self.gps):
            gp.update(X_new, Y_new[:, i])
        self._shared = None
        return self

    def __repr__(self) -> str:
        return (f"MultiOutputSparseGP(input_dim={self.input_dim}, output_dim={self.output_dim}, "
                f"n_inducing={self.n_inducing})")

"""`FastGPPredictor` family with the reference's API (src/gp/fast_gp.py), fit and predict on the B200.

These are the reference's "fast" predictors: SE-ARD GP **without** y-normalisation, noise and jitter folded into one
diagonal term, one variance shared by all outputs and floored at 0 (fast_gp.py:113-213).  `benchmark_gp_prediction`
(src/utils/profiler.py:487-530) times exactly `fit(X, Y)` + `predict(X_test)`.  Same device kernels as ExactGP.
"""
from __future__ import annotations

import time
from dataclasses import dataclass
from typing import List, Optional, Tuple

import numpy as np

from . import engine as E
from .kernels import _is_dev

HAS_NUMBA = False  # the reference's optional Numba kernel (fast_gp.py:466-489) has no role on the device


@dataclass
class FastGPConfig:
    """fast_gp.py:40-61."""

    jitter: float = 1e-6
    cache_kernel_matrix: bool = True
    cache_cholesky: bool = True
    use_sparse: bool = False
    n_inducing: int = 100
    batch_size: int = 100
    use_numba: bool = False


def _host(t):
    return t.cpu().numpy()


class FastGPPredictor:
    """fast_gp.py:64-270."""

    def __init__(self, config: Optional[FastGPConfig] = None):
        self.config = config or FastGPConfig()
        self._X = None
        self._Y = None
        self._n_train = 0
        self._n_outputs = 0
        self._lengthscales = None
        self._variance = 1.0
        self._noise = 0.01
        self._model: Optional[E.ExactModelDev] = None

    @property
    def _L(self):
        return None if self._model is None else _host(self._model.f.L[: self._n_train, : self._n_train])

    @property
    def _alpha(self):
        return None if self._model is None else _host(self._model.alpha[:, : self._n_train]).T.copy()

    def fit(self, X, Y, lengthscales=None, variance: float = 1.0, noise: float = 0.01) -> float:
        """fast_gp.py:113-169: K + (noise + jitter) I, one retry with + 1e-4 I; alpha = K^{-1} Y (no normalisation)."""
        t0 = time.perf_counter()
        X = np.ascontiguousarray(np.asarray(X, dtype=np.float64))
        Y = np.ascontiguousarray(np.asarray(Y, dtype=np.float64).reshape(len(X), -1))
        self._X, self._Y = X, Y
        self._n_train, self._n_outputs = X.shape[0], Y.shape[1]
        self._variance, self._noise = variance, noise
        self._lengthscales = (np.ones(X.shape[1]) if lengthscales is None
                              else np.ascontiguousarray(np.asarray(lengthscales, dtype=np.float64)))
        diag = self._noise + self.config.jitter
        f = E.ExactFactor(E.SE, self._lengthscales, self._variance, diag)
        try:
            f.fit(X, jitters=(0.0, 1e-4))   # K += 1e-4 I once on failure (:155-160), then LinAlgError
        except ValueError as e:
            raise np.linalg.LinAlgError(str(e)) from e
        self._model = E.ExactModelDev(f).fit_targets(Y, normalize=False)
        E.torch_().cuda.current_stream().synchronize()
        return (time.perf_counter() - t0) * 1000

    def predict_device(self, Xq, return_var: bool = True):
        """-> (mean (Q,m), var (Q,m) | None) CUDA tensors."""
        if self._model is None:
            raise RuntimeError("Must call fit() before predict()")
        mean, var = self._model.predict(E.to_dev(Xq), want_var=return_var, var_floor=0.0)
        return mean, var

    def predict(self, X_test, return_var: bool = True) -> Tuple[object, Optional[object]]:
        """fast_gp.py:171-213: var = max(variance - sum v^2, 0), the same for every output."""
        dev_in = _is_dev(X_test)
        Xq = X_test if dev_in else np.atleast_2d(np.asarray(X_test, dtype=np.float64))
        if dev_in and Xq.dim() == 1:
            Xq = Xq.reshape(1, -1)
        mean, var = self.predict_device(Xq, return_var)
        if dev_in:
            return mean, var
        return _host(mean), (None if var is None else _host(var))

    def predict_single(self, x) -> Tuple[np.ndarray, np.ndarray]:
        """fast_gp.py:215-228."""
        mean, var = self.predict(np.asarray(x, dtype=np.float64).reshape(1, -1), return_var=True)
        return mean[0], var[0]

    def _compute_kernel_matrix(self, X1, X2):
        """fast_gp.py:230-252 (device Gram, returned on the host)."""
        return _host(E.gram(E.SE, X1, X2, self._lengthscales, self._variance))

    def update(self, X_new, Y_new) -> float:
        """fast_gp.py:254-270: refit on the stacked data."""
        X_all = np.vstack([self._X, np.atleast_2d(X_new)])
        Y_all = np.vstack([self._Y, np.asarray(Y_new, dtype=np.float64).reshape(-1, self._n_outputs)])
        return self.fit(X_all, Y_all, self._lengthscales, self._variance, self._noise)


class CachedGPPredictor(FastGPPredictor):
    """fast_gp.py:273-335: FIFO cache of single-point queries (host-side bookkeeping, linear scan as in the reference)."""

    def __init__(self, config: Optional[FastGPConfig] = None, cache_size: int = 1000, cache_tolerance: float = 1e-4):
        super().__init__(config)
        self._cache_size = cache_size
        self._cache_tol = cache_tolerance
        self._cache: List[Tuple[np.ndarray, np.ndarray, np.ndarray]] = []
        self._cache_hits = 0
        self._cache_misses = 0

    def predict_single(self, x):
        x = np.asarray(x, dtype=np.float64)
        for xc, mc, vc in self._cache:
            if np.linalg.norm(x - xc) < self._cache_tol:
                self._cache_hits += 1
                return mc, vc
        self._cache_misses += 1
        mean, var = super().predict_single(x)
        if len(self._cache) >= self._cache_size:
            self._cache.pop(0)
        self._cache.append((x.copy(), mean, var))
        return mean, var

    def clear_cache(self) -> None:
        self._cache.clear()
        self._cache_hits = 0
        self._cache_misses = 0

    def get_cache_stats(self) -> dict:
        total = self._cache_hits + self._cache_misses
        return {"hits": self._cache_hits, "misses": self._cache_misses,
                "hit_rate": self._cache_hits / max(1, total), "cache_size": len(self._cache)}


class SparseGPPredictor:
    """fast_gp.py:338-462: FITC-style predictor with a random inducing subset (`np.random.choice`, same RNG stream as
    the reference) and Woodbury factor chol(K_mm + K_mn Lambda^{-1} K_nm + jitter I).  The reference forms
    Q_nn = diag(K_nm K_mm^{-1} K_mn) through an explicit inverse (:396); here it is |L_m^{-1} K_mn|^2 column-wise --
    the same quantity through the Cholesky factor the class already holds."""

    def __init__(self, n_inducing: int = 100, config: Optional[FastGPConfig] = None):
        self.n_inducing = n_inducing
        self.config = config or FastGPConfig()
        self._Z = None
        self._lengthscales = None
        self._variance = 1.0
        self._noise = 0.01
        self._f: Optional[E.SparseFactor] = None
        self._W = None       # mean weights, device (m_out, M_pad)
        self._n_outputs = 0

    @property
    def _Lm(self):
        return None if self._f is None else _host(self._f.L_uu[: self._f.M, : self._f.M])

    @property
    def _A(self):
        return None if self._f is None else _host(self._f.L_B[: self._f.M, : self._f.M])

    @property
    def _mean_weights(self):
        return None if self._W is None else _host(self._W[:, : self._f.M]).T.copy()

    def fit(self, X, Y, lengthscales=None, variance: float = 1.0, noise: float = 0.01) -> float:
        t0 = time.perf_counter()
        torch = E.torch_()
        X = np.asarray(X, dtype=np.float64)
        n = len(X)
        m = min(self.n_inducing, n)
        indices = np.random.choice(n, m, replace=False)
        self._Z = X[indices].copy()
        self._lengthscales = np.ones(X.shape[1]) if lengthscales is None else np.asarray(lengthscales, dtype=np.float64)
        self._variance, self._noise = variance, noise
        Y = np.asarray(Y, dtype=np.float64).reshape(n, -1)
        self._n_outputs = Y.shape[1]
        f = E.SparseFactor(E.SE, self._lengthscales, variance, noise, method="fitc", jitter=self.config.jitter)
        f.fit(X, self._Z)                                    # L_uu = chol(K_mm + jitter I), qff = Q_nn diagonal
        P, Mp, npad = self._n_outputs, f.M_pad, f.n_pad
        Yt = torch.zeros((P, npad), dtype=torch.float64, device="cuda")
        Yt[:, :n] = E.to_dev(Y).T
        f.Lam = torch.empty((npad,), dtype=torch.float64, device="cuda")
        Ks = torch.empty_like(f.Kuf)
        Ylam = torch.empty_like(Yt)
        E.call("gpmpc_fitc_scale", E.ptr(f.Kuf), npad, Mp, n, npad, E.ptr(f.qff), variance, noise, -1e300, E.ptr(Yt), P,
               E.ptr(f.Lam), E.ptr(Ks), npad, E.ptr(Ylam), E.stream())
        Am = f.Kuu.clone()                                   # K_mm (+ jitter) + K_mn Lambda^{-1} K_nm + jitter I  (:404-405)
        E.call("gpmpc_add_diag", E.ptr(Am), f.M, Mp, self.config.jitter, E.stream())
        E.call("gpmpc_dgemm_nt", E.ptr(Ks), npad, E.ptr(Ks), npad, E.ptr(Am), Mp, Mp, Mp, npad, 1.0, 1.0, 0, E.stream())
        E.call("gpmpc_add_diag", E.ptr(Am), f.M, Mp, self.config.jitter, E.stream())
        f.W_B, info = E.potrf(Am)
        if int(info.item()) != 0:
            raise np.linalg.LinAlgError("Matrix is not positive definite")
        f.L_B = Am
        cT = torch.zeros((Mp, P), dtype=torch.float64, device="cuda")
        zeros = torch.zeros((P,), dtype=torch.float64, device="cuda")
        ones = torch.ones((P,), dtype=torch.float64, device="cuda")
        E.call("gpmpc_predict_mean", E.ptr(f.Kuf), npad, f.M, n, E.ptr(Ylam), npad, P, E.ptr(zeros), E.ptr(ones),
               E.ptr(cT), P, 0, E.stream())
        self._W = f.solve_B(cT.T.contiguous())
        self._f, self._zeros, self._ones = f, zeros, ones
        torch.cuda.current_stream().synchronize()
        return (time.perf_counter() - t0) * 1000

    def predict(self, X_test, return_var: bool = True):
        """fast_gp.py:424-449: var = max(variance - |L_m^{-1} k|^2 + |L_A^{-1} k|^2, 0), shared by all outputs."""
        if self._f is None:
            raise RuntimeError("Must call fit() before predict()")
        torch = E.torch_()
        f = self._f
        dev_in = _is_dev(X_test)
        Xq = E.to_dev(X_test if dev_in else np.atleast_2d(np.asarray(X_test, dtype=np.float64)))
        Q, P = Xq.shape[0], self._n_outputs
        Xqs, sqq = E.scale_inputs(Xq, f.ls, DP=f.DP)
        Ks = E.gram_scaled(E.SE, Xqs, sqq, Q, f.Zs, f.sqz, f.M, f.sv)
        mean = torch.empty((Q, P), dtype=torch.float64, device="cuda")
        E.call("gpmpc_predict_mean", E.ptr(Ks), Ks.stride(0), Q, f.M, E.ptr(self._W), f.M_pad, P, E.ptr(self._zeros),
               E.ptr(self._ones), E.ptr(mean), P, 0, E.stream())
        var = None
        if return_var:
            Ks2 = Ks.clone()
            ssv, ssw = (torch.empty((Ks.shape[0],), dtype=torch.float64, device="cuda") for _ in range(2))
            E.call("gpmpc_trsm_sumsq", E.ptr(f.L_uu), f.M_pad, f.M_pad, E.ptr(f.W_uu), E.ptr(Ks), Ks.shape[0],
                   Ks.stride(0), E.ptr(ssv), E.stream())
            E.call("gpmpc_trsm_sumsq", E.ptr(f.L_B), f.M_pad, f.M_pad, E.ptr(f.W_B), E.ptr(Ks2), Ks2.shape[0],
                   Ks2.stride(0), E.ptr(ssw), E.stream())
            var = torch.empty((Q, P), dtype=torch.float64, device="cuda")
            E.call("gpmpc_predict_var", E.ptr(ssv), E.ptr(ssw), Q, P, f.sv, 0.0, E.ptr(self._ones), E.ptr(var), P, 0,
                   E.stream())
        if dev_in:
            return mean, var
        return _host(mean), (None if var is None else _host(var))

    def _compute_kernel(self, X1, X2):
        return _host(E.gram(E.SE, X1, X2, self._lengthscales, self._variance))


class NumbaGPPredictor(FastGPPredictor):
    """fast_gp.py:492-509: in the reference a Numba-jitted Gram; on the device it is the same fused Gram kernel."""


def create_fast_gp(gp_type: str = "exact", n_inducing: int = 100, use_cache: bool = True,  # noqa: ARG001
                   use_numba: bool = True):
    """fast_gp.py:512-539."""
    config = FastGPConfig(use_numba=use_numba and HAS_NUMBA)
    if gp_type == "sparse":
        return SparseGPPredictor(n_inducing, config)
    if gp_type == "cached":
        return CachedGPPredictor(config)
    if gp_type == "numba" and HAS_NUMBA:
        return NumbaGPPredictor(config)
    return FastGPPredictor(config)

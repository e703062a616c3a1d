"""Exact GP regression with the reference's API (src/gp/exact_gp.py), fit and predict on the B200.

``fit`` = fused Gram build + blocked DMMA Cholesky (jitter escalation as in exact_gp.py:163-175) + cho_solve + LML,
``predict`` = cross-Gram + mean GEMV + blocked triangular solve with fused column sum of squares -- all in
libgpmpc_b200.so.  ``MultiOutputExactGP`` factorises ONCE when its outputs share hyper-parameters (the reference
default, exact_gp.py:464-474, computes the same K and L three times; the result is bit-identical).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Callable, Optional, Tuple

import numpy as np

from . import engine as E
from .kernels import Kernel, SquaredExponentialARD, _is_dev


@dataclass
class GPPrediction:
    """exact_gp.py:33-44."""

    mean: object
    variance: object
    std: object

    @property
    def confidence_bounds(self):
        return self.mean - 1.96 * self.std, self.mean + 1.96 * self.std


def _host(t):
    return t.cpu().numpy()


def make_factor(kernel: Kernel, noise_variance: float) -> E.ExactFactor:
    kind = getattr(kernel, "kind", None)
    if kind is not None:
        return E.ExactFactor(kind, kernel.lengthscales, kernel.signal_variance, noise_variance)
    return E.ExactFactor(None, [1.0], 1.0, noise_variance, kernel=kernel)


class ExactGP:
    """exact_gp.py:47-421."""

    def __init__(self, kernel: Kernel, noise_variance: float = 1e-4,
                 mean_function: Optional[Callable] = None, normalize_y: bool = True):
        self.kernel = kernel
        self._noise_variance = noise_variance
        self.mean_function = mean_function
        self.normalize_y = normalize_y
        self.X_train = None
        self.y_train = None
        self.n_train = 0
        self._y_mean = 0.0
        self._y_std = 1.0
        self._model: Optional[E.ExactModelDev] = None   # device state (factor, alpha, normalisation)
        self._col = 0                                   # this GP's column in a shared multi-output model
        self._log_marginal_likelihood = None

    # ---- reference attribute surface -------------------------------------------------------------------------------
    @property
    def noise_variance(self) -> float:
        return self._noise_variance

    @noise_variance.setter
    def noise_variance(self, value: float) -> None:
        assert value > 0, "Noise variance must be positive"
        self._noise_variance = value
        self._invalidate_cache()

    def _invalidate_cache(self) -> None:
        self._model = None
        self._log_marginal_likelihood = None

    @property
    def _L(self):
        """Lower Cholesky factor (n, n) as numpy (device -> host copy), None before fit."""
        if self._model is None:
            return None
        n = self.n_train
        return _host(self._model.f.L[:n, :n])

    @property
    def _alpha(self):
        if self._model is None:
            return None
        return _host(self._model.alpha[self._col, : self.n_train])

    # ---- fit -----------------------------------------------------------------------------------------------------------
    def _attach(self, model: E.ExactModelDev, col: int, X, y_host_norm, ymean, ystd, lml):
        self._model, self._col = model, col
        self.X_train = X
        self.y_train = y_host_norm
        self.n_train = model.f.n
        self._y_mean, self._y_std = ymean, ystd
        self._log_marginal_likelihood = lml

    def fit(self, X, y) -> "ExactGP":
        """exact_gp.py:118-184."""
        Xk = X if _is_dev(X) else np.atleast_2d(np.asarray(X, dtype=np.float64))
        if _is_dev(Xk) and Xk.dim() == 1:
            Xk = Xk.reshape(1, -1)
        yk = y.flatten() if _is_dev(y) else np.atleast_1d(np.asarray(y, dtype=np.float64)).flatten()
        assert Xk.shape[0] == yk.shape[0], "X and y must have same number of samples"
        factor = make_factor(self.kernel, self._noise_variance).fit(Xk)
        pm = None
        if self.mean_function is not None:
            pm = np.asarray(self.mean_function(_host(Xk) if _is_dev(Xk) else Xk), dtype=np.float64)
        model = E.ExactModelDev(factor).fit_targets(yk, normalize=self.normalize_y, prior_mean=pm)
        self._attach(model, 0, X if _is_dev(X) else Xk, _host(model.Yn[0, : factor.n]), float(model.ymean[0].item()),
                     float(model.ystd[0].item()), float(model.lml[0].item()))
        return self

    def update(self, X_new, y_new) -> "ExactGP":
        """Add observations (the exact-GP counterpart of SparseGP.update, sparse_gp.py:328-353, which refits on the
        augmented data).  Same result as `fit` on the concatenated data -- targets are re-normalised, alpha and the LML
        recomputed -- but the factor is EXTENDED by the new rows (ExactFactor.append, O(n^2 m)) when the
        hyper-parameters are those of the cached factor; otherwise, or when the extension is not positive definite, this
        is a full refit."""
        if np.size(_host(y_new) if _is_dev(y_new) else y_new) == 0:
            return self                      # nothing to add
        if self._model is None:
            return self.fit(X_new, y_new)
        dev_x = _is_dev(self.X_train)
        Xn = X_new if _is_dev(X_new) else np.atleast_2d(np.asarray(X_new, dtype=np.float64))
        yn = _host(y_new).flatten() if _is_dev(y_new) else np.atleast_1d(np.asarray(y_new, dtype=np.float64)).flatten()
        # raw targets of the old rows: y_train holds (y - y_mean) / y_std - m(X) / y_std (exact_gp.py:141-154)
        y_old = self.y_train * self._y_std + self._y_mean
        if self.mean_function is not None:
            Xo = _host(self.X_train) if dev_x else self.X_train
            y_old = y_old + np.asarray(self.mean_function(Xo), dtype=np.float64).flatten()
        y_all = np.concatenate([y_old, yn])
        if dev_x:
            X_all = E.torch_().cat([self.X_train, E.to_dev(Xn)], dim=0)
        else:
            X_all = np.vstack([self.X_train, _host(Xn) if _is_dev(Xn) else Xn])
        f = self._model.f
        k = self.kernel
        same = (f.kind is not None and getattr(k, "kind", None) == f.kind and f.noise == float(self._noise_variance)
                and f.sv == float(k.signal_variance) and np.array_equal(f.ls_host, np.asarray(k.lengthscales, dtype=np.float64))
                and self._model.P == 1)
        if not same:
            return self.fit(X_all, y_all)
        try:
            f.append(Xn)
        except ValueError:
            return self.fit(X_all, y_all)
        pm = None
        if self.mean_function is not None:
            pm = np.asarray(self.mean_function(_host(X_all) if dev_x else X_all), dtype=np.float64)
        model = E.ExactModelDev(f).fit_targets(y_all, normalize=self.normalize_y, prior_mean=pm)
        self._attach(model, 0, X_all, _host(model.Yn[0, : f.n]), float(model.ymean[0].item()),
                     float(model.ystd[0].item()), float(model.lml[0].item()))
        return self

    @property
    def log_marginal_likelihood(self) -> float:
        if self._log_marginal_likelihood is None:
            raise RuntimeError("Must call fit() before accessing log_marginal_likelihood")
        return self._log_marginal_likelihood

    # ---- predict ----------------------------------------------------------------------------------------------------
    def predict(self, X, return_std: bool = True, return_cov: bool = False):
        """exact_gp.py:213-268."""
        if self._model is None:
            raise RuntimeError("Must call fit() before predict()")
        dev_in = _is_dev(X)
        Xq = X if dev_in else np.atleast_2d(np.asarray(X, dtype=np.float64))
        if dev_in and Xq.dim() == 1:
            Xq = Xq.reshape(1, -1)
        Xd = E.to_dev(Xq)
        Q = Xd.shape[0]
        out = (lambda t: t) if dev_in else _host
        if return_cov:
            mean, _, V = self._model.predict(Xd, want_var=False, col=self._col, keep_v=True, chunk=1 << 30)
            mean = self._add_prior_mean(mean[:, 0], Xq)
            Kss = E.to_dev(self.kernel._gram_dev(Xd))
            torch = E.torch_()
            qp = V.shape[0]
            VVt = torch.empty((qp, qp), dtype=torch.float64, device=V.device)
            E.call("gpmpc_dgemm_nt", E.ptr(V), V.stride(0), E.ptr(V), V.stride(0), E.ptr(VVt), qp, qp, qp, V.shape[1],
                   1.0, 0.0, 0, E.stream())
            cov = (Kss - VVt[:Q, :Q]) * self._y_std**2
            return out(mean), out(cov)
        mean, var = self._model.predict(Xd, want_var=return_std, col=self._col)
        mean = self._add_prior_mean(mean[:, 0], Xq)
        if return_std:
            var = var[:, 0]
            return GPPrediction(mean=out(mean), variance=out(var), std=out(var.sqrt()))
        m = out(mean)
        z = m * 0
        return GPPrediction(mean=m, variance=z, std=z.copy() if not dev_in else z.clone())

    def _add_prior_mean(self, mean, Xq):
        # (K* alpha + m(X*)/y_std) * y_std + y_mean: the device epilogue applied K* alpha * y_std + y_mean already
        if self.mean_function is None:
            return mean
        m = np.asarray(self.mean_function(_host(Xq) if _is_dev(Xq) else Xq), dtype=np.float64)
        return mean + E.to_dev(m / self._y_std) * self._y_std

    def predict_f(self, X) -> Tuple[object, object]:
        pred = self.predict(X, return_std=True)
        return pred.mean, pred.variance

    # ---- sampling (host RNG for seed parity with np.random; factorisation on the device) -------------------------------
    def _chol_host(self, K):
        torch = E.torch_()
        n = K.shape[0]
        npad = E.pad_to(n, 128)
        A = torch.eye(npad, dtype=torch.float64, device="cuda")
        A[:n, :n] = E.to_dev(K)
        _, info = E.potrf(A)
        if int(info.item()) != 0:
            raise np.linalg.LinAlgError("Matrix is not positive definite")
        return _host(A[:n, :n])

    def sample_prior(self, X, n_samples: int = 1, random_state: Optional[int] = None):
        """exact_gp.py:291-320."""
        if random_state is not None:
            np.random.seed(random_state)
        X = np.atleast_2d(X)
        K = self.kernel(X)
        K = K + 1e-10 * np.eye(K.shape[0])
        L = self._chol_host(K)
        samples = L @ np.random.randn(X.shape[0], n_samples)
        if self.mean_function is not None:
            samples = samples + self.mean_function(X)[:, None]
        return samples.T

    def sample_posterior(self, X, n_samples: int = 1, random_state: Optional[int] = None):
        """exact_gp.py:322-355."""
        if self._model is None:
            raise RuntimeError("Must call fit() before sample_posterior()")
        if random_state is not None:
            np.random.seed(random_state)
        X = np.atleast_2d(X)
        mean, cov = self.predict(X, return_cov=True)
        cov = cov + 1e-10 * np.eye(cov.shape[0])
        L = self._chol_host(cov)
        samples = mean[:, None] + L @ np.random.randn(X.shape[0], n_samples)
        return samples.T

    # ---- hyper-parameters -------------------------------------------------------------------------------------------------
    def _analytic_gradient_available(self) -> bool:
        k = self.kernel
        return (getattr(k, "kind", None) == E.SE and getattr(k, "learn_signal_variance", True)
                and getattr(k, "learn_lengthscales", True) and k.input_dim <= 32)

    def log_marginal_likelihood_gradient(self) -> np.ndarray:
        """d LML / d[kernel log-parameters..., log noise_variance] at the fitted hyper-parameters (SE-ARD), the
        parameter order of `optimize_hyperparameters` (exact_gp.py:380-384).  Device pass: gpmpc_lml_grad_se."""
        if self._model is None:
            raise RuntimeError("Must call fit() before log_marginal_likelihood_gradient()")
        if not self._analytic_gradient_available():
            raise ValueError("analytic LML gradients need an SE-ARD kernel with all parameters learnable")
        return _host(self._model.lml_gradient()[self._col])

    def optimize_hyperparameters(self, n_restarts: int = 5, verbose: bool = False,
                                 analytic_gradients: Optional[bool] = None) -> dict:
        """exact_gp.py:357-421: L-BFGS-B on the negative LML; every objective evaluation is a device refit.

        analytic_gradients (extension; None = when available): pass the device-computed gradient to L-BFGS-B instead of
        letting scipy difference the objective (the reference's behaviour, n_params + 1 refits per gradient)."""
        from scipy.optimize import minimize
        if self.X_train is None:
            raise RuntimeError("Must call fit() before optimize_hyperparameters()")
        use_grad = self._analytic_gradient_available() if analytic_gradients is None else bool(analytic_gradients)
        if use_grad and not self._analytic_gradient_available():
            raise ValueError("analytic LML gradients need an SE-ARD kernel with all parameters learnable")

        def objective(params):
            n_kernel = self.kernel.n_params
            self.kernel.set_params(params[:n_kernel])
            self._noise_variance = np.exp(params[n_kernel])
            try:
                self.fit(self.X_train, self.y_train * self._y_std + self._y_mean)
                if use_grad:
                    return -self._log_marginal_likelihood, -self.log_marginal_likelihood_gradient()
                return -self._log_marginal_likelihood
            except (np.linalg.LinAlgError, ValueError):
                return (np.inf, np.zeros(len(params))) if use_grad else np.inf

        initial_params = np.concatenate([self.kernel.get_params(), [np.log(self._noise_variance)]])
        best_result, best_nll = None, np.inf
        for restart in range(n_restarts):
            params0 = initial_params if restart == 0 else initial_params + 0.5 * np.random.randn(len(initial_params))
            try:
                result = minimize(objective, params0, method="L-BFGS-B", jac=True if use_grad else None,
                                  options={"maxiter": 100, "disp": True} if verbose else {"maxiter": 100})
                if result.fun < best_nll:
                    best_nll, best_result = result.fun, result
            except Exception as e:  # noqa: BLE001 (reference behaviour)
                if verbose:
                    print(f"Restart {restart} failed: {e}")
        if best_result is not None:
            objective(best_result.x)
        return {
            "success": best_result is not None and best_result.success,
            "log_marginal_likelihood": -best_nll if best_result else None,
            "n_iterations": best_result.nit if best_result else 0,
        }

    def __repr__(self) -> str:
        return f"ExactGP(kernel={self.kernel}, noise_variance={self._noise_variance:.6f}, n_train={self.n_train})"


def _same_hypers(gps) -> bool:
    k0, nv0 = gps[0].kernel, gps[0]._noise_variance
    for g in gps[1:]:
        k = g.kernel
        if type(k) is not type(k0) or g._noise_variance != nv0 or g.mean_function is not None:
            return False
        if getattr(k, "kind", None) is None and k is not k0:
            return False
        if getattr(k, "kind", None) is not None and not (
                k.signal_variance == k0.signal_variance and np.array_equal(k.lengthscales, k0.lengthscales)):
            return False
    return gps[0].mean_function is None and all(g.normalize_y == gps[0].normalize_y for g in gps)


class MultiOutputExactGP:
    """exact_gp.py:427-535: independent GPs per output (one shared factor while their hyper-parameters coincide)."""

    def __init__(self, input_dim: int, output_dim: int, kernel: Optional[Kernel] = None,
                 noise_variance: float = 1e-4, share_hyperparameters: bool = False):
        self.input_dim = input_dim
        self.output_dim = output_dim
        self.share_hyperparameters = share_hyperparameters
        self.gps: list[ExactGP] = []
        for _ in range(output_dim):
            if kernel is None:
                k = SquaredExponentialARD(input_dim)
            elif share_hyperparameters:
                k = kernel
            else:  # exact_gp.py:467-472: any passed kernel silently becomes SE-ARD with its parameter vector
                k = SquaredExponentialARD(input_dim)
                k.set_params(kernel.get_params())
            self.gps.append(ExactGP(k, noise_variance=noise_variance))
        self._shared: Optional[E.ExactModelDev] = None

    def fit(self, X, Y) -> "MultiOutputExactGP":
        dev_in = _is_dev(X)
        Xk = X if dev_in else np.atleast_2d(np.asarray(X, dtype=np.float64))
        Yk = Y if _is_dev(Y) else np.atleast_2d(np.asarray(Y, dtype=np.float64))
        if Yk.shape[1] != self.output_dim:
            if Yk.shape[0] == self.output_dim:
                Yk = Yk.T
            else:
                raise ValueError(f"Y must have {self.output_dim} columns")
        self._shared = None
        if _same_hypers(self.gps):
            g0 = self.gps[0]
            factor = make_factor(g0.kernel, g0._noise_variance).fit(Xk)
            model = E.ExactModelDev(factor).fit_targets(Yk, normalize=g0.normalize_y)
            Yn, ym, ys, lml = _host(model.Yn[:, : factor.n]), _host(model.ymean), _host(model.ystd), _host(model.lml)
            for i, gp in enumerate(self.gps):
                gp._attach(model, i, Xk, Yn[i], float(ym[i]), float(ys[i]), float(lml[i]))
            self._shared = model
        else:
            for i, gp in enumerate(self.gps):
                gp.fit(Xk, Yk[:, i])
        return self

    def predict(self, X) -> Tuple[object, object]:
        """-> mean (M, P), variance (M, P)."""
        dev_in = _is_dev(X)
        Xq = X if dev_in else np.atleast_2d(np.asarray(X, dtype=np.float64))
        if dev_in and Xq.dim() == 1:
            Xq = Xq.reshape(1, -1)
        if self._shared is not None and all(g._model is self._shared for g in self.gps):
            mean, var = self._shared.predict(E.to_dev(Xq))
            return (mean, var) if dev_in else (_host(mean), _host(var))
        M = Xq.shape[0]
        if dev_in:
            torch = E.torch_()
            means = torch.zeros((M, self.output_dim), dtype=torch.float64, device=Xq.device)
            variances = torch.zeros_like(means)
        else:
            means = np.zeros((M, self.output_dim))
            variances = np.zeros((M, self.output_dim))
        for i, gp in enumerate(self.gps):
            pred = gp.predict(Xq)
            means[:, i] = pred.mean
            variances[:, i] = pred.variance
        return means, variances

    def predict_f(self, X):
        return self.predict(X)

    @property
    def log_marginal_likelihood(self) -> float:
        return sum(gp.log_marginal_likelihood for gp in self.gps)

    @property
    def device_model(self) -> Optional[E.ExactModelDev]:
        """The shared (factor, alpha) device state consumed by the fused horizon kernels, if the outputs share one."""
        if self._shared is not None and all(g._model is self._shared for g in self.gps):
            return self._shared
        return None

    def __repr__(self) -> str:
        return f"MultiOutputExactGP(input_dim={self.input_dim}, output_dim={self.output_dim})"

"""Covariance functions with the reference's API (src/gp/kernels.py), evaluated on the B200.

Every class keeps the reference's constructor, properties and log-space parameter vector; ``__call__`` and
``gradients`` run the fused Gram kernel of libgpmpc_b200.so (csrc/gram.cu: DMMA cross term + exp / Matern epilogue,
same operation order as kernels.py:205-262).  numpy in -> numpy out; CUDA ``torch.float64`` tensors in -> CUDA
tensors out (no host round trip).  There is no CPU path.
"""
from __future__ import annotations

from abc import ABC, abstractmethod
from typing import List, Optional

import numpy as np

from . import engine as E


def _is_dev(a) -> bool:
    return hasattr(a, "is_cuda")


def _out(K, like):
    """Mirror the input kind: numpy caller gets numpy back."""
    return K if _is_dev(like) else K.cpu().numpy()


def _atleast_2d(X):
    if _is_dev(X):
        return X if X.dim() == 2 else X.reshape(1, -1)
    return np.atleast_2d(np.asarray(X, dtype=np.float64))


class Kernel(ABC):
    """kernels.py:33-124."""

    #: device kernel id (engine.SE / M32 / M52) when the Gram can be built by one fused launch, else None
    kind: Optional[int] = None

    @abstractmethod
    def __call__(self, X1, X2=None): ...

    @abstractmethod
    def diagonal(self, X): ...

    @abstractmethod
    def gradients(self, X1, X2=None) -> dict: ...

    @property
    @abstractmethod
    def n_params(self) -> int: ...

    @property
    @abstractmethod
    def param_names(self) -> List[str]: ...

    @abstractmethod
    def get_params(self): ...

    @abstractmethod
    def set_params(self, params) -> None: ...

    def __add__(self, other: "Kernel") -> "SumKernel":
        return SumKernel(self, other)

    def __mul__(self, other: "Kernel") -> "ProductKernel":
        return ProductKernel(self, other)

    # device-side evaluation used by the GP classes (always returns a CUDA tensor)
    def _gram_dev(self, X1, X2=None):
        return E.to_dev(self(E.to_dev(_atleast_2d(X1)), None if X2 is None else E.to_dev(_atleast_2d(X2))))


class _ARDKernel(Kernel):
    """Shared implementation of the three ARD kernels (one fused device launch each)."""

    def __init__(self, input_dim: int, signal_variance: float = 1.0, lengthscales=None):
        self.input_dim = input_dim
        self._signal_variance = signal_variance
        if lengthscales is None:
            self._lengthscales = np.ones(input_dim)
        else:
            self._lengthscales = np.asarray(lengthscales, dtype=np.float64).flatten()

    @property
    def signal_variance(self) -> float:
        return self._signal_variance

    @property
    def lengthscales(self):
        return self._lengthscales

    def __call__(self, X1, X2=None):
        X1 = _atleast_2d(X1)
        if X2 is not None:
            X2 = _atleast_2d(X2)
        K = E.gram(self.kind, X1, X2, self._lengthscales, float(self._signal_variance))
        return _out(K, X1)

    def diagonal(self, X):
        X = _atleast_2d(X)
        if _is_dev(X):
            return E.torch_().full((X.shape[0],), float(self._signal_variance), dtype=X.dtype, device=X.device)
        return np.full(X.shape[0], self._signal_variance)

    def gradients(self, X1, X2=None) -> dict:
        # kernels.py:551-558 / :644-650: the Matern classes only return the signal-variance gradient
        return {"log_signal_variance": self(X1, X2)}

    @property
    def n_params(self) -> int:
        return 1 + self.input_dim

    @property
    def param_names(self) -> List[str]:
        return ["log_signal_variance"] + [f"log_lengthscale_{i}" for i in range(self.input_dim)]

    def get_params(self):
        return np.concatenate([[np.log(self._signal_variance)], np.log(self._lengthscales)])

    def set_params(self, params) -> None:
        params = np.asarray(params, dtype=np.float64)
        self._signal_variance = np.exp(params[0])
        self._lengthscales = np.exp(params[1:])


class SquaredExponentialARD(_ARDKernel):
    """kernels.py:130-379: k = s2 exp(-0.5 sum_i (x_i - x'_i)^2 / l_i^2)."""

    kind = E.SE

    def __init__(self, input_dim: int, signal_variance: float = 1.0, lengthscales=None,
                 learn_signal_variance: bool = True, learn_lengthscales: bool = True):
        super().__init__(input_dim, signal_variance, lengthscales)
        assert len(self._lengthscales) == input_dim
        self.learn_signal_variance = learn_signal_variance
        self.learn_lengthscales = learn_lengthscales

    @property
    def signal_variance(self) -> float:
        return self._signal_variance

    @signal_variance.setter
    def signal_variance(self, value: float) -> None:
        assert value > 0, "Signal variance must be positive"
        self._signal_variance = value

    @property
    def lengthscales(self):
        return self._lengthscales

    @lengthscales.setter
    def lengthscales(self, value) -> None:
        value = np.asarray(value, dtype=np.float64).flatten()
        assert len(value) == self.input_dim
        assert np.all(value > 0), "Lengthscales must be positive"
        self._lengthscales = value

    def gradients(self, X1, X2=None) -> dict:
        """kernels.py:279-318: dK/dlog s2 = K ; dK/dlog l_i = K (x_i - x'_i)^2 / l_i^2."""
        X1 = _atleast_2d(X1)
        if X2 is not None:
            X2 = _atleast_2d(X2)
        grads = {}
        if self.learn_signal_variance:
            grads["log_signal_variance"] = self(X1, X2)
        if self.learn_lengthscales:
            for i in range(self.input_dim):
                G = E.gram_grad_lengthscale(X1, X2, self._lengthscales, float(self._signal_variance), i)
                grads[f"log_lengthscale_{i}"] = _out(G, X1)
        return grads

    @property
    def n_params(self) -> int:
        return (1 if self.learn_signal_variance else 0) + (self.input_dim if self.learn_lengthscales else 0)

    @property
    def param_names(self) -> List[str]:
        names = []
        if self.learn_signal_variance:
            names.append("log_signal_variance")
        if self.learn_lengthscales:
            names += [f"log_lengthscale_{i}" for i in range(self.input_dim)]
        return names

    def get_params(self):
        params = []
        if self.learn_signal_variance:
            params.append(np.log(self._signal_variance))
        if self.learn_lengthscales:
            params.extend(np.log(self._lengthscales))
        return np.array(params)

    def set_params(self, params) -> None:
        params = np.asarray(params, dtype=np.float64).flatten()
        idx = 0
        if self.learn_signal_variance:
            self._signal_variance = np.exp(params[idx])
            idx += 1
        if self.learn_lengthscales:
            self._lengthscales = np.exp(params[idx: idx + self.input_dim])

    def __repr__(self) -> str:
        return (f"SquaredExponentialARD(input_dim={self.input_dim}, signal_variance={self._signal_variance:.4f}, "
                f"lengthscales={self._lengthscales})")


RBF = SquaredExponentialARD
SE_ARD = SquaredExponentialARD


class Matern32(_ARDKernel):
    """kernels.py:482-576: k = s2 (1 + sqrt3 r) exp(-sqrt3 r)."""

    kind = E.M32

    def __repr__(self) -> str:
        return f"Matern32(input_dim={self.input_dim}, σ²={self._signal_variance:.4f})"


class Matern52(_ARDKernel):
    """kernels.py:579-668: k = s2 (1 + sqrt5 r + 5 r^2 / 3) exp(-sqrt5 r)."""

    kind = E.M52

    def __repr__(self) -> str:
        return f"Matern52(input_dim={self.input_dim}, σ²={self._signal_variance:.4f})"


class SquaredExponential(Kernel):
    """kernels.py:390-474: isotropic SE, k = s2 exp(-|x-x'|^2 / (2 l^2)) on the *unscaled* squared distance."""

    def __init__(self, signal_variance: float = 1.0, lengthscale: float = 1.0):
        self._signal_variance = signal_variance
        self._lengthscale = lengthscale

    @property
    def signal_variance(self) -> float:
        return self._signal_variance

    @property
    def lengthscale(self) -> float:
        return self._lengthscale

    def _dist_sq(self, X1, X2):
        X1 = _atleast_2d(X1)
        X2 = None if X2 is None else _atleast_2d(X2)
        D = X1.shape[1]
        return E.gram(E.DIST2, X1, X2, np.ones(D), 1.0), X1

    def __call__(self, X1, X2=None):
        d2, X1 = self._dist_sq(X1, X2)
        K = self._signal_variance * E.torch_().exp(-d2 / (2 * self._lengthscale**2))
        return _out(K, X1)

    def diagonal(self, X):
        X = _atleast_2d(X)
        if _is_dev(X):
            return E.torch_().full((X.shape[0],), float(self._signal_variance), dtype=X.dtype, device=X.device)
        return np.full(X.shape[0], self._signal_variance)

    def gradients(self, X1, X2=None) -> dict:
        d2, X1 = self._dist_sq(X1, X2)
        K = self._signal_variance * E.torch_().exp(-d2 / (2 * self._lengthscale**2))
        return {"log_signal_variance": _out(K, X1), "log_lengthscale": _out(K * d2 / (self._lengthscale**2), X1)}

    @property
    def n_params(self) -> int:
        return 2

    @property
    def param_names(self) -> List[str]:
        return ["log_signal_variance", "log_lengthscale"]

    def get_params(self):
        return np.array([np.log(self._signal_variance), np.log(self._lengthscale)])

    def set_params(self, params) -> None:
        self._signal_variance = np.exp(params[0])
        self._lengthscale = np.exp(params[1])

    def __repr__(self) -> str:
        return f"SquaredExponential(σ²={self._signal_variance:.4f}, l={self._lengthscale:.4f})"


class _Composite(Kernel):
    def __init__(self, k1: Kernel, k2: Kernel):
        self.k1 = k1
        self.k2 = k2

    @property
    def n_params(self) -> int:
        return self.k1.n_params + self.k2.n_params

    @property
    def param_names(self) -> List[str]:
        return [f"k1_{n}" for n in self.k1.param_names] + [f"k2_{n}" for n in self.k2.param_names]

    def get_params(self):
        return np.concatenate([self.k1.get_params(), self.k2.get_params()])

    def set_params(self, params) -> None:
        n1 = self.k1.n_params
        self.k1.set_params(params[:n1])
        self.k2.set_params(params[n1:])


class SumKernel(_Composite):
    """kernels.py:676-726."""

    def __call__(self, X1, X2=None):
        return self.k1(X1, X2) + self.k2(X1, X2)

    def diagonal(self, X):
        return self.k1.diagonal(X) + self.k2.diagonal(X)

    def gradients(self, X1, X2=None) -> dict:
        grads = {}
        for name, grad in self.k1.gradients(X1, X2).items():
            grads[f"k1_{name}"] = grad
        for name, grad in self.k2.gradients(X1, X2).items():
            grads[f"k2_{name}"] = grad
        return grads

    def __repr__(self) -> str:
        return f"SumKernel({self.k1}, {self.k2})"


class ProductKernel(_Composite):
    """kernels.py:729-782."""

    def __call__(self, X1, X2=None):
        return self.k1(X1, X2) * self.k2(X1, X2)

    def diagonal(self, X):
        return self.k1.diagonal(X) * self.k2.diagonal(X)

    def gradients(self, X1, X2=None) -> dict:
        K1 = self.k1(X1, X2)
        K2 = self.k2(X1, X2)
        grads = {}
        for name, grad in self.k1.gradients(X1, X2).items():
            grads[f"k1_{name}"] = grad * K2
        for name, grad in self.k2.gradients(X1, X2).items():
            grads[f"k2_{name}"] = K1 * grad
        return grads

    def __repr__(self) -> str:
        return f"ProductKernel({self.k1}, {self.k2})"


class WhiteNoise(Kernel):
    """kernels.py:790-844: s2 * I for a self-Gram, zeros for any cross-covariance."""

    def __init__(self, noise_variance: float = 1e-6):
        self._noise_variance = noise_variance

    @property
    def noise_variance(self) -> float:
        return self._noise_variance

    def __call__(self, X1, X2=None):
        X1 = _atleast_2d(X1)
        if _is_dev(X1):
            torch = E.torch_()
            if X2 is None:
                return self._noise_variance * torch.eye(X1.shape[0], dtype=X1.dtype, device=X1.device)
            return torch.zeros((X1.shape[0], _atleast_2d(X2).shape[0]), dtype=X1.dtype, device=X1.device)
        if X2 is None:
            return self._noise_variance * np.eye(X1.shape[0])
        return np.zeros((X1.shape[0], np.atleast_2d(X2).shape[0]))

    def diagonal(self, X):
        X = _atleast_2d(X)
        if _is_dev(X):
            return E.torch_().full((X.shape[0],), float(self._noise_variance), dtype=X.dtype, device=X.device)
        return np.full(X.shape[0], self._noise_variance)

    def gradients(self, X1, X2=None) -> dict:
        return {"log_noise_variance": self(X1, X2)}

    @property
    def n_params(self) -> int:
        return 1

    @property
    def param_names(self) -> List[str]:
        return ["log_noise_variance"]

    def get_params(self):
        return np.array([np.log(self._noise_variance)])

    def set_params(self, params) -> None:
        self._noise_variance = np.exp(params[0])

    def __repr__(self) -> str:
        return f"WhiteNoise(σ²={self._noise_variance:.6f})"


def create_se_ard_kernel(input_dim: int, signal_variance: float = 1.0, lengthscales=None) -> SquaredExponentialARD:
    """kernels.py:852-873."""
    return SquaredExponentialARD(input_dim=input_dim, signal_variance=signal_variance, lengthscales=lengthscales)


def create_matern_kernel(input_dim: int, nu: float = 2.5, signal_variance: float = 1.0, lengthscales=None) -> Kernel:
    """kernels.py:876-898."""
    if nu in (1.5, 3 / 2):
        return Matern32(input_dim, signal_variance, lengthscales)
    if nu in (2.5, 5 / 2):
        return Matern52(input_dim, signal_variance, lengthscales)
    raise ValueError(f"Unsupported nu={nu}. Use 1.5 (Matern32) or 2.5 (Matern52)")

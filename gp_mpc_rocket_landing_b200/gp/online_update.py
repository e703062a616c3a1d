"""Online data path in front of `fit` (reference src/gp/online_update.py): novelty-gated training buffer, update
scheduling, residual collection -- same classes, methods and return values -- with the two pieces that are data-parallel
moved to the device:

* novelty gating of a whole batch of candidate observations (one scenario each, as the Monte Carlo loop produces them) is
  ONE batched GP variance query (`add_observations`) instead of a predict call per point (online_update.py:316-325);
* `update()` extends the Cholesky factor by the new rows (ExactGP.update -> ExactFactor.append, O(n^2 m)) when the model
  is an exact GP whose training set is a prefix of the buffer; the reference refits from scratch every
  `update_interval` points (online_update.py:383-391, sparse_gp.py:328-353).  Sparse GPs are refit, as in the reference
  (their fit is O(n M^2) already).

The buffer is a preallocated ring (the reference uses four deques and rebuilds the arrays with np.array(list(...)) on
every access, online_update.py:190-204).
"""
from __future__ import annotations

import time
from dataclasses import dataclass
from typing import Callable, List, Optional, Tuple

import numpy as np

from .kernels import _is_dev


@dataclass
class OnlineUpdateConfig:
    """online_update.py:36-61."""

    buffer_size: int = 1000
    min_data_for_fit: int = 20
    use_novelty_filter: bool = True
    novelty_threshold: float = 0.3
    min_distance: float = 0.01
    update_interval: int = 10
    refit_interval: int = 100
    max_update_time_ms: float = 5.0
    add_inducing_on_novel: bool = True
    max_inducing_points: int = 100
    inducing_update_threshold: float = 0.5


@dataclass
class DataPoint:
    """online_update.py:64-72."""

    x: np.ndarray
    u: np.ndarray
    d: np.ndarray
    timestamp: float
    novelty: float = 0.0


class DataBuffer:
    """online_update.py:75-229: FIFO of at most `max_size` (features, targets) rows, oldest evicted first."""

    def __init__(self, max_size: int = 1000, feature_dim: int = 10, target_dim: int = 3):
        self.max_size = max_size
        self.feature_dim = feature_dim
        self.target_dim = target_dim
        self._Z = np.zeros((max_size, feature_dim))
        self._D = np.zeros((max_size, target_dim))
        self._nov = np.zeros(max_size)
        self._ts = np.zeros(max_size)
        self._head = 0          # index of the oldest row
        self._size = 0
        self._total_added = 0
        self._total_rejected = 0
        self.evictions = 0      # rows dropped at the front so far (lets a consumer detect that its prefix moved)

    @property
    def size(self) -> int:
        return self._size

    @property
    def is_empty(self) -> bool:
        return self._size == 0

    @property
    def is_full(self) -> bool:
        return self._size >= self.max_size

    def _order(self):
        return (self._head + np.arange(self._size)) % self.max_size

    def add(self, features, targets, novelty: float = 0.0, timestamp: Optional[float] = None) -> bool:
        if timestamp is None:
            timestamp = time.time()
        if self._size == self.max_size:
            slot = self._head
            self._head = (self._head + 1) % self.max_size
            self.evictions += 1
        else:
            slot = (self._head + self._size) % self.max_size
            self._size += 1
        self._Z[slot] = features
        self._D[slot] = targets
        self._nov[slot] = novelty
        self._ts[slot] = timestamp
        self._total_added += 1
        return True

    def add_if_novel(self, features, targets, novelty: float, threshold: float = 0.3,
                     min_distance: Optional[float] = None) -> bool:
        """online_update.py:154-188: novelty threshold, then minimum Euclidean distance to the stored features."""
        if novelty < threshold:
            self._total_rejected += 1
            return False
        if min_distance is not None and self._size > 0:
            Z = self._Z[self._order()]
            if np.min(np.linalg.norm(Z - features, axis=1)) < min_distance:
                self._total_rejected += 1
                return False
        return self.add(features, targets, novelty)

    def get_features(self) -> np.ndarray:
        if self.is_empty:
            return np.empty((0, self.feature_dim))
        return self._Z[self._order()].copy()

    def get_targets(self) -> np.ndarray:
        if self.is_empty:
            return np.empty((0, self.target_dim))
        return self._D[self._order()].copy()

    def get_data(self) -> Tuple[np.ndarray, np.ndarray]:
        return self.get_features(), self.get_targets()

    def get_recent(self, n: int) -> Tuple[np.ndarray, np.ndarray]:
        n = min(n, self._size)
        idx = self._order()[self._size - n:]
        return self._Z[idx].copy(), self._D[idx].copy()

    def clear(self) -> None:
        self._head = 0
        self._size = 0

    def get_statistics(self) -> dict:
        return {
            "size": self.size,
            "max_size": self.max_size,
            "total_added": self._total_added,
            "total_rejected": self._total_rejected,
            "acceptance_rate": self._total_added / max(1, self._total_added + self._total_rejected),
            "mean_novelty": float(np.mean(self._nov[self._order()])) if self.size > 0 else 0.0,
        }


def _is_fitted(gp) -> bool:
    """online_update.py:316 tests `gp._L_B is not None`; here both SparseGP and ExactGP hold their fitted state in
    `_model` (reading `_L_B` would copy the factor to the host)."""
    if hasattr(gp, "_model"):
        return gp._model is not None
    return getattr(gp, "_L_B", None) is not None


def _targets_1d(D: np.ndarray) -> np.ndarray:
    """online_update.py:386-391: single-output GPs get column 0, or the row mean of a multi-column target."""
    if D.ndim == 1:
        return D
    return D[:, 0] if D.shape[1] == 1 else D.mean(axis=1)


class OnlineGPUpdater:
    """online_update.py:232-425 for a SparseGP or an ExactGP."""

    def __init__(self, gp, config: Optional[OnlineUpdateConfig] = None,
                 feature_extractor: Optional[Callable[[np.ndarray, np.ndarray], np.ndarray]] = None):
        self.gp = gp
        self.config = config or OnlineUpdateConfig()
        self.feature_extractor = feature_extractor or (lambda x, u: np.concatenate([x, u]))
        self._feature_dim: Optional[int] = None
        self._target_dim: int = 1
        self._buffer: Optional[DataBuffer] = None
        self._points_since_update = 0
        self._points_since_refit = 0
        self._total_updates = 0
        self._last_update_time = 0.0
        self._update_times: List[float] = []
        self._fitted_rows = 0          # buffer rows [0, _fitted_rows) are the GP's training set ...
        self._fitted_evictions = -1    # ... as long as the buffer has not evicted since
        self.incremental_updates = 0

    def _ensure_buffer(self, feature_dim: int, target_dim: int = 1) -> None:
        if self._buffer is None:
            self._feature_dim = feature_dim
            self._target_dim = target_dim
            self._buffer = DataBuffer(max_size=self.config.buffer_size, feature_dim=feature_dim, target_dim=target_dim)

    def _novelty(self, Z: np.ndarray) -> np.ndarray:
        """variance / prior variance per row (online_update.py:318-321), one batched device query."""
        if not (self.config.use_novelty_filter and _is_fitted(self.gp)):
            return np.ones(Z.shape[0])
        try:
            pred = self.gp.predict(Z)
            var = np.asarray(pred.variance, dtype=np.float64).reshape(Z.shape[0], -1)
            return var.mean(axis=1) / float(self.gp.kernel.signal_variance)
        except Exception:  # noqa: BLE001 (reference behaviour: assume novel when the prediction fails)
            return np.ones(Z.shape[0])

    def _gate(self, z, d, novelty) -> bool:
        if self.config.use_novelty_filter:
            added = self._buffer.add_if_novel(z, d, float(novelty), threshold=self.config.novelty_threshold,
                                              min_distance=self.config.min_distance)
        else:
            added = self._buffer.add(z, d, float(novelty))
        if added:
            self._points_since_update += 1
            self._points_since_refit += 1
        return added

    def add_observation(self, x, u, d) -> bool:
        """online_update.py:293-345."""
        z = np.asarray(self.feature_extractor(x, u), dtype=np.float64)
        d = np.atleast_1d(np.asarray(d, dtype=np.float64))
        self._ensure_buffer(len(z), len(d))
        return self._gate(z, d, self._novelty(z[None])[0])

    def add_observations(self, X, U, D) -> np.ndarray:
        """Batch form of `add_observation` (one row per scenario of the Monte Carlo loop): the novelty scores of all rows
        come from one batched variance query; the threshold / minimum-distance rules are then applied in row order, so
        the accepted set equals that of the row-by-row calls (the GP does not change between them)."""
        X, U = np.atleast_2d(np.asarray(X, dtype=np.float64)), np.atleast_2d(np.asarray(U, dtype=np.float64))
        if X.size == 0:
            return np.zeros(0, dtype=bool)
        D = np.asarray(D, dtype=np.float64).reshape(X.shape[0], -1)
        Z = np.stack([np.asarray(self.feature_extractor(X[i], U[i]), dtype=np.float64) for i in range(X.shape[0])])
        self._ensure_buffer(Z.shape[1], D.shape[1])
        nov = self._novelty(Z)
        return np.array([self._gate(Z[i], D[i], nov[i]) for i in range(Z.shape[0])], dtype=bool)

    def should_update(self) -> bool:
        if self._buffer is None or self._buffer.size < self.config.min_data_for_fit:
            return False
        return self._points_since_update >= self.config.update_interval

    def should_refit(self) -> bool:
        if self._buffer is None:
            return False
        return self._points_since_refit >= self.config.refit_interval

    def update(self, force: bool = False) -> dict:
        """online_update.py:361-408.  An exact GP whose training rows are still the front of the buffer is extended by
        the new rows; every `refit_interval` points (and always for sparse GPs) the model is refit on the whole buffer."""
        if self._buffer is None or self._buffer.is_empty:
            return {"status": "no_data"}
        if not force and not self.should_update():
            return {"status": "skipped"}
        start = time.perf_counter()
        Z, D = self._buffer.get_data()
        y = _targets_1d(D)
        refit = self.should_refit()
        # extension is only valid for an exact GP whose training set is still exactly the front of the buffer: a sparse GP
        # is refit as in the reference (online_update.py:383-391), and an external gp.fit() on other data in between (its
        # n_train no longer equals the rows this updater fitted) falls back to a full refit as well
        from .exact_gp import ExactGP
        incremental = isinstance(self.gp, ExactGP) or getattr(self.gp, "supports_incremental_update", False)
        can_extend = (not refit and incremental and getattr(self.gp, "_model", None) is not None
                      and self._fitted_evictions == self._buffer.evictions and 0 < self._fitted_rows <= Z.shape[0]
                      and int(getattr(self.gp, "n_train", -1)) == self._fitted_rows)
        mode = "refit"
        try:
            if can_extend and self._fitted_rows < Z.shape[0]:
                self.gp.update(Z[self._fitted_rows:], y[self._fitted_rows:])
                mode = "extended"
                self.incremental_updates += 1
            elif not (can_extend and self._fitted_rows == Z.shape[0]):
                self.gp.fit(Z, y)
            self._fitted_rows, self._fitted_evictions = Z.shape[0], self._buffer.evictions
            status = "success"
        except Exception as e:  # noqa: BLE001 (reference behaviour)
            status = f"error: {e}"
            self._fitted_rows, self._fitted_evictions = 0, -1
        elapsed = (time.perf_counter() - start) * 1000
        self._update_times.append(elapsed)
        self._last_update_time = elapsed
        self._total_updates += 1
        self._points_since_update = 0
        if refit or mode == "refit":   # any full fit restarts the refit interval (the reference only resets the counter)
            self._points_since_refit = 0
        return {"status": status, "time_ms": elapsed, "n_data": self._buffer.size, "total_updates": self._total_updates,
                "mode": mode}

    def get_statistics(self) -> dict:
        stats = {"total_updates": self._total_updates, "points_since_update": self._points_since_update,
                 "last_update_time_ms": self._last_update_time, "incremental_updates": self.incremental_updates}
        if self._buffer is not None:
            stats.update(self._buffer.get_statistics())
        if self._update_times:
            stats["mean_update_time_ms"] = float(np.mean(self._update_times))
            stats["max_update_time_ms"] = float(np.max(self._update_times))
        return stats


class OnlineStructuredGPUpdater:
    """online_update.py:428-557: novelty gate (`gp.is_novel`) in front of `StructuredRocketGP.add_data`, scheduled fit."""

    def __init__(self, gp, config: Optional[OnlineUpdateConfig] = None):
        self.gp = gp
        self.config = config or OnlineUpdateConfig()
        self._points_since_update = 0
        self._total_observations = 0
        self._total_updates = 0
        self._update_times: List[float] = []

    def _accept(self, x, u, d_v, d_omega) -> None:
        self.gp.add_data(np.reshape(x, (1, -1)), np.reshape(u, (1, -1)), np.reshape(d_v, (1, -1)),
                         np.reshape(d_omega, (1, -1)))
        self._points_since_update += 1
        self._total_observations += 1

    def add_observation(self, x, u, d_v, d_omega) -> bool:
        if self.config.use_novelty_filter and self.gp._is_fitted and not self.gp.is_novel(x, u):
            return False
        self._accept(x, u, d_v, d_omega)
        return True

    def add_observations(self, X, U, D_v, D_omega) -> np.ndarray:
        """Batch form: the novelty rule of structured_gp.py:344-357 (max variance > threshold * signal variance)
        evaluated for all rows by one batched predict.  `add_data` marks the GP unfitted, after which the reference's gate
        accepts everything (`gp._is_fitted` is False) -- same here: rows after the first accepted one are all kept."""
        X, U = np.atleast_2d(np.asarray(X, dtype=np.float64)), np.atleast_2d(np.asarray(U, dtype=np.float64))
        D_v, D_omega = np.atleast_2d(D_v), np.atleast_2d(D_omega)
        B = X.shape[0]
        keep = np.ones(B, dtype=bool)
        if self.config.use_novelty_filter and self.gp._is_fitted:
            _, _, vv, vw = self.gp.predict_batch(X, U)
            vv, vw = (np.asarray(v.cpu().numpy() if _is_dev(v) else v) for v in (vv, vw))
            novel = np.maximum(vv.max(axis=1), vw.max(axis=1)) > self.gp.config.novelty_threshold * self.gp.config.signal_variance
            first = int(np.argmax(novel)) if novel.any() else B
            keep[:first] = False
        for i in np.nonzero(keep)[0]:
            self._accept(X[i], U[i], D_v[i], D_omega[i])
        return keep

    def should_update(self) -> bool:
        if self.gp.n_data < self.config.min_data_for_fit:
            return False
        return self._points_since_update >= self.config.update_interval

    def update(self, force: bool = False) -> dict:
        if not force and not self.should_update():
            return {"status": "skipped"}
        start = time.perf_counter()
        try:
            self.gp.fit()
            status = "success"
        except Exception as e:  # noqa: BLE001 (reference behaviour)
            status = f"error: {e}"
        elapsed = (time.perf_counter() - start) * 1000
        self._update_times.append(elapsed)
        self._total_updates += 1
        self._points_since_update = 0
        return {"status": status, "time_ms": elapsed, "n_data": self.gp.n_data, "total_updates": self._total_updates}

    def get_statistics(self) -> dict:
        stats = {"total_observations": self._total_observations, "total_updates": self._total_updates,
                 "points_since_update": self._points_since_update, "n_data": self.gp.n_data,
                 "gp_fitted": self.gp._is_fitted}
        if self._update_times:
            stats["mean_update_time_ms"] = float(np.mean(self._update_times))
            stats["max_update_time_ms"] = float(np.max(self._update_times))
        return stats


class ResidualCollector:
    """online_update.py:559-685: d = (x_actual - f_nominal(x, u)) / dt in the velocity / body-rate rows.
    `record_batch` does the same for a batch of transitions with one call of a batched nominal model
    (e.g. `Rocket6DoFDynamics.step_batch`, which runs `gpmpc_step_6dof` on the device)."""

    def __init__(self, nominal_dynamics: Callable, max_samples: int = 10000):
        self.nominal_dynamics = nominal_dynamics
        self.max_samples = max_samples
        self.states: List[np.ndarray] = []
        self.controls: List[np.ndarray] = []
        self.residuals_v: List[np.ndarray] = []
        self.residuals_omega: List[np.ndarray] = []

    def _push(self, x, u, d_v, d_omega) -> None:
        if len(self.states) >= self.max_samples:
            for lst in (self.states, self.controls, self.residuals_v, self.residuals_omega):
                lst.pop(0)
        self.states.append(np.array(x, dtype=np.float64))
        self.controls.append(np.array(u, dtype=np.float64))
        self.residuals_v.append(d_v)
        self.residuals_omega.append(d_omega)

    def record(self, x, u, x_next_actual, dt: float) -> None:
        residual = np.asarray(x_next_actual) - np.asarray(self.nominal_dynamics(x, u, dt))
        self._push(x, u, residual[4:7] / dt, residual[11:14] / dt)

    def record_batch(self, X, U, X_next_actual, dt: float, batched_dynamics: Optional[Callable] = None) -> None:
        """B transitions at once; `batched_dynamics(X, U, dt) -> X_next_nominal` (B, 14) defaults to a row loop over
        `nominal_dynamics`."""
        X, U, Xn = (np.atleast_2d(np.asarray(a, dtype=np.float64)) for a in (X, U, X_next_actual))
        if batched_dynamics is not None:
            nom = batched_dynamics(X, U, dt)
            nom = nom.cpu().numpy() if _is_dev(nom) else np.asarray(nom)
        else:
            nom = np.stack([self.nominal_dynamics(X[i], U[i], dt) for i in range(X.shape[0])])
        R = Xn - nom
        for i in range(X.shape[0]):
            self._push(X[i], U[i], R[i, 4:7] / dt, R[i, 11:14] / dt)

    def get_training_data(self) -> Tuple[np.ndarray, np.ndarray, np.ndarray, np.ndarray]:
        if not self.states:
            return np.empty((0, 14)), np.empty((0, 3)), np.empty((0, 3)), np.empty((0, 3))
        return (np.array(self.states), np.array(self.controls), np.array(self.residuals_v),
                np.array(self.residuals_omega))

    @property
    def n_samples(self) -> int:
        return len(self.states)

    def clear(self) -> None:
        for lst in (self.states, self.controls, self.residuals_v, self.residuals_omega):
            lst.clear()

    def get_statistics(self) -> dict:
        if not self.states:
            return {"n_samples": 0}
        D_v, D_w = np.array(self.residuals_v), np.array(self.residuals_omega)
        return {"n_samples": self.n_samples, "d_v_mean": np.mean(D_v, axis=0).tolist(),
                "d_v_std": np.std(D_v, axis=0).tolist(), "d_omega_mean": np.mean(D_w, axis=0).tolist(),
                "d_omega_std": np.std(D_w, axis=0).tolist()}

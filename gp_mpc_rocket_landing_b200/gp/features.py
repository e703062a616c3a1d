"""(x, u) -> GP input features with the reference's API (src/gp/features.py), computed by the device kernel
`gpmpc_features` (csrc/gram.cu) in the reference's operation order.

The device kernel implements the reference defaults (exponential atmosphere rho0 = 1.225, H = 8500 m, reference
velocity 10, altitude and density features included) -- the only configuration the reference's GP classes ever
construct (structured_gp.py:104-107, :430-432).  Other settings raise: there is no CPU path.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional

import numpy as np

from .. import _lib
from . import engine as E
from .kernels import _is_dev


@dataclass
class AtmosphereModel:
    """features.py:44-63 (parameter carrier; the density itself is evaluated inside the device kernels)."""

    rho_0: float = 1.225
    scale_height: float = 8500.0

    def density(self, altitude):
        return self.rho_0 * np.exp(-altitude / self.scale_height)

    def density_array(self, altitude):
        return self.rho_0 * np.exp(-np.asarray(altitude) / self.scale_height)


class RocketFeatureExtractor:
    """features.py:71-146."""

    _which = -1
    _base_names: List[str] = []
    _nx = 14

    def __init__(self, atmosphere: Optional[AtmosphereModel] = None, include_altitude: bool = True,
                 include_density: bool = True, reference_velocity: float = 10.0):
        self.atmosphere = atmosphere or AtmosphereModel()
        self.include_altitude = include_altitude
        self.include_density = include_density
        self.v_ref = reference_velocity
        if (self.atmosphere.rho_0, self.atmosphere.scale_height, self.v_ref) != (1.225, 8500.0, 10.0) or not (
                include_altitude and include_density):
            raise _lib.GpmpcError("the device feature kernels implement the reference defaults only "
                                  "(rho0=1.225, H=8500, v_ref=10, altitude and density included)")
        self._compute_feature_dim()

    def _compute_feature_dim(self) -> None:
        self._feature_names = list(self._base_names)
        self._n_features = len(self._feature_names)

    @property
    def n_features(self) -> int:
        return self._n_features

    @property
    def feature_names(self) -> List[str]:
        return self._feature_names

    def extract_batch(self, X, U):
        """(N, nx), (N, 3) -> (N, n_features); one device launch instead of the reference's Python loop (:129-146)."""
        dev_in = _is_dev(X)
        Xk = X if dev_in else np.atleast_2d(np.asarray(X, dtype=np.float64))
        Uk = U if _is_dev(U) else np.atleast_2d(np.asarray(U, dtype=np.float64))
        Z = E.features(self._which, Xk, Uk)
        return Z if dev_in else Z.cpu().numpy()

    def extract(self, x, u):
        Z = self.extract_batch(x.reshape(1, -1) if _is_dev(x) else np.asarray(x, dtype=np.float64).reshape(1, -1),
                               u.reshape(1, -1) if _is_dev(u) else np.asarray(u, dtype=np.float64).reshape(1, -1))
        return Z[0]


class TranslationalFeatureExtractor(RocketFeatureExtractor):
    """features.py:149-270 (13 features)."""

    _which = 1
    _base_names = ["v_x", "v_y", "v_z", "speed", "q_dyn", "alpha", "beta", "T_x", "T_y", "T_z", "T_mag", "altitude",
                   "density"]


class RotationalFeatureExtractor(RocketFeatureExtractor):
    """features.py:273-366 (12 features)."""

    _which = 2
    _base_names = ["omega_x", "omega_y", "omega_z", "omega_mag", "T_x", "T_y", "T_z", "v_Bx", "v_By", "v_Bz", "speed",
                   "q_dyn"]


class Simple3DoFFeatureExtractor(RocketFeatureExtractor):
    """features.py:368-444 (11 features)."""

    _which = 0
    _nx = 7
    _base_names = ["v_x", "v_y", "v_z", "speed", "q_dyn", "T_x", "T_y", "T_z", "T_mag", "altitude", "density"]


class CombinedFeatureExtractor:
    """features.py:447-491."""

    def __init__(self, atmosphere: Optional[AtmosphereModel] = None, reference_velocity: float = 10.0):
        self.translational = TranslationalFeatureExtractor(atmosphere=atmosphere, reference_velocity=reference_velocity)
        self.rotational = RotationalFeatureExtractor(atmosphere=atmosphere, reference_velocity=reference_velocity)

    @property
    def n_features_translational(self) -> int:
        return self.translational.n_features

    @property
    def n_features_rotational(self) -> int:
        return self.rotational.n_features

    def extract_translational(self, x, u):
        return self.translational.extract(x, u)

    def extract_rotational(self, x, u):
        return self.rotational.extract(x, u)

    def extract_batch_translational(self, X, U):
        return self.translational.extract_batch(X, U)

    def extract_batch_rotational(self, X, U):
        return self.rotational.extract_batch(X, U)

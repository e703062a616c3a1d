"""`src.gp` surface of the hot path (reference src/gp/__init__.py:27-128), B200-native."""
from . import engine  # noqa: F401  (imports _lib: fails loudly when libgpmpc_b200.so is missing)
from .exact_gp import ExactGP, GPPrediction, MultiOutputExactGP
from .fast_gp import (CachedGPPredictor, FastGPConfig, FastGPPredictor, NumbaGPPredictor, SparseGPPredictor,
                      create_fast_gp)
from .features import (AtmosphereModel, CombinedFeatureExtractor, RocketFeatureExtractor, RotationalFeatureExtractor,
                       Simple3DoFFeatureExtractor, TranslationalFeatureExtractor)
from .kernels import (RBF, SE_ARD, Kernel, Matern32, Matern52, ProductKernel, SquaredExponential,
                      SquaredExponentialARD, SumKernel, WhiteNoise, create_matern_kernel, create_se_ard_kernel)
from .online_update import (DataBuffer, DataPoint, OnlineGPUpdater, OnlineStructuredGPUpdater, OnlineUpdateConfig,
                            ResidualCollector)
from .sparse_gp import MultiOutputSparseGP, SparseGP
from .structured_gp import Simple3DoFGP, StructuredGPConfig, StructuredRocketGP

__all__ = [
    "ExactGP", "GPPrediction", "MultiOutputExactGP", "AtmosphereModel", "CombinedFeatureExtractor",
    "RocketFeatureExtractor", "RotationalFeatureExtractor", "Simple3DoFFeatureExtractor",
    "TranslationalFeatureExtractor", "RBF", "SE_ARD", "Kernel", "Matern32", "Matern52", "ProductKernel",
    "SquaredExponential", "SquaredExponentialARD", "SumKernel", "WhiteNoise", "create_matern_kernel",
    "create_se_ard_kernel", "MultiOutputSparseGP", "SparseGP", "CachedGPPredictor", "FastGPConfig",
    "FastGPPredictor", "NumbaGPPredictor", "SparseGPPredictor", "create_fast_gp", "Simple3DoFGP", "StructuredGPConfig", "StructuredRocketGP",
    "DataBuffer", "DataPoint", "OnlineGPUpdater", "OnlineStructuredGPUpdater", "OnlineUpdateConfig", "ResidualCollector",
]

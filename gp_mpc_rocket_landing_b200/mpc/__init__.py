"""`src.mpc` surface of the hot path (reference src/mpc/__init__.py:24-128), B200-native."""
from . import engine  # noqa: F401  (imports _lib: fails loudly when libgpmpc_b200.so is missing)
from .constraints import ConstraintParams, TightenedConstraints
from .cost_functions import CostWeights
from .gp_mpc import GPMPC, GPMPCConfig
from .gp_mpc6 import GPMPC6DoF, GPMPC6DoFConfig
from .nominal_mpc import BatchMPCSolution, MPCConfig, MPCSolution, NominalMPC3DoF
from .osqp_rti import BatchRTISolution, FastRTI3DoF, OSQPRTIConfig, OSQPRTIMPC, OSQPRTISolution
from .uncertainty_prop import (ConstraintTightening, PropagatedUncertainty, TubeBasedRobustness,
                               UncertaintyPropagator)

__all__ = [
    "ConstraintParams", "TightenedConstraints", "CostWeights", "GPMPC", "GPMPCConfig", "GPMPC6DoF", "GPMPC6DoFConfig", "BatchMPCSolution", "MPCConfig",
    "MPCSolution", "NominalMPC3DoF", "BatchRTISolution", "FastRTI3DoF", "OSQPRTIConfig", "OSQPRTIMPC",
    "OSQPRTISolution", "PropagatedUncertainty", "UncertaintyPropagator", "ConstraintTightening", "TubeBasedRobustness",
]

"""gp_mpc_rocket_landing_b200 -- B200-native (sm_100a CUDA) implementation of the per-control-step numeric core of
shiivashaakeri/gp-mpc-rocket-landing, batched over Monte Carlo descent scenarios.

Sub-packages mirror the reference's module layout for the hot path:
  gp   <- src/gp   (kernels, exact / sparse / fast GP, features, Simple3DoFGP / StructuredRocketGP)
  mpc  <- src/mpc  (OSQPRTIMPC, GPMPC, NominalMPC3DoF, uncertainty propagation, configs) + solve_batch / step_batch
All numerics run in libgpmpc_b200.so (include/gpmpc_b200.h); there is no CPU fallback: importing `gp` or `mpc`
fails loudly when the library has not been built (`python -m gp_mpc_rocket_landing_b200.build`).
"""
__version__ = "0.1.0"

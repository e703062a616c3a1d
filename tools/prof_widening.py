"""Profiling target for the kernels of the widening rows: one LML gradient, one factor extension, one
blocked potrs, one refit-sweep LML solve, one unscented propagation -- between cudaProfilerStart/Stop."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from gp_mpc_rocket_landing_b200.gp import ExactGP, Simple3DoFGP, SquaredExponentialARD, engine as E
from gp_mpc_rocket_landing_b200.dynamics import Rocket3DoFDynamics
from gp_mpc_rocket_landing_b200.mpc.uncertainty_prop import UncertaintyPropagator
from gp_mpc_rocket_landing_b200 import workloads

torch = E.torch_()
n = 4000
rng = np.random.default_rng(0)
X = rng.normal(size=(n + 10, 11)); y = np.sin(X[:, 0]) + 0.1 * rng.normal(size=n + 10)
gp = ExactGP(SquaredExponentialARD(11, 1.0, np.full(11, 3.0)), noise_variance=1e-2).fit(X[:n], y[:n])
gp.log_marginal_likelihood_gradient(); gp.update(X[n:n + 5], y[n:n + 5])          # warm-up
Xr, Ur, Dr = workloads.residual_dataset(1000, seed=0)
g3 = Simple3DoFGP(noise_variance=1e-2, use_sparse=False); g3.add_data(Xr, Ur, Dr); g3.fit()
up = UncertaintyPropagator(Rocket3DoFDynamics(), g3, method="unscented")
X0, _ = workloads.sample_initial_conditions(0, 1024)
U = np.zeros((1024, 20, 3)); U[:, :, 0] = X0[:, :1]
up.propagate_batch(X0, U, None, 0.1)
torch.cuda.synchronize()
torch.cuda.profiler.start()
gp.log_marginal_likelihood_gradient()
gp.update(X[n + 5:n + 10], y[n + 5:n + 10])
up.propagate_batch(X0, U, None, 0.1)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("done")

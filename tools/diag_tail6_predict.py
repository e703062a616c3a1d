"""Diagnostic (not part of the product): is the 6-DoF ADMM iteration count predictable before the launch?  Correlates the
per-QP count with quantities known at assembly time: the count of the previous control step, the equality residual of
the primal warm start, its box violation, the carried rho."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import gp_mpc_rocket_landing_b200.gp as gp
import gp_mpc_rocket_landing_b200.mpc as mpc
from gp_mpc_rocket_landing_b200 import dynamics, workloads

X, U, Dv, Dw = workloads.residual_dataset_6dof(1000, seed=0)
sg = gp.StructuredRocketGP(gp.StructuredGPConfig(n_inducing=100, use_sparse=True, noise_variance=1e-3))
np.random.seed(0)
sg.add_data(X, U, Dv, Dw); sg.fit()
dyn = dynamics.Rocket6DoFDynamics()
ctl = mpc.GPMPC6DoF(dyn, sg, mpc.GPMPC6DoFConfig(N=30, max_sqp_iter=1, use_gp_uncertainty=False))
X0, XT = workloads.sample_initial_conditions_6dof(0, 1024)
x = gp.engine.to_dev(X0); xt = gp.engine.to_dev(XT)
prev = None
rows = []
for k in range(24):
    st = ctl._batch
    rho_in = st.ws.rho.clone() if st is not None else None
    s = ctl.solve_batch(x, xt)
    st = ctl._batch
    it = st.ws.iters.cpu().numpy().astype(np.float64)
    # quantities of THIS QP known before its solve: stage data and the warm start it was given are gone (x holds the solution),
    # so use what the next QP will see: residual of the solution shifted... keep it simple: the assembled rhs magnitude
    AB = st.AB; beq = st.beq
    N = 30; nx, ns = 14, 17
    z = st.ws.x
    zs = z[:, :N * ns].view(-1, N, ns); xn = torch.cat([z[:, ns:N * ns].view(-1, N - 1, ns)[:, :, :nx], z[:, N * ns:].view(-1, 1, nx)], 1)
    res = torch.einsum("bkij,bkj->bki", AB, zs) - xn - beq[:, nx:].view(-1, N, nx)
    feats = {"prev_iters": prev, "rho_in": None if rho_in is None else rho_in.cpu().numpy(),
             "beq_norm": beq[:, nx:].abs().amax(1).cpu().numpy(), "x0_dist": (x - xt).abs().amax(1).cpu().numpy(),
             "sol_eq_res": res.abs().amax((1, 2)).cpu().numpy()}
    if k >= 4:
        rows.append({n: (None if v is None else float(np.corrcoef(v, it)[0, 1])) for n, v in feats.items()})
    prev = it
    x = dyn.step_batch(x, s.u0.contiguous(), 0.1)
out = {n: float(np.mean([r[n] for r in rows if r[n] is not None])) for n in rows[0]}
print(json.dumps({"mean_correlation_with_iteration_count": out}))

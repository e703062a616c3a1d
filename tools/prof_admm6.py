"""GPU timing / profiling target (not part of the product): 6-DoF QP data (n_x = 14, N = 30) from a closed-loop step, then
the ADMM kernel alone with fixed iteration counts (termination checks and rho adaptation off) -> per-iteration and set-up
cost; with ITERS set: ONE launch between cudaProfilerStart/Stop for `ncu --profile-from-start off`."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import gp_mpc_rocket_landing_b200.gp as gp
import gp_mpc_rocket_landing_b200.mpc as mpc
from gp_mpc_rocket_landing_b200 import _lib, dynamics, workloads
from gp_mpc_rocket_landing_b200.mpc import engine as me

B = int(os.environ.get("B", 1024))
X, U, Dv, Dw = workloads.residual_dataset_6dof(1000, seed=0)
sg = gp.StructuredRocketGP(gp.StructuredGPConfig(n_inducing=100, use_sparse=True, noise_variance=1e-3))
import numpy as np
np.random.seed(0)
sg.add_data(X, U, Dv, Dw); sg.fit()
dyn = dynamics.Rocket6DoFDynamics()
ctl = mpc.GPMPC6DoF(dyn, sg, mpc.GPMPC6DoFConfig(N=30, max_sqp_iter=1, use_gp_uncertainty=False))
X0, XT = workloads.sample_initial_conditions_6dof(0, B)
x = gp.engine.to_dev(X0); xt = gp.engine.to_dev(XT)
for _ in range(3):
    s = ctl.solve_batch(x, xt)
    x = dyn.step_batch(x, s.u0.contiguous(), 0.1)
st = ctl._batch


def run(iters, adaptive=0, check=0):
    S = _lib.AdmmSettings.reference_defaults(max_iter=iters, check_every=check, adaptive_rho=adaptive, warm_x=1, warm_y=1)
    x0, y0 = st.ws.x.clone(), st.ws.y.clone()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    best = 1e9
    for _ in range(3):
        st.ws.x.copy_(x0); st.ws.y.copy_(y0)
        torch.cuda.synchronize()
        e0.record()
        me.admm_solve(ctl._qp, st.AB, st.beq, st.q, S, st.ws)
        e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    st.ws.x.copy_(x0); st.ws.y.copy_(y0)
    return best


if "ITERS" in os.environ:
    S = _lib.AdmmSettings.reference_defaults(max_iter=int(os.environ["ITERS"]), check_every=0, adaptive_rho=0, warm_x=1, warm_y=1)
    torch.cuda.synchronize()
    torch.cuda.profiler.start()
    me.admm_solve(ctl._qp, st.AB, st.beq, st.q, S, st.ws)
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()
    print("done")
else:
    t = {n: run(n) for n in (1, 50, 100, 200)}
    out = {"B": B, "ms": t, "per_iteration_us_batch": (t[200] - t[100]) * 10, "setup_ms": t[1],
           "with_checks_every_25_200it": run(200, 0, 25), "with_adaptive_rho_200it": run(200, 1, 25)}
    print(json.dumps(out))

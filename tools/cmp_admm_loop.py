"""GPU A/B (not part of the product): ONE closed-loop control step from the same steady-state controller state with the
second- and third-generation ADMM kernels (GPMPCConfig.qp_kernel_variant 2 / 3): ADMM phase time, iteration histogram."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import bench
from gp_mpc_rocket_landing_b200.mpc import engine as me


class A:
    batch, sqp_iters, qp_max_iter = 4096, 1, 1000


dyn, gp, ctl, mc = bench.build_problem(A, 0)
for _ in range(int(os.environ.get("STEPS", 60))):
    mc.step()
torch.cuda.synchronize()
st = ctl._batch
keep = {k: getattr(st, k).clone() for k in ("U_warm", "X_warm", "has_warm", "U", "X")}
keep.update({"ws_" + k: getattr(st.ws, k).clone() for k in ("x", "y", "rho")})
mcs = {k: getattr(mc, k).clone() for k in ("x", "t_step", "active", "outcome", "counts", "m0", "x_final", "fail_steps") if hasattr(mc, k)}


def restore():
    for k, v in keep.items():
        (getattr(st.ws, k[3:]) if k.startswith("ws_") else getattr(st, k)).copy_(v)
    for k, v in mcs.items():
        getattr(mc, k).copy_(v)


out = {}
for rep in range(2):
    for variant in (2, 3):
        restore()
        ctl.config.qp_kernel_variant = variant
        tm = me.EventTimers(capacity=64)
        ctl.timers = tm
        torch.cuda.synchronize()
        mc.step()
        torch.cuda.synchronize()
        ph = tm.totals()
        ctl.timers = None
        it = st.ws.iters.cpu().numpy()
        out[f"gen{variant}_rep{rep}"] = {"admm_ms": ph["admm"][0], "mean_iters": float(it.mean()),
                                        "hist": {int(k): int(v) for k, v in zip(*np.unique(it, return_counts=True))}}
print(json.dumps(out))

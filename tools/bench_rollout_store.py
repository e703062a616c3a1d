"""GPU measurement (not part of the product): what the K* store costs the rollout kernel, i.e. the most a fusion of the
cross-Gram into the variance GEMM could save on the rollout side.  Prints one JSON line."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from gp_mpc_rocket_landing_b200.mpc import engine as me


class A:
    batch, sqp_iters, qp_max_iter = 4096, 1, 1000


dyn, gp, ctl, mc = bench.build_problem(A, 0)
for _ in range(3):
    mc.step()
st = ctl._batch
torch.cuda.synchronize()
mdl = ctl._gp_model_struct() if hasattr(ctl, "_gp_model_struct") else ctl._model_struct


def run(Ks):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(3):
        me.rollout(mdl, ctl._dyn, mc.x, st.U, st.X, st.dmean, Ks, None)
    e0.record()
    for _ in range(20):
        me.rollout(mdl, ctl._dyn, mc.x, st.U, st.X, st.dmean, Ks, None)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / 20


print(json.dumps({"rollout_ms_with_Ks_store": run(st.Ks), "rollout_ms_without_Ks_store": run(None), "B": 4096, "N": 20,
                  "Ks_bytes": st.Ks.numel() * 8}))

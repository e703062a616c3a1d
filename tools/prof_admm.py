"""GPU profiling target (not part of the product): realistic QP data from a few closed-loop steps, then ONE ADMM launch
with a fixed iteration count (termination checks and rho adaptation off) for `ncu -k regex:admm`."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from gp_mpc_rocket_landing_b200 import _lib
from gp_mpc_rocket_landing_b200.mpc import engine as me

class A: pass
a = A(); a.batch = int(os.environ.get("B", 4096)); a.sqp_iters = 1; a.qp_max_iter = 1000
dyn, gp, ctl, mc = bench.build_problem(a, 0)
for k in range(3):
    mc.step()
st_ = ctl._batch
S = _lib.AdmmSettings.reference_defaults(max_iter=int(os.environ.get("ITERS", 100)), check_every=0, adaptive_rho=0,
                                         warm_x=1, warm_y=1, kernel_variant=int(os.environ.get("VARIANT", 0)))
torch.cuda.synchronize()
torch.cuda.profiler.start()
me.admm_solve(ctl._qp, st_.AB, st_.beq, st_.q, S, st_.ws)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("done")

"""GPU diagnostic (not part of the product): ADMM iteration histogram per closed-loop step and the per-iteration cost
of the ADMM kernel (fixed iteration count, termination disabled)."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from gp_mpc_rocket_landing_b200 import _lib
from gp_mpc_rocket_landing_b200.mpc import engine as me

class A: pass
a = A(); a.batch = int(os.environ.get("B", 4096)); a.sqp_iters = 1; a.qp_max_iter = 1000
dyn, gp, ctl, mc = bench.build_problem(a, 0)
if os.environ.get("CHECK"):
    ctl.config.qp_check_every = int(os.environ["CHECK"])
if os.environ.get("RHO_KEEP"):
    _orig = ctl._settings
    def _s():
        S = _orig(); S.rho_from_input = 1; return S
    ctl._settings = _s
for k in range(int(os.environ.get("STEPS", 30))):
    torch.cuda.synchronize(); _t=time.perf_counter()
    sol = mc.step()
    torch.cuda.synchronize(); _dt=(time.perf_counter()-_t)*1e3
    it = sol.qp_iterations.cpu().numpy(); st = sol.qp_status.cpu().numpy()
    h = np.bincount(np.minimum(it // 25, 41))
    print(k, "%.2f ms" % _dt, "iters mean %.1f max %d" % (it.mean(), it.max()), "hist/25:", {int(i) * 25: int(c) for i, c in enumerate(h) if c},
          "status:", {int(s): int((st == s).sum()) for s in np.unique(st)}, flush=True)
# per-iteration cost: fixed iterations, no termination checks, no adaptive rho
st_ = ctl._batch
for iters in (50, 100, 200):
    S = _lib.AdmmSettings.reference_defaults(max_iter=iters, check_every=0, adaptive_rho=0, warm_x=1, warm_y=1)
    me.admm_solve(ctl._qp, st_.AB, st_.beq, st_.q, S, st_.ws)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); me.admm_solve(ctl._qp, st_.AB, st_.beq, st_.q, S, st_.ws); e1.record(); torch.cuda.synchronize()
    print("fixed", iters, "iters:", e0.elapsed_time(e1), "ms for", a.batch, "QPs", flush=True)
S = _lib.AdmmSettings.reference_defaults(max_iter=0, check_every=0, adaptive_rho=0, warm_x=1, warm_y=1)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); me.admm_solve(ctl._qp, st_.AB, st_.beq, st_.q, S, st_.ws); e1.record(); torch.cuda.synchronize()
print("setup only (scaling+factor+final checks):", e0.elapsed_time(e1), "ms", flush=True)

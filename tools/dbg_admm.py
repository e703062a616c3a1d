"""Debug aid (not part of the product): new ADMM kernel vs the second generation after a fixed, small number of iterations."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench
from gp_mpc_rocket_landing_b200 import _lib
from gp_mpc_rocket_landing_b200.mpc import engine as me
class A: pass
a = A(); a.batch = int(os.environ.get("B", 64)); a.sqp_iters = 1; a.qp_max_iter = 1000
dyn, gp, ctl, mc = bench.build_problem(a, 0)
for k in range(3):
    mc.step()
st = ctl._batch; ws = st.ws; qp = ctl._qp
x0, y0, r0 = ws.x.clone(), ws.y.clone(), ws.rho.clone()
def solve(variant, **kw):
    S = _lib.AdmmSettings.reference_defaults(max_iter=1000, warm_x=1, warm_y=1, rho_from_input=1, kernel_variant=variant)
    for k, v in kw.items(): setattr(S, k, v)
    ws.x.copy_(x0); ws.y.copy_(y0); ws.rho.copy_(r0)
    me.admm_solve(qp, st.AB, st.beq, st.q, S, ws); torch.cuda.synchronize()
    return ws.x.cpu().numpy().copy(), ws.y.cpu().numpy().copy(), ws.status.cpu().numpy().copy(), ws.pri.cpu().numpy().copy()
for it in (0, 1, 2, 3, 10, 26, 60):
    xa, ya, sa, pa = solve(2, max_iter=it, check_every=0, adaptive_rho=0)
    xb, yb, sb, pb = solve(int(os.environ.get("VARIANT", 0)), max_iter=it, check_every=0, adaptive_rho=0)
    nan = np.isnan(xb).any(axis=1)
    dx = np.abs(xa - xb).max(axis=1); dy = np.abs(ya - yb).max(axis=1)
    print(f"iters {it:3d}: nan QPs {int(nan.sum())}/{len(nan)}  max|dx| {np.nanmax(dx):.3e}  max|dy| {np.nanmax(dy):.3e}  status {np.unique(sa)} {np.unique(sb)}", flush=True)
    if it == 1:
        b = int(np.argmax(np.where(np.isnan(dx), np.inf, dx)))
        d = np.abs(xa[b] - xb[b])
        print("   worst QP", b, "worst vars", np.argsort(-np.where(np.isnan(d), np.inf, d))[:12], "diffs", np.sort(np.where(np.isnan(d), np.inf, d))[::-1][:6])

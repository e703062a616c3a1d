"""Phase-latency trace of the third-generation ADMM kernel (needs a library built with -DADMM3_TRACE; not part of the product)."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench
from gp_mpc_rocket_landing_b200 import _lib
from gp_mpc_rocket_landing_b200.mpc import engine as me
class A: pass
a = A(); a.batch = int(os.environ.get("B", 4096)); a.sqp_iters = 1; a.qp_max_iter = 1000
dyn, gp, ctl, mc = bench.build_problem(a, 0)
for k in range(3):
    mc.step()
st_ = ctl._batch
lib = _lib.lib
lib.gpmpc_admm3_trace_read.restype = C.c_int
buf = (C.c_longlong * 4096)()
torch.cuda.synchronize()
lib.gpmpc_admm3_trace_read(buf, 4096)   # clear
S = _lib.AdmmSettings.reference_defaults(max_iter=int(os.environ.get("ITERS", 30)), check_every=0, adaptive_rho=0, warm_x=1, warm_y=1)
me.admm_solve(ctl._qp, st_.AB, st_.beq, st_.q, S, st_.ws)
torch.cuda.synchronize()
n = lib.gpmpc_admm3_trace_read(buf, 4096)
ev = [(buf[i] >> 8, buf[i] & 255) for i in range(n)]
names = {0: "iter-start(L2 of prev + L1)", 1: "rhs->P1 barrier", 2: "P1", 3: "C1", 4: "P2", 5: "P3", 6: "P4", 7: "C2", 8: "P5", 20: "setup(load+scale)", 21: "factorize", 30: "start", 31: "load", 32: "init regs (Pdiag,q loads)", 33: "ruiz iter", 34: "ruiz done", 22: "fact phase1 (M rows)", 23: "fact chains", 24: "fact middle", 40: "loop end", 41: "final"}
import collections
acc = collections.defaultdict(list)
for (t0, i0), (t1, i1) in zip(ev[:-1], ev[1:]):
    acc[(i0, i1)].append(t1 - t0)
for k, v in sorted(acc.items()):
    print(f"{k[0]:3d}->{k[1]:3d}  n={len(v):4d}  median {int(np.median(v)):7d}  min {min(v):7d}  max {max(v):7d}   (ends: {names.get(k[1], '')})")

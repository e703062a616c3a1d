"""GPU diagnostic (not part of the product): phase times of one ExactSweep.fit (Gram build / batched potrf / TRSM+LML)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gp_mpc_rocket_landing_b200.gp import engine as E, refit
from gp_mpc_rocket_landing_b200._lib import call, ptr, stream

for n in (1000, 4000, 8000):
    rng = np.random.default_rng(1)
    X = rng.normal(0, 0.4, (n, 11)); Y = rng.normal(size=(n, 3))
    ls, sv, nz = refit.default_grid(11)
    G = 16
    sw = E.ExactSweep(0, X, Y)
    npad = sw.n_pad
    L = torch.empty((G, npad, npad), dtype=torch.float64, device="cuda")
    Winv = torch.empty((G, npad // 64, 64, 64), dtype=torch.float64, device="cuda")
    info = torch.empty((G,), dtype=torch.int32, device="cuda")
    def ev(): return torch.cuda.Event(enable_timing=True)
    for rep in range(2):
        e = [ev() for _ in range(4)]
        e[0].record()
        for g in range(G): sw._build(L, g, ls[g], sv[g], nz[g])
        e[1].record()
        call("gpmpc_potrf", ptr(L), npad, npad, npad * npad, G, ptr(Winv), ptr(info), stream())
        e[2].record()
        r = sw.fit(ls[:1], sv[:1], nz[:1])
        e[3].record(); torch.cuda.synchronize()
    fl = G * npad ** 3 / 3
    print(n, "gram %.2f ms  potrf %.2f ms (%.1f TF/s)  one-candidate fit %.2f ms" % (
        e[0].elapsed_time(e[1]), e[1].elapsed_time(e[2]), fl / e[1].elapsed_time(e[2]) / 1e9, e[2].elapsed_time(e[3])), flush=True)
    del L, Winv, sw
    torch.cuda.empty_cache()

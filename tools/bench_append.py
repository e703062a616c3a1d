"""Exact-GP online update: extend the factor by m new rows (ExactGP.update) against the reference's scheme, a full refit
on the augmented data (online_update.py:383-391).  Prints one JSON line per n."""
import json, sys, time
import numpy as np
sys.path.insert(0, ".")
from gp_mpc_rocket_landing_b200.gp import ExactGP, SquaredExponentialARD  # noqa: E402
from gp_mpc_rocket_landing_b200.gp import engine as E  # noqa: E402

torch = E.torch_()
m = 10
for n in (1000, 4000, 8000):
    rng = np.random.default_rng(n)
    X = rng.normal(size=(n + 20 * m, 11)); y = np.sin(X[:, 0]) + 0.1 * rng.normal(size=len(X))
    k = SquaredExponentialARD(11, 1.0, np.full(11, 3.0))
    gp = ExactGP(k, noise_variance=1e-2).fit(X[:n], y[:n])
    gp.update(X[n:n + m], y[n:n + m])            # warm-up
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for j in range(1, 11):
        gp.update(X[n + j * m:n + (j + 1) * m], y[n + j * m:n + (j + 1) * m])
    torch.cuda.synchronize(); t_inc = (time.perf_counter() - t0) / 10
    full = ExactGP(k, noise_variance=1e-2)
    full.fit(X[:n + m], y[:n + m])
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for j in range(1, 6):
        full.fit(X[:n + (j + 1) * m], y[:n + (j + 1) * m])
    torch.cuda.synchronize(); t_full = (time.perf_counter() - t0) / 5
    print(json.dumps({"n": n, "m_new": m, "extend_ms": round(t_inc * 1e3, 3), "refit_ms": round(t_full * 1e3, 3),
                      "speedup": round(t_full / t_inc, 2)}))

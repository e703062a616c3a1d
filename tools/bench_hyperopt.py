"""Hyper-parameter optimisation of the config-1 residual GP (1000 points, 11 features): L-BFGS-B with the device LML
gradient (gpmpc_lml_grad_se) against the reference's scheme (scipy differences the objective: n_params + 1 refits per
gradient, exact_gp.py:357-421).  Prints one JSON line.  Usage: python tools/bench_hyperopt.py [n]"""
import json
import sys
import time

import numpy as np

sys.path.insert(0, ".")
from gp_mpc_rocket_landing_b200 import workloads  # noqa: E402
from gp_mpc_rocket_landing_b200.gp import ExactGP, SquaredExponentialARD  # noqa: E402
from gp_mpc_rocket_landing_b200.gp import engine as E  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
    torch = E.torch_()
    X, U, D = workloads.residual_dataset(n, seed=0)
    Z = E.features(0, X, U).cpu().numpy()
    y = D[:, 0]
    out = {"n": n, "D": Z.shape[1]}
    for name, mode in (("device_gradient", True), ("finite_difference", False)):
        gp = ExactGP(SquaredExponentialARD(Z.shape[1]), noise_variance=1e-2).fit(Z, y)
        nfit = [0]
        fit0 = gp.fit

        def counted(*a, _f=fit0, **k):
            nfit[0] += 1
            return _f(*a, **k)

        gp.fit = counted
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        res = gp.optimize_hyperparameters(n_restarts=1, analytic_gradients=mode)
        torch.cuda.synchronize()
        out[name] = {"seconds": round(time.perf_counter() - t0, 3), "refits": nfit[0], "lbfgs_iterations": int(res["n_iterations"]),
                     "lml": float(res["log_marginal_likelihood"])}
    # cost of one gradient pass next to one refit
    gp = ExactGP(SquaredExponentialARD(Z.shape[1]), noise_variance=1e-2).fit(Z, y)
    for what, fn in (("refit_ms", lambda: gp.fit(Z, y)), ("gradient_ms", gp.log_marginal_likelihood_gradient)):
        fn(); torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(10):
            fn()
        torch.cuda.synchronize()
        out[what] = round((time.perf_counter() - t0) * 100, 3)
    print(json.dumps(out))


if __name__ == "__main__":
    main()

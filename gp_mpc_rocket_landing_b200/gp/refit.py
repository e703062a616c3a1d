"""Batched hyper-parameter refit sweep (BASELINE config 3): the consumer-facing wrapper of `engine.ExactSweep`.

The reference evaluates hyper-parameter candidates one `ExactGP.fit` at a time (exact_gp.py:357-421 inside L-BFGS-B;
learning/hyperparameter_tuner.py:361-395 by finite differences).  `refit_sweep` evaluates a whole grid in one batched
Gram + Cholesky + solve + LML pass and returns the log marginal likelihood of every (candidate, output); with
torch.distributed initialised the grid is scattered over the ranks and the LMLs all-gathered."""
from __future__ import annotations

from typing import Dict, Optional

import numpy as np

from . import engine as E
from .kernels import Kernel


def default_grid(input_dim: int, n_lengthscales: int = 4, noises=(1e-4, 1e-3, 1e-2, 1e-1)):
    """SURVEY 8(d) config 3 grid: shared scalar lengthscale log-uniform in [0.1, 10] x noise in {1e-4..1e-1}; sigma_f^2 = 1."""
    ell = np.logspace(-1, 1, n_lengthscales)
    ls = np.repeat(ell, len(noises))[:, None] * np.ones((1, input_dim))
    nz = np.tile(np.asarray(noises, dtype=np.float64), n_lengthscales)
    return ls, np.ones(ls.shape[0]), nz


def refit_sweep(X, Y, lengthscales, signal_vars, noises, kind: int = E.SE, normalize_y: bool = True,
                max_batch_bytes: float = 64e9, distributed: bool = False) -> Dict[str, np.ndarray]:
    """-> {"lml": (G,P), "jitter": (G,), "best": (P,) argmax per output, "best_shared": argmax of the summed LML}.
    Candidates are processed in chunks whose factors fit `max_batch_bytes` of HBM."""
    ls = np.atleast_2d(np.asarray(lengthscales, dtype=np.float64))
    G = ls.shape[0]
    sv = np.broadcast_to(np.asarray(signal_vars, dtype=np.float64), (G,)).copy()
    nz = np.broadcast_to(np.asarray(noises, dtype=np.float64), (G,)).copy()
    lo, hi = 0, G
    if distributed:
        from .. import distributed as D
        sh = D.shard_scenarios(G)
        lo, hi = sh.first, sh.first + sh.count
    sweep = E.ExactSweep(kind, X, Y, normalize=normalize_y)
    per = 8.0 * sweep.n_pad * sweep.n_pad
    chunk = max(1, int(max_batch_bytes // per))
    lml = np.zeros((hi - lo, sweep.P))
    jit = np.zeros(hi - lo)
    for s in range(lo, hi, chunk):
        e = min(hi, s + chunk)
        r = sweep.fit(ls[s:e], sv[s:e], nz[s:e])
        lml[s - lo:e - lo] = r["lml"].cpu().numpy()
        jit[s - lo:e - lo] = r["jitter"]
    if distributed:
        from .. import distributed as D
        lml = D.gather_records(lml)
        jit = D.gather_records(jit)
    return {"lml": lml, "jitter": jit, "best": np.argmax(lml, axis=0), "best_shared": int(np.argmax(lml.sum(axis=1)))}


def apply_best(kernel: Kernel, lengthscales, signal_vars, index: int) -> Kernel:
    """Install candidate `index` into a reference-API kernel object (kernels.py set via attributes)."""
    kernel.lengthscales = np.asarray(lengthscales, dtype=np.float64)[index].copy()
    kernel.signal_variance = float(np.broadcast_to(signal_vars, (len(lengthscales),))[index])
    return kernel


def refine_shared(X, Y, lengthscales0, signal_var0: float, noise0: float, normalize_y: bool = True, maxiter: int = 100,
                  bounds_log=(-9.0, 9.0)) -> Dict[str, object]:
    """Gradient refinement of ONE hyper-parameter set shared by all P outputs (the model the Monte Carlo driver broadcasts):
    L-BFGS-B on -sum_p LML_p over theta = [log s2, log l_1..l_D, log noise], started from e.g. the best candidate of
    `refit_sweep`.  Every evaluation is one device fit of the shared factor plus one `gpmpc_lml_grad_se` pass that returns
    the gradients of all P outputs -- against n_params + 1 refits per output for the reference's differenced L-BFGS-B
    (exact_gp.py:357-421) or the tuner's finite differences (learning/hyperparameter_tuner.py:361-395).
    -> {"lengthscales", "signal_variance", "noise_variance", "lml" (P,), "n_evaluations", "n_iterations", "success"}."""
    from scipy.optimize import minimize
    Xd = E.to_dev(np.atleast_2d(np.asarray(X, dtype=np.float64)) if not hasattr(X, "is_cuda") else X)
    Yd = E.to_dev(np.asarray(Y, dtype=np.float64) if not hasattr(Y, "is_cuda") else Y)
    D = Xd.shape[1]
    theta0 = np.log(np.concatenate([[float(signal_var0)], np.broadcast_to(np.asarray(lengthscales0, dtype=np.float64), (D,)),
                                    [float(noise0)]]))
    state = {"n": 0, "best": (np.inf, None, None)}

    def objective(theta):
        state["n"] += 1
        sv, ls, nz = float(np.exp(theta[0])), np.exp(theta[1:1 + D]), float(np.exp(theta[1 + D]))
        try:
            f = E.ExactFactor(E.SE, ls, sv, nz).fit(Xd)
        except ValueError:
            return np.inf, np.zeros_like(theta)
        m = E.ExactModelDev(f).fit_targets(Yd, normalize=normalize_y)
        lml = m.lml.cpu().numpy()
        g = m.lml_gradient().sum(dim=0).cpu().numpy()
        if -lml.sum() < state["best"][0]:
            state["best"] = (-lml.sum(), theta.copy(), lml.copy())
        return -float(lml.sum()), -g

    res = minimize(objective, theta0, jac=True, method="L-BFGS-B", bounds=[bounds_log] * len(theta0),
                   options={"maxiter": maxiter})
    _, theta, lml = state["best"]
    return {"lengthscales": np.exp(theta[1:1 + D]), "signal_variance": float(np.exp(theta[0])),
            "noise_variance": float(np.exp(theta[1 + D])), "lml": lml, "n_evaluations": state["n"],
            "n_iterations": int(res.nit), "success": bool(res.success)}


def sweep_and_refine(X, Y, lengthscales=None, signal_vars=None, noises=None, **kw) -> Dict[str, object]:
    """Grid then gradient: `refit_sweep` over the candidates (default: `default_grid`), then `refine_shared` from the
    candidate with the largest summed LML.  Adds "grid_best_index" and "grid_best_lml" to the refinement's result."""
    Xh = np.atleast_2d(X.cpu().numpy() if hasattr(X, "is_cuda") else np.asarray(X, dtype=np.float64))
    if lengthscales is None:
        lengthscales, signal_vars, noises = default_grid(Xh.shape[1])
    sw = refit_sweep(X, Y, lengthscales, signal_vars, noises)
    b = sw["best_shared"]
    sv = np.broadcast_to(np.asarray(signal_vars, dtype=np.float64), (len(lengthscales),))
    nz = np.broadcast_to(np.asarray(noises, dtype=np.float64), (len(lengthscales),))
    out = refine_shared(X, Y, np.asarray(lengthscales)[b], sv[b], nz[b], **kw)
    out["grid_best_index"] = b
    out["grid_best_lml"] = sw["lml"][b]
    return out

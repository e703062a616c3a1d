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

"""Multi-GPU plumbing of the Monte Carlo path: one process per GPU, scenarios block-sharded by index, and
the only two collectives the path has --

  * `reduce_stats`: SUM all-reduce of the outcome counters / per-step sums (tens of bytes, latency class);
  * `broadcast_hyperparameters`: rank 0 sends theta = [log sigma_f^2, log l_1..l_D, log sigma_n^2] per output plus a
    version counter; every rank then refits its own replica (the fit is deterministic, so replicas stay bit-identical).

There is no data-path collective: scenarios never interact (reference src/experiments/monte_carlo.py:606-616).  The
functions take whatever process group is initialised -- NCCL with CUDA tensors on the GPU box, gloo with CPU tensors
in tests/test_distributed.py -- and degrade to the identity when torch.distributed is not initialised (world size 1).
"""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import Optional, Sequence, Tuple

import numpy as np


def _dist():
    import torch.distributed as dist
    return dist


def is_initialized() -> bool:
    d = _dist()
    return d.is_available() and d.is_initialized()


def world() -> Tuple[int, int]:
    """(rank, world_size) from the process group, else from the torchrun environment, else (0, 1)."""
    if is_initialized():
        d = _dist()
        return d.get_rank(), d.get_world_size()
    return int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))


def init_process_group(backend: Optional[str] = None) -> Tuple[int, int]:
    """Initialise from the torchrun environment (RANK / WORLD_SIZE / MASTER_ADDR / MASTER_PORT).  backend defaults to
    nccl when CUDA is present (and binds the process to cuda:LOCAL_RANK), else gloo."""
    import torch
    rank, ws = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    if ws == 1 or is_initialized():
        return world()
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    if backend is None:
        backend = "nccl" if torch.cuda.is_available() else "gloo"
    if backend == "nccl":
        local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(local)
        _dist().init_process_group("nccl", device_id=torch.device("cuda", local))
    else:
        _dist().init_process_group(backend)
    return rank, ws


@dataclass(frozen=True)
class ScenarioShard:
    """Contiguous block [first, first + count) of the global scenario index range owned by one rank."""

    first: int
    count: int
    rank: int
    world_size: int


def shard_scenarios(total: int, rank: Optional[int] = None, world_size: Optional[int] = None) -> ScenarioShard:
    """Block partition with the remainder spread over the first ranks: every index is owned exactly once, and the
    initial condition of scenario i depends only on (seed, i), so results are independent of the world size."""
    r, w = world()
    rank = r if rank is None else rank
    world_size = w if world_size is None else world_size
    base, rem = divmod(int(total), world_size)
    count = base + (1 if rank < rem else 0)
    first = rank * base + min(rank, rem)
    return ScenarioShard(first, count, rank, world_size)


def _as_tensor(a, device):
    import torch
    if isinstance(a, torch.Tensor):
        return a
    return torch.as_tensor(np.asarray(a), device=device)


def _collective_device():
    import torch
    if is_initialized() and _dist().get_backend() == "nccl":
        return torch.device("cuda", torch.cuda.current_device())
    return torch.device("cpu")


def reduce_stats(counts, sums=None):
    """SUM over ranks of the integer outcome counters (and optionally a float vector of sums: fuel, steps, QP
    iterations ...).  Tensors are reduced in place and returned; numpy inputs come back as numpy."""
    import torch
    dev = _collective_device()
    outs = []
    for a in (counts, sums):
        if a is None:
            outs.append(None)
            continue
        was_np = not isinstance(a, torch.Tensor)
        t = _as_tensor(a, dev)
        if is_initialized() and t.device != dev:
            t = t.to(dev)
        if is_initialized():
            _dist().all_reduce(t, op=_dist().ReduceOp.SUM)
        outs.append(t.cpu().numpy() if was_np else t)
    return outs[0] if sums is None else tuple(outs)


class AsyncStatsReducer:
    """The per-step statistics gather of the Monte Carlo path (north_star: "NCCL ... only to gather per-step statistics"):
    a SUM all-reduce of the rank's cumulative counter vector (8 int64: runs ended per outcome, solves, ADMM iterations),
    issued every control step on a SIDE stream with device tensors only -- no `.item()`, no numpy round trip, and the
    main stream never waits for it.  `total()` is the latest global vector (device tensor, valid after `wait()`).
    With CPU tensors (gloo tests) or without a process group it degrades to a synchronous reduce / a copy."""

    def __init__(self, counts):
        import torch
        self.src = counts
        self.cuda = counts.is_cuda
        self.buf = torch.zeros_like(counts)
        self.glob = torch.zeros_like(counts)
        self.side = torch.cuda.Stream() if self.cuda else None
        self.n = 0
        self._timed = []

    def reset(self):
        self.n = 0
        self._timed = []

    def submit(self):
        import torch
        self.n += 1
        if not self.cuda:
            self.buf.copy_(self.src)
            if is_initialized():
                _dist().all_reduce(self.buf, op=_dist().ReduceOp.SUM)
            self.glob.copy_(self.buf)
            return
        ev = torch.cuda.Event()
        ev.record()                                   # after this step's counters were written on the main stream
        with torch.cuda.stream(self.side):
            self.side.wait_event(ev)
            timed = self.n <= 64
            if timed:
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record(self.side)
            self.buf.copy_(self.src, non_blocking=True)
            if is_initialized():
                _dist().all_reduce(self.buf, op=_dist().ReduceOp.SUM)
            self.glob.copy_(self.buf, non_blocking=True)
            if timed:
                b.record(self.side)
                self._timed.append((a, b))

    def wait(self):
        if self.cuda:
            import torch
            torch.cuda.current_stream().wait_stream(self.side)

    def total(self):
        self.wait()
        return self.glob

    def device_ms(self) -> float:
        """Device time of the side-stream work (copy + all-reduce + copy), extrapolated from the first 64 timed steps."""
        if not self._timed:
            return 0.0
        import torch
        torch.cuda.synchronize()
        t = sum(a.elapsed_time(b) for a, b in self._timed)
        return t / len(self._timed) * self.n


def gather_records(local: np.ndarray) -> np.ndarray:
    """All-gather of per-scenario terminal records (rows) in rank order -> the global array on every rank.  Shards may
    have different lengths (remainder ranks)."""
    local = np.ascontiguousarray(local)
    if not is_initialized():
        return local
    parts = [None] * _dist().get_world_size()
    _dist().all_gather_object(parts, local)
    return np.concatenate(parts, axis=0)


# ---------------------------------------------------------------------------------------------------- hyper-parameters
def _gps_of(model) -> Sequence:
    """The single-output GPs of a Simple3DoFGP / MultiOutput*GP / ExactGP / SparseGP."""
    inner = getattr(model, "gp", model)
    return list(getattr(inner, "gps", [inner]))


def pack_hyperparameters(model) -> np.ndarray:
    """(P, n_kernel_params + 1): kernel.get_params() (log-space, reference kernels.py:341-371) + log noise_variance --
    the parameter vector ExactGP.optimize_hyperparameters works on (reference exact_gp.py:375-377)."""
    rows = [np.concatenate([np.asarray(g.kernel.get_params(), dtype=np.float64), [np.log(g.noise_variance)]])
            for g in _gps_of(model)]
    return np.stack(rows)


def apply_hyperparameters(model, theta: np.ndarray) -> None:
    """Inverse of `pack_hyperparameters`; invalidates the fitted state (the caller refits)."""
    gps = _gps_of(model)
    theta = np.asarray(theta, dtype=np.float64).reshape(len(gps), -1)
    for g, row in zip(gps, theta):
        g.kernel.set_params(row[:-1])
        g.noise_variance = float(np.exp(row[-1]))
    if hasattr(model, "_is_fitted"):
        model._is_fitted = False


def broadcast_hyperparameters(model, version: int = 0, src: int = 0, refit: bool = True) -> int:
    """Rank `src` publishes its model's theta and `version`; every other rank installs them and (optionally) refits.
    Returns the version now held by every rank.  Message = P*(D+2)+1 doubles."""
    import torch
    theta = pack_hyperparameters(model)
    if not is_initialized():
        return int(version)
    dev = _collective_device()
    buf = torch.empty(theta.size + 1, dtype=torch.float64, device=dev)
    if _dist().get_rank() == src:
        buf[:-1] = torch.from_numpy(theta.ravel()).to(dev)
        buf[-1] = float(version)
    _dist().broadcast(buf, src=src)
    host = buf.cpu().numpy()
    if _dist().get_rank() != src:
        apply_hyperparameters(model, host[:-1].reshape(theta.shape))
        if refit and hasattr(model, "fit") and getattr(model, "n_data", 1) > 0:
            model.fit()
    return int(round(host[-1]))

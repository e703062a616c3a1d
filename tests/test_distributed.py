"""CPU: the N>1 path of the Monte Carlo driver under `gloo`, world_size 2 -- scenario sharding (independent of the world
size), the stats all-reduce, the terminal-record gather and the hyper-parameter broadcast.  No compute
kernels run here (no GPU): only the host logic that surrounds them."""
import multiprocessing as mp
import os
import socket

import numpy as np

from conftest import ROOT


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world_size, port, q):
    import sys
    sys.path.insert(0, ROOT)
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world_size), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port),
                      LOCAL_RANK=str(rank))
    import torch
    from gp_mpc_rocket_landing_b200 import distributed as D, workloads
    from gp_mpc_rocket_landing_b200.gp import Simple3DoFGP
    out = {}
    try:
        r, w = D.init_process_group("gloo")
        assert (r, w) == (rank, world_size) and D.world() == (rank, world_size)
        # --- sharding: 11 scenarios over 2 ranks -> 6 + 5, ICs depend on (seed, index) only
        sh = D.shard_scenarios(11)
        X0, XT = workloads.sample_initial_conditions(sh.first, sh.count, seed=42)
        allX0 = D.gather_records(X0)
        out["shard"] = (sh.first, sh.count)
        out["X0_all"] = allX0
        # --- stats: integer counters (tensor, in place) and float sums (numpy in -> numpy out)
        counts = torch.zeros(8, dtype=torch.int64)
        counts[1 + rank] = 3 + rank
        sums = np.array([1.5 * (rank + 1), 10.0])
        c, s = D.reduce_stats(counts, sums)
        out["counts"] = c.numpy().copy()
        out["sums"] = s
        # --- the per-step statistics reducer the timed loop uses (bench.py): cumulative counters, one reduce per step
        loc = torch.zeros(8, dtype=torch.int64)
        red = D.AsyncStatsReducer(loc)
        for step in range(3):
            loc[0] += 10 * (rank + 1)
            loc[1 + rank] += 1
            red.submit()
        out["stream_total"] = red.total().numpy().copy()
        # --- hyper-parameter broadcast: rank 0 holds refit values, rank 1 the defaults
        gp = Simple3DoFGP(use_sparse=False, noise_variance=1e-4)
        if rank == 0:
            for i, g in enumerate(gp.gp.gps):
                g.kernel.set_params(np.log(np.r_[0.5 + i, np.linspace(0.3, 2.0, 11) + i]))
                g.noise_variance = 10.0 ** (-3 + i)
        ver = D.broadcast_hyperparameters(gp, version=7 if rank == 0 else 0, refit=False)
        out["version"] = ver
        out["theta"] = D.pack_hyperparameters(gp)
        torch.distributed.barrier()
        torch.distributed.destroy_process_group()
        out["ok"] = True
    except Exception as e:  # noqa: BLE001
        import traceback
        out["ok"] = False
        out["err"] = traceback.format_exc() + repr(e)
    q.put((rank, out))


def test_world_size_2_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=240) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    for r in (0, 1):
        assert res[r]["ok"], res[r].get("err")
    assert res[0]["shard"] == (0, 6) and res[1]["shard"] == (6, 5)
    from gp_mpc_rocket_landing_b200 import workloads
    X0, _ = workloads.sample_initial_conditions(0, 11, seed=42)
    for r in (0, 1):
        np.testing.assert_array_equal(res[r]["X0_all"], X0)          # shards concatenate to the unsharded batch
        np.testing.assert_array_equal(res[r]["counts"], [0, 3, 4, 0, 0, 0, 0, 0])
        np.testing.assert_allclose(res[r]["sums"], [4.5, 20.0])
        np.testing.assert_array_equal(res[r]["stream_total"], [90, 3, 3, 0, 0, 0, 0, 0])
        assert res[r]["version"] == 7
    np.testing.assert_array_equal(res[0]["theta"], res[1]["theta"])  # replicas hold bit-identical hyper-parameters
    assert res[0]["theta"].shape == (3, 13)
    np.testing.assert_allclose(np.exp(res[1]["theta"][2, -1]), 1e-1)


def test_single_process_identity():
    """Without a process group every collective is the identity and the shard is the whole range."""
    from gp_mpc_rocket_landing_b200 import distributed as D
    for k in ("RANK", "WORLD_SIZE"):
        os.environ.pop(k, None)
    sh = D.shard_scenarios(4096)
    assert (sh.first, sh.count, sh.world_size) == (0, 4096, 1)
    parts = [D.shard_scenarios(65536 + 3, r, 8) for r in range(8)]
    assert parts[0].first == 0 and sum(p.count for p in parts) == 65539
    assert all(parts[i].first + parts[i].count == parts[i + 1].first for i in range(7))
    c = D.reduce_stats(np.arange(8))
    np.testing.assert_array_equal(c, np.arange(8))
    np.testing.assert_array_equal(D.gather_records(np.ones((3, 2))), np.ones((3, 2)))
    import torch
    loc = torch.arange(8)
    red = D.AsyncStatsReducer(loc)
    red.submit()
    np.testing.assert_array_equal(red.total().numpy(), np.arange(8))

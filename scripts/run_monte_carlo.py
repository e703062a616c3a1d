#!/usr/bin/env python
"""Batched Monte Carlo driver (the reference's scripts/run_monte_carlo.py is a 2-line placeholder; the sequential loop
it stands for is MonteCarloSimulator.run, src/experiments/monte_carlo.py:585-636).

    python scripts/run_monte_carlo.py --scenarios 4096 --controller gpmpc
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 scripts/run_monte_carlo.py --scenarios 65536

Every rank simulates its block of the scenario index range to termination on its own GPU; rank 0 prints the pooled
statistics (outcome counts via one all-reduce, terminal records via one gather at the end).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--scenarios", type=int, default=4096, help="total scenarios over all ranks")
    ap.add_argument("--controller", default="gpmpc", choices=["gpmpc", "nominal", "rti"])
    ap.add_argument("--horizon", type=int, default=20)
    ap.add_argument("--n-train", type=int, default=1000, help="GP training points (synthetic drag residual)")
    ap.add_argument("--seed", type=int, default=42)
    ap.add_argument("--max-steps", type=int, default=None)
    ap.add_argument("--drag-cda", type=float, default=0.25)
    ap.add_argument("--rebroadcast-every", type=int, default=0, help="broadcast GP hyper-parameters every K steps")
    ap.add_argument("--sqp-iters", type=int, default=1, help="SQP iterations per solve (1 = RTI; gp_mpc.py:293 uses 10)")
    ap.add_argument("--target-mode", default=None, choices=["per_step", "fixed"],
                    help="default: what monte_carlo.py:492-510 does for the controller type (`.solve` -> per_step)")
    ap.add_argument("--on-solve-failure", default="continue", choices=["continue", "divergence"],
                    help="divergence = monte_carlo.py:507-510 (with an RTI every run then ends at step 0, DESIGN 5b)")
    ap.add_argument("--convex-rows", action="store_true", help="thrust-cone / glide-slope rows in the QP (DESIGN 11g)")
    ap.add_argument("--safety-filter", action="store_true", help="predictive safety filter on u0 (monte_carlo.py:519-525)")
    a = ap.parse_args()

    from gp_mpc_rocket_landing_b200 import distributed as D, workloads
    from gp_mpc_rocket_landing_b200.dynamics import Rocket3DoFDynamics
    from gp_mpc_rocket_landing_b200.gp import Simple3DoFGP
    from gp_mpc_rocket_landing_b200.monte_carlo import OUTCOME_NAMES, BatchedMonteCarlo
    from gp_mpc_rocket_landing_b200.mpc import GPMPC, FastRTI3DoF, GPMPCConfig, OSQPRTIConfig

    rank, world = D.init_process_group()
    shard = D.shard_scenarios(a.scenarios)
    dyn = Rocket3DoFDynamics()
    gp = None
    if a.controller == "gpmpc":
        X, U, Dres = workloads.residual_dataset(a.n_train, seed=0)
        gp = Simple3DoFGP(noise_variance=1e-4, use_sparse=False)
        gp.add_data(X, U, Dres)
        gp.fit()
        D.broadcast_hyperparameters(gp, version=1)       # replicas start from rank 0's hyper-parameters
    if a.controller == "rti":
        ctl = FastRTI3DoF(dyn, OSQPRTIConfig(N=a.horizon, dt=0.1, osqp_max_iter=1000))
    else:
        ctl = GPMPC(dyn, gp, GPMPCConfig(N=a.horizon, dt=0.1, max_sqp_iter=a.sqp_iters, qp_max_iter=1000,
                                         convex_constraints=a.convex_rows))
    filt = None
    if a.safety_filter:
        from gp_mpc_rocket_landing_b200.safety import PredictiveSafetyFilter, SafetyFilterConfig
        from gp_mpc_rocket_landing_b200.mpc import ConstraintParams
        filt = PredictiveSafetyFilter(dyn, SafetyFilterConfig(N=10, dt=0.1), constraint_params=ConstraintParams())
        p = dyn.params
        filt.initialize(np.array([1.5, 0.5, 0, 0, 0, 0, 0.0]), np.array([1.5 * p.g0, 0, 0.0]), backup_alpha=50.0)
    mc = BatchedMonteCarlo(dyn, ctl, drag_cda=a.drag_cda, target_mode=a.target_mode,
                           on_solve_failure=None if a.controller == "rti" else a.on_solve_failure, safety_filter=filt)
    mc.reset(shard.first, shard.count, a.seed)
    limit = mc.max_steps if a.max_steps is None else a.max_steps
    t0 = time.perf_counter()
    steps = 0
    for k in range(limit):
        mc.step()
        steps += 1
        if a.rebroadcast_every and gp is not None and (k + 1) % a.rebroadcast_every == 0:
            D.broadcast_hyperparameters(gp, version=2 + k // a.rebroadcast_every, refit=False)
        if k % 10 == 9:
            left = D.reduce_stats(np.array([mc.n_active()]))     # stop together: every rank sees the global count
            if int(left[0]) == 0:
                break
    res = mc.results()
    wall = time.perf_counter() - t0
    counts = D.reduce_stats(np.array([res.counts.get(n, 0) for n in OUTCOME_NAMES[1:]], dtype=np.int64))
    rec = D.gather_records(np.column_stack([res.outcome, res.n_steps, res.fuel_used, res.final_state]))
    if rank == 0:
        ok = rec[:, 0] == 1
        done = rec[:, 0] != 0
        print(json.dumps({
            "scenarios": int(rec.shape[0]), "ranks": world, "controller": a.controller, "control_steps": steps,
            "wall_s": wall, "outcomes": {n: int(c) for n, c in zip(OUTCOME_NAMES[1:], counts)},
            "success_rate": float(ok.sum() / max(1, done.sum())),
            "fuel_mean_success": float(rec[ok, 2].mean()) if ok.any() else None,
            "landing_error_mean": float(np.linalg.norm(rec[ok][:, 5:7], axis=1).mean()) if ok.any() else None,
        }))
    if D.is_initialized():
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

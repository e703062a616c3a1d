#!/usr/bin/env python
"""bench.py -- GP-MPC solves/sec (3-DoF, N=20, 1000-point exact GP, B=4096 scenarios per GPU) on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--batch B] [--sqp-iters S]

A "step" is one closed-loop control step of the Monte Carlo batch: `GPMPC.solve_batch` for every local scenario
(GP rollout + GP variance + covariance propagation + S x (QP assembly + ADMM + update)), the plant step and the
termination / respawn bookkeeping.  A "solve" is one scenario x one control step.  `value` is timed with the scenario
states resident in HBM; `e2e` repeats the loop through the public numpy API (host buffers in, host results out).
`--impl reference` times the CPU restatement of the same path (oracle/, the reference's algorithm on numpy/scipy --
the reference's own GPMPC cannot run: CasADi/IPOPT/simdyn are absent) on all host cores.
See DESIGN.md "Measurement" for the roofline arithmetic.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

N_HORIZON, DT, N_TRAIN, NOISE = 20, 0.1, 1000, 1e-4
# dram__bytes_read.sum + dram__bytes_write.sum per step from the committed `ncu --set full` captures (profiles/), B = 4096
NCU_TRAFFIC = {"prologue": 2.2e6, "rollout": 740.6e6, "gp_variance": 691.6e6, "cov": 17.7e6, "assemble": 26.8e6,
               "admm": 83.2e6, "update": 18.5e6}       # profiles/r2_step_kernels_full_v2.md (one closed-loop step)
METRIC = "GP-MPC solves/sec (3-DoF, N=20, 1000-pt exact GP)"
CPU_GP_LEG = None     # which GP code the CPU legs ran (set by cpu_reference_run)
CPU_CTL_LEG = None    # which controller loop the CPU legs ran


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)   # 200 steps after 20 warm-up
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=4096, help="scenarios per GPU (weak scaling)")
    ap.add_argument("--sqp-iters", type=int, default=1, help="SQP iterations per solve (1 = RTI; reference loop: 10)")
    ap.add_argument("--qp-max-iter", type=int, default=1000)
    ap.add_argument("--cpu-scenarios", type=int, default=0, help="reference arm: scenarios per step (default: #cores)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--convex-rows", action="store_true",
                    help="solve the QP with the polyhedral thrust-cone / glide-slope rows (GPMPCConfig.convex_constraints)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--admm-variant", type=int, default=0, help="GpmpcAdmmSettings.kernel_variant (0 newest, 2 / 1 older)")
    ap.add_argument("--no-extra", action="store_true", help="skip the short S=10 / config-3 / config-4 lines")
    ap.add_argument("--target-mode", default="per_step", choices=["per_step", "fixed"],
                    help="per_step: the reference's Monte Carlo target for `.solve` controllers (monte_carlo.py:497-499)")
    ap.add_argument("--fail-policy", default="continue", choices=["continue", "divergence"],
                    help="divergence: a solve with success == False ends the run (monte_carlo.py:507-510)")
    ap.add_argument("--theta-every", type=int, default=100,
                    help="broadcast the GP hyper-parameters and refit every K steps inside the timed loop (0: never)")
    return ap.parse_args()


# ====================================================================================================== CPU baseline
def _cpu_worker(idx, conn, fits_blob, warmup, steps, sqp_iters, qp_max_iter):
    """One host process = one scenario in closed loop on the oracle (numpy/scipy restatement of the reference path)."""
    from threadpoolctl import threadpool_limits
    threadpool_limits(1)   # one BLAS thread per process: the parallelism is one process per scenario
    from gp_mpc_rocket_landing_b200 import workloads
    from oracle import gp as ogp, mpc as ompc
    if fits_blob[0] == "reference":     # the reference's own fitted Simple3DoFGP (oracle/_ref), inherited through fork
        gp = fits_blob[1]
    else:
        gp = ogp.Simple3DoFGPOracle(noise_variance=NOISE)
        gp.X, gp.U, gp.D, gp.fits = fits_blob[1]
    gm = None
    if sqp_iters == 10 and fits_blob[0] == "reference":
        # the reference's SQP loop (gp_mpc.py:293 hard-codes 10 iterations): run the reference's OWN GPMPC.solve out of
        # oracle/_ref, with only the absent third-party parts replaced (oracle/ref_controller.py)
        from oracle import ref_controller, ref_gp
        gm = ref_gp.load_gpmpc()
    if gm is not None:
        rc = ref_controller.ReferenceGPMPC(gm, gp.predict, N=N_HORIZON, dt=DT, settings=ompc.OSQPSettings(max_iter=qp_max_iter, linsys="sparse"))

        class _Ctl:
            @staticmethod
            def solve(x, xt):
                u0, _, _ = rc.solve(x, xt)
                return type("R", (), {"U_opt": u0[None, :]})
        ctl = _Ctl
    else:
        ctl = ompc.GPMPCOracle(gp.predict, N=N_HORIZON, dt=DT, settings=ompc.OSQPSettings(max_iter=qp_max_iter, linsys="sparse"),
                               max_sqp_iter=sqp_iters)
    X0, XT = workloads.sample_initial_conditions(idx, 1)
    x = X0[0]
    p = ompc.Rocket3DoFParams()

    def one():   # one control step of monte_carlo.py:455-549 for a `.solve` controller (per-step target), drag plant
        nonlocal x
        xt = x.copy(); xt[4:7] = 0; xt[1] = max(0.5, x[1] - 2.0)
        r = ctl.solve(x, xt)
        xn = ompc.step_3dof(x, r.U_opt[0], DT, p)
        rho = 1.225 * np.exp(-x[1] / 8500.0); sp = np.sqrt(x[4] ** 2 + x[5] ** 2 + x[6] ** 2)
        xn[4:7] = xn[4:7] + DT * (-0.5 * rho * 0.25 * sp * x[4:7] / x[0])
        x = xn

    for _ in range(warmup):
        one()
    conn.send("ready")
    conn.recv()
    t0 = time.perf_counter()
    for _ in range(steps):
        one()
    conn.send(time.perf_counter() - t0)
    conn.close()


def cpu_reference_run(n_scen, warmup, steps, sqp_iters, qp_max_iter):
    """Closed loop of `n_scen` scenarios, one process each (scenarios are independent), `steps` control steps.
    Returns (solves/s, wall seconds, cores used)."""
    import multiprocessing as mp
    from gp_mpc_rocket_landing_b200 import workloads
    from oracle import gp as ogp, ref_gp
    X, U, D = workloads.residual_dataset(N_TRAIN, seed=0)
    ref = ref_gp.load()
    if ref is not None:     # GP leg = the reference's own src/gp code (exact_gp.py, structured_gp.py), fitted once here
        gp = ref.Simple3DoFGP(use_sparse=False, noise_variance=NOISE)
        gp.add_data(X, U, D)
        gp.fit()
        blob = ("reference", gp)
    else:
        gp = ogp.Simple3DoFGPOracle(noise_variance=NOISE)
        gp.add_data(X, U, D)
        gp.fit()
        blob = ("port", (gp.X, gp.U, gp.D, gp.fits))
    global CPU_GP_LEG, CPU_CTL_LEG
    CPU_CTL_LEG = ("reference GPMPC.solve (oracle/_ref, unmodified; _solve_qp -> OSQP-form QP + OSQP restatement, dynamics -> "
                   "in-repo Euler statement)" if (sqp_iters == 10 and ref is not None and ref_gp.load_gpmpc() is not None)
                   else "numpy restatement oracle/mpc.py::GPMPCOracle")
    CPU_GP_LEG = ("reference src/gp (oracle/_ref: Simple3DoFGP / ExactGP, unmodified)" if ref is not None
                  else "numpy restatement oracle/gp.py (oracle/_ref absent)")
    ctx = mp.get_context("fork")
    procs = []
    for i in range(n_scen):
        a, b = ctx.Pipe()
        pr = ctx.Process(target=_cpu_worker, args=(i, b, blob, warmup, steps, sqp_iters, qp_max_iter))
        pr.start()
        procs.append((pr, a))
    for _, a in procs:
        a.recv()
    t0 = time.perf_counter()
    for _, a in procs:
        a.send("go")
    for _, a in procs:
        a.recv()
    wall = time.perf_counter() - t0
    for pr, _ in procs:
        pr.join()
    return n_scen * steps / wall, wall, n_scen


def config_dict(args, world):
    return {"workload": "BASELINE configs[1]: 3-DoF GP-MPC Monte Carlo batch, perturbed initial states "
                        "(sample_initial_condition seed 42+i), closed loop with respawn",
            "scenarios_per_gpu": args.batch, "scenarios_total": args.batch * world, "horizon_N": N_HORIZON, "dt": DT,
            "gp": f"Simple3DoFGP exact, n={N_TRAIN} synthetic residual points, SE-ARD, noise {NOISE}",
            "sqp_iters_per_solve": args.sqp_iters, "qp": f"OSQP-form QP n=207 m=354, ADMM max_iter {args.qp_max_iter}, "
            "eps 1e-4, warm start", "parallelism": f"scenario sharding x{world} (no data-path collective)",
            "l2": "working set per step (K* rows 0.67 GB + QP state) exceeds the 126 MB L2; no explicit flush"}


def run_reference_arm(args):
    """`--impl reference`: the CPU restatement of the path on all host cores; rank 0 only."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    cores = os.cpu_count() or 1
    n_scen = args.cpu_scenarios or max(1, min(cores, 64))
    # bound the run: one oracle solve costs ~0.25 s per core and step; keep the whole arm within a few minutes
    steps, warmup = min(max(1, args.steps), 20), min(max(1, args.warmup), 10)
    sps, wall, used = cpu_reference_run(n_scen, warmup, steps, args.sqp_iters, args.qp_max_iter)
    line = {
        "impl": "reference", "metric": METRIC, "value": sps, "unit": "solves/s", "n_gpus": args.gpus, "steps": steps,
        "warmup": warmup, "ms_per_step": wall / steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": config_dict(args, args.gpus),
        "cpu_baseline": {"value": sps, "unit": "solves/s", "cores": used, "kind": "port",
                         "sample": f"{n_scen} scenarios x {steps} closed-loop control steps, one process per scenario "
                                   "(numpy/scipy oracle of the reference algorithm; the reference's own GPMPC needs "
                                   "CasADi/IPOPT/simdyn, absent)", "gp_leg": CPU_GP_LEG, "controller_leg": CPU_CTL_LEG},
        "e2e": {"value": sps, "unit": "solves/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ====================================================================================================== clocks
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.path = index, None, None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.index)], stdout=open(self.path, "w"),
                                         stderr=subprocess.DEVNULL)
        except Exception:  # noqa: BLE001
            self.proc = None

    def mark(self):
        """Start of the timed region: samples written from here on are the ones reported."""
        try:
            self.offset = os.path.getsize(self.path)
        except Exception:  # noqa: BLE001
            self.offset = 0

    def stop(self):
        if self.proc is None:
            return None
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:  # noqa: BLE001
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        text = open(self.path).read()
        timed = text[getattr(self, "offset", 0):]
        window = "timed region"
        if len([ln for ln in timed.splitlines() if ln.count(",") >= 8]) == 0:
            timed, window = text, "warm-up + timed region (the timed region was shorter than one 100 ms sample)"
        for ln in timed.splitlines():
            f = [t.strip() for t in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        os.unlink(self.path)
        if not sm:
            return None
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons),
                "samples": len(sm), "window": window}


# ====================================================================================================== GPU arm
def fp64_peak_tflops(torch):
    """cuBLAS DGEMM 8192^3 (best of 5): the FP64 tensor ceiling for the DMMA kernels.  MEASURED_PEAKS.json holds only
    HBM GB/s and bf16 TF/s, so the FP64 denominator is measured here, in the same run."""
    n = 8192
    a = torch.randn((n, n), dtype=torch.float64, device="cuda")
    b = torch.randn((n, n), dtype=torch.float64, device="cuda")
    torch.matmul(a, b)
    best = 1e9
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); torch.matmul(a, b); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    del a, b
    return 2.0 * n**3 / (best * 1e-3) / 1e12


def build_problem(args, rank):
    from gp_mpc_rocket_landing_b200 import workloads
    from gp_mpc_rocket_landing_b200.dynamics import Rocket3DoFDynamics
    from gp_mpc_rocket_landing_b200.gp import Simple3DoFGP
    from gp_mpc_rocket_landing_b200.monte_carlo import BatchedMonteCarlo
    from gp_mpc_rocket_landing_b200.mpc import GPMPC, GPMPCConfig
    X, U, D = workloads.residual_dataset(N_TRAIN, seed=0)
    gp = Simple3DoFGP(noise_variance=NOISE, use_sparse=False)
    gp.add_data(X, U, D)
    gp.fit()
    dyn = Rocket3DoFDynamics()
    cfg = GPMPCConfig(N=N_HORIZON, dt=DT, max_sqp_iter=args.sqp_iters, qp_max_iter=args.qp_max_iter,
                      qp_kernel_variant=int(getattr(args, "admm_variant", 0)),
                      convex_constraints=bool(getattr(args, "convex_rows", False)))
    cp = None
    if getattr(args, "gamma_gs", None):
        from gp_mpc_rocket_landing_b200.mpc import ConstraintParams
        cp = ConstraintParams(gamma_gs=float(args.gamma_gs))
    ctl = GPMPC(dyn, gp, cfg, constraint_params=cp)
    ctl.setup()
    # the reference's loop for a `.solve` controller (per-step target, monte_carlo.py:497-499); a failed SQP step-size
    # test is counted, not fatal (an RTI with S = 1 never meets it; `--fail-policy divergence` = monte_carlo.py:507-510)
    mc = BatchedMonteCarlo(dyn, ctl, drag_cda=0.25, respawn=True, target_mode=getattr(args, "target_mode", "per_step"),
                           on_solve_failure=getattr(args, "fail_policy", "continue"))
    mc.reset(first=rank * args.batch, count=args.batch, seed=42)
    return dyn, gp, ctl, mc


def run_b200_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    cpu_base = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        # before CUDA is initialised in this process (the workers are forked)
        cores = os.cpu_count() or 1
        n_scen = max(1, min(cores, 32))
        sps, wall, used = cpu_reference_run(n_scen, 1, 6, args.sqp_iters, args.qp_max_iter)
        cpu_base = {"value": sps, "unit": "solves/s", "cores": used, "kind": "port",
                    "sample": f"{n_scen} scenarios x 6 closed-loop control steps of the same workload, one process per "
                              f"scenario, {wall:.1f} s wall (numpy/scipy oracle of the reference algorithm)",
                    "gp_leg": CPU_GP_LEG, "controller_leg": CPU_CTL_LEG}
        if not args.no_extra and args.sqp_iters != 10:
            # CPU side of `extra.sqp10`: the reference's own GPMPC.solve (10 SQP iterations), bounded sample
            try:
                sps10, wall10, used10 = cpu_reference_run(n_scen, 1, 2, 10, args.qp_max_iter)
                cpu_base["sqp10"] = {"value": sps10, "unit": "solves/s", "cores": used10, "kind": "port",
                                     "sample": f"{n_scen} scenarios x 2 closed-loop control steps at 10 SQP iterations per "
                                               f"solve, {wall10:.1f} s wall", "gp_leg": CPU_GP_LEG,
                                     "controller_leg": CPU_CTL_LEG}
            except Exception as e:  # noqa: BLE001
                cpu_base["sqp10"] = {"error": repr(e)}
    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py --impl b200 needs a CUDA device (there is no CPU fallback)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from gp_mpc_rocket_landing_b200 import distributed as D
    from gp_mpc_rocket_landing_b200.mpc import engine as me
    dyn, gp, ctl, mc = build_problem(args, rank)
    B, K, W = args.batch, args.steps, max(args.warmup, 3)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident loop ---------------------------------------------------------------------------------
    # Every rank steps its own block of scenarios; the two collectives of the path run INSIDE the timed loop:
    #   * the per-step statistics all-reduce (8 counters) on a side stream, device tensors only (D.AsyncStatsReducer);
    #   * every `--theta-every` steps rank 0 broadcasts the GP hyper-parameters and every rank refits its replica
    #     (Gram + Cholesky + alpha + inverse factor) and re-reads it into the controller.
    peak64 = fp64_peak_tflops(torch) if rank == 0 else None
    clocks = ClockSampler(local)
    clocks.start()
    stats = D.AsyncStatsReducer(mc.counts)
    for _ in range(W):
        mc.step()
        stats.submit()
    stats.reset()
    timers = me.EventTimers(capacity=(K + 2) * (4 + 8 * args.sqp_iters))
    ctl.timers = timers
    n_theta, theta_events = 0, []
    barrier()
    mc.counts.zero_()
    clocks.mark()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    prof_range = bool(os.environ.get("GPMPC_PROFILE_RANGE"))    # `ncu --profile-from-start off`: the timed steps only
    if prof_range:
        torch.cuda.profiler.start()
    e0.record()
    launches = 0
    for k in range(K):
        mc.step()
        launches += ctl.last_kernel_launches + 1      # + mc_advance3_kernel (plant step, termination, respawn)
        stats.submit()
        # hyper-parameter broadcast + local refit every `--theta-every` steps; a run shorter than that does it once, half way
        if args.theta_every and ((k + 1) % args.theta_every == 0 or (K < args.theta_every and k + 1 == max(1, K // 2))):
            ta, tb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            ta.record()
            D.broadcast_hyperparameters(gp, version=2 + k // args.theta_every, refit=False)
            gp.fit()                                   # every rank refits (deterministic: replicas stay identical)
            ctl.refresh_gp()
            tb.record()
            theta_events.append((ta, tb))
            n_theta += 1
    e1.record()
    if prof_range:
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
    stats.wait()
    barrier()
    clk = clocks.stop()
    ms = e0.elapsed_time(e1)
    phases = timers.totals()
    step_ms = np.array(timers.step_times_ms())
    ctl.timers = None
    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    value = world * B * K / (ms * 1e-3)
    total = stats.total().cpu().numpy()               # summed over ranks by the in-loop all-reduces
    side_ms = stats.device_ms()
    theta_ms = sum(a.elapsed_time(b) for a, b in theta_events)
    # the same loop without any collective (no stats reduce, no broadcast / refit): what the collectives cost the step
    Kn = max(5, min(K, 40))
    barrier()
    n0, n1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n0.record()
    for _ in range(Kn):
        mc.step()
    n1.record()
    barrier()
    tn = torch.tensor([n0.elapsed_time(n1) / Kn], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tn, op=dist.ReduceOp.MAX)
    ms_nocomm = float(tn.item())

    # ---- end-to-end loop through the public numpy API ------------------------------------------------------------
    e2e = None
    if not args.no_e2e:
        from gp_mpc_rocket_landing_b200 import workloads
        X0, XT = workloads.sample_initial_conditions(rank * B, B)
        ctl.reset_warm_start()
        x = X0.copy()
        Ke = min(K, 40)
        per_step = mc.target_mode == "per_step"

        def e2e_step(x):
            s = ctl.solve_batch(x, XT, per_step_target=per_step)   # H2D: x (, XT); D2H: u0, cost, flags
            xn = dyn.step_batch(x, s.u0, DT)                        # H2D: x, u0 ; D2H: x_next
            # the termination rules of the Monte Carlo loop (experiments/monte_carlo.py:460-488), on the host copy:
            # finished scenarios restart from their initial condition, exactly as the device-resident loop respawns them
            done = (xn[:, 1] < 0.0) | (xn[:, 0] <= 1.01) | ((xn[:, 1] < 1.0) & (np.abs(xn[:, 4]) < 5.0)) | \
                   ~np.all(np.abs(xn) <= 1e6, axis=1)
            if done.any():
                xn[done] = X0[done]
                ctl.reset_scenarios(done)
            return xn

        for _ in range(W):
            x = e2e_step(x)
        barrier()
        t0 = time.perf_counter()
        for _ in range(Ke):
            x = e2e_step(x)
        barrier()
        te = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
        h2d = B * (7 + (0 if per_step else 7) + 7 + 3) * 8
        # u0 (3) + cost (1) doubles, iteration / status / flag arrays, and the next state from step_batch; the planned
        # trajectories X_opt / U_opt stay on the device until the caller reads them (BatchMPCSolution defers that copy)
        d2h = B * (3 + 1) * 8 + B * (4 + 4 + 4 + 4) + B * 7 * 8
        e2e = {"value": world * B * Ke / float(te.item()), "unit": "solves/s", "h2d_bytes_per_step": h2d,
               "d2h_bytes_per_step": d2h, "steps": Ke, "timer": "host wall clock around synchronised steps",
               "result_read": "u0, cost, status / iteration flags, next state (trajectories on request)"}

    extra = {}
    if rank == 0 and world == 1 and not args.no_extra:
        extra = run_extras(args, torch)

    if rank == 0:
        # ---- rooflines: algorithmic work per launch / measured launch time ------------------------
        share = {k: v[0] / ms for k, v in phases.items()}
        Q = B * N_HORIZON
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:  # noqa: BLE001
            pass
        hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
        hbm_src = "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback 6650 GB/s (B200_PROFILING.md)"
        n_pad = ((N_TRAIN + 127) // 128) * 128
        n_admm_it = float(total[7]) / max(1.0, float(total[0]))
        work = {  # phase -> (bound, algorithmic units per launch: flops or bytes, kernel description)
            "gp_variance": ("tensor", Q * float(N_TRAIN) ** 2,
                            "dgemm_nt_kernel x1: triangular DMMA GEMM |L^-1 K*^T|^2 against the explicit inverse factor, "
                            "sum of squares fused, V never stored (n^2 P_L flop per query with n = 1000, SURVEY 8d; the "
                            f"kernel runs on n_pad = {n_pad})"),
            "admm": ("hbm", B * 22e3, "admm kernel<7,3,20>: 22 kB compulsory per QP (SURVEY 8d); dependency-chain / "
                                      "shared-memory bound, neither HBM nor FLOP limited (DESIGN.md section 4, 7)"),
            "rollout": ("hbm", Q * 8.0 * n_pad + B * 8.0 * (7 + 60 + 147 + 60),
                        "rollout3_smem_kernel: K* rows written once (8 n_pad bytes per query) + trajectories"),
            "cov": ("hbm", B * 8.0 * (21 * 49 + 60 + 60 + 147 + 60) + Q * 8.0 * (n_pad // 64),
                    "var_cov3_kernel: partial sums -> variances -> covariance chain"),
            "assemble": ("hbm", B * 8.0 * (1400 + 147 + 207 + 207 + 147 + 60 + 60), "qp_assemble3_kernel (+ fused z pack)"),
            "update": ("hbm", B * 8.0 * (207 + 2 * 207 + 2 + 207), "mpc_update3_kernel (+ fused warm-start store)"),
            "prologue": ("hbm", B * 8.0 * (147 + 60 + 60 + 7), "solve_prologue3_kernel"),
        }
        by_kernel = {}
        for name, (tot, cnt) in phases.items():
            if name not in work or cnt == 0:
                continue
            bound, units, desc = work[name]
            avg_s = tot / cnt * 1e-3
            if bound == "tensor":
                r = {"kernel": desc, "bound": "tensor", "achieved": units / avg_s / 1e12, "peak": peak64,
                     "unit": "TFLOP/s", "peak_source": "cuBLAS DGEMM 8192^3 FP64 measured in this run (MEASURED_PEAKS.json "
                                                       "has no FP64 entry; FP64 has no tcgen05 kind: DMMA is the tensor path)"}
            else:
                r = {"kernel": desc, "bound": "hbm", "achieved": units / avg_s / 1e9, "peak": hbm_peak, "unit": "GB/s",
                     "peak_source": hbm_src}
            r["frac"] = r["achieved"] / r["peak"] if r["peak"] else None
            r["avg_ms_per_launch"] = avg_s * 1e3
            r["share_of_step"] = share[name]
            r["traffic"] = NCU_TRAFFIC.get(name)
            by_kernel[name] = r
        if "admm" in by_kernel:
            # "report achieved HBM GB/s against the compulsory-byte figure AND FP64 FMA utilisation": the
            # kernel is bound by its dependency chains, so a flop view and a latency model are printed next to the bytes
            r = by_kernel["admm"]
            flops = B * (53e3 + n_admm_it * 25e3)           # factor + 25 kFLOP per iteration
            r["fp64_tflops"] = flops / (r["avg_ms_per_launch"] * 1e-3) / 1e12
            r["fp64_fma_frac_of_dgemm_peak"] = r["fp64_tflops"] / peak64 if peak64 else None
            r["mean_iterations"] = n_admm_it
            m = dict(ADMM_MODEL)
            m["model_ms_per_launch"] = (m["setup_ms"] + n_admm_it * m["us_per_iteration_of_batch"] * 1e-3 * B / 4096.0) \
                * m["tail_factor"]
            m["cycles_per_qp_iteration"] = m["us_per_iteration_of_batch"] * 1e-6 * 1.965e9 * m["resident_qps"] / 4096.0
            m["chain_share_of_iteration"] = (m["dependent_chain_steps_per_iteration"] * m["cycles_per_chain_step"]
                                             / m["cycles_per_qp_iteration"])
            r["latency_model"] = m
        dom = max(by_kernel, key=lambda k: by_kernel[k]["share_of_step"])
        roof = dict(by_kernel[dom])
        roof["phase"] = dom
        outcomes = {n: int(total[i]) for i, n in enumerate(
            ("", "success", "crash", "fuel_exhausted", "constraint_violation", "timeout", "divergence")) if n}
        line = {
            "metric": METRIC, "value": value, "unit": "solves/s", "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": config_dict(args, world), "e2e": e2e, "gpu_launches": launches,
            "launches_per_step": launches / K,
            "roofline": roof, "roofline_by_kernel": by_kernel, "cpu_baseline": cpu_base, "clocks": clk,
            "phases_share_of_step": share, "phase_ms_per_step": {k: v[0] / K for k, v in phases.items()},
            "phases_sum_over_step": sum(v[0] for v in phases.values()) / ms,
            "mean_admm_iters_per_solve": n_admm_it,
            "p50_step_latency_ms": float(np.median(step_ms)) if step_ms.size else ms / K,
            "p99_step_latency_ms": float(np.percentile(step_ms, 99)) if step_ms.size else None,
            "max_step_latency_ms": float(step_ms.max()) if step_ms.size else None,
            "comm": {"stats_allreduce_per_step": 1, "theta_broadcast_every": args.theta_every,
                     "theta_broadcasts_and_refits_in_timed_region": n_theta,
                     "theta_broadcast_refit_ms_total_on_main_stream": theta_ms,
                     "ms_per_step_same_loop_without_collectives": ms_nocomm,
                     "comm_ms_per_step": ms / K - ms_nocomm,
                     "comm_within_run_to_run_noise": bool(abs(ms / K - ms_nocomm) < 0.03 * ms / K),   # the two loops see
                     # different scenario mixes (the re-run continues the closed loop), so a few % either way is noise
                     "stats_allreduce_side_stream_elapsed_ms_per_step": side_ms / K,
                     "note": "collectives are issued inside the timed loop; the stats all-reduce runs on a side stream (its "
                             "elapsed time includes waiting for a free SM next to the step's kernels and is off the critical "
                             "path); comm_ms_per_step = timed loop minus the same loop re-run without collectives"},
            "fp64_peak_tflops_measured": peak64,
            "outcomes_all_ranks": outcomes, "runs_finished_all_ranks": int(sum(outcomes.values())),
            "success_rate": outcomes["success"] / max(1, sum(outcomes.values())),
            "extra": extra,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


# latency model of the ADMM kernel: filled from the kernel's structure, printed with the roofline
# Latency model of the 3-DoF ADMM launch: fixed-iteration runs of the same kernel on the bench QPs
# (tools/cmp_admm.py, profiles/r2_admm_gen3_vs_gen2_3dof.json) give the set-up and per-iteration cost of a 4096-QP launch;
# the chain figures are the micro-benchmarks of tools/ubench (LDS -> 5 dependent DFMA -> STS -> warp barrier).
ADMM_MODEL = {"resident_qps": 444, "setup_ms": 0.64, "us_per_iteration_of_batch": 39.5, "tail_factor": 1.12,
              "dependent_chain_steps_per_iteration": 21, "cycles_per_chain_step": 190,
              "note": "launch time = (set-up + mean iterations x per-iteration cost) x tail; adaptive-rho refactorisations "
                      "are inside the tail factor; the kernel is bound by these dependent steps, not by HBM or FP64 rate"}


def run_extras(args, torch):
    """Short driver-run lines for the other BASELINE configs (rank 0, one GPU): S = 10 SQP iterations (gp_mpc.py:296,343),
    config 3 (refit sweep, n = 4000) and config 4 (6-DoF, 1024 scenarios).  Each is bounded to a few seconds."""
    out = {}
    try:
        import copy
        a10 = copy.copy(args)
        a10.sqp_iters, a10.batch = 10, args.batch
        dyn, gp, ctl, mc = build_problem(a10, 0)
        for _ in range(3):
            mc.step()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n = 10
        mc.counts.zero_()
        e0.record()
        for _ in range(n):
            mc.step()
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / n
        c = mc.counts.cpu().numpy()
        out["sqp10"] = {"config": "same workload with the reference's SQP loop, up to 10 iterations per solve "
                                  "(gp_mpc.py:296,343), uncertainty re-propagated after every iteration (:347-353)",
                        "ms_per_step": ms, "solves_per_s": args.batch / (ms * 1e-3),
                        "qp_solves_per_s": args.batch * 10 / (ms * 1e-3),
                        "mean_admm_iters_per_solve": float(c[7]) / max(1.0, float(c[0])), "launches_per_step": ctl.last_kernel_launches + 1}
        del dyn, gp, ctl, mc
        torch.cuda.empty_cache()
    except Exception as e:  # noqa: BLE001
        out["sqp10"] = {"error": repr(e)}
    try:
        import copy
        ac = copy.copy(args)
        ac.convex_rows = True
        dyn, gp, ctl, mc = build_problem(ac, 0)
        for _ in range(3):
            mc.step()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n = 10
        mc.counts.zero_()
        e0.record()
        for _ in range(n):
            mc.step()
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / n
        c = mc.counts.cpu().numpy()
        out["convex_rows"] = {"config": "same workload, QP with the polyhedral thrust-cone / glide-slope rows of gp_mpc.py:417-429 "
                                        "(augmented stage form n = 354, m = 648; DESIGN.md 11g)",
                              "ms_per_step": ms, "solves_per_s": args.batch / (ms * 1e-3),
                              "mean_admm_iters_per_solve": float(c[7]) / max(1.0, float(c[0])),
                              "launches_per_step": ctl.last_kernel_launches + 1}
        del dyn, gp, ctl, mc
        torch.cuda.empty_cache()
    except Exception as e:  # noqa: BLE001
        out["convex_rows"] = {"error": repr(e)}
    try:
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import bench_refit
        out["config3_refit_sweep_n4000"] = bench_refit.measure(4000, reps=2, kernels=("se_ard",))[0]
        torch.cuda.empty_cache()
    except Exception as e:  # noqa: BLE001
        out["config3_refit_sweep_n4000"] = {"error": repr(e)}
    try:
        import bench_6dof
        out["config4_6dof_1024"] = bench_6dof.measure(batch=1024, steps=10, warmup=3)
        torch.cuda.empty_cache()
    except Exception as e:  # noqa: BLE001
        out["config4_6dof_1024"] = {"error": repr(e)}
    return out


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        run_reference_arm(a)
    else:
        run_b200_arm(a)

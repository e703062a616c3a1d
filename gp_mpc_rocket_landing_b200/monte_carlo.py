"""Batched Monte Carlo driver: the reference's `MonteCarloSimulator.run/run_single` loop
(src/experiments/monte_carlo.py:401-636) turned inside-out -- all B scenarios advance one control step per
`controller.solve_batch` / `step_batch`, with per-scenario termination masks kept on the device.

Sharding: scenarios are independent, so rank r of W simulates the block [r*B/W, (r+1)*B/W) of the scenario
index range; initial conditions are generated per index from `seed + i` (monte_carlo.py:606-616), so the results do
not depend on W.  The only collectives are the stats gather and the hyper-parameter broadcast (parallel.py).
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Any, Dict, Optional

import numpy as np

from .gp import engine as ge
from .mpc import engine as me
from .workloads import SimulationConfig, sample_initial_conditions

OUTCOME_NAMES = ("running", "success", "crash", "fuel_exhausted", "constraint_violation", "timeout", "divergence")


@dataclass
class BatchMonteCarloResults:
    """Per-scenario terminal records (monte_carlo.py:133-183 `SimulationResult` fields that the batch keeps)."""

    outcome: np.ndarray          # (B,) LandingOutcome code
    final_state: np.ndarray      # (B,7)
    n_steps: np.ndarray          # (B,)
    fuel_used: np.ndarray        # (B,) m0 - m_final
    solve_fail_steps: np.ndarray  # (B,) control steps whose solve reported success == False
    counts: Dict[str, int] = field(default_factory=dict)

    @property
    def success_rate(self) -> float:
        done = self.outcome != 0
        return float(np.mean(self.outcome[done] == 1)) if done.any() else 0.0

    def get_statistics(self) -> Dict[str, Any]:
        """monte_carlo.py:186-250 subset."""
        done = self.outcome != 0
        ok = self.outcome == 1
        return {
            "n_runs": int(self.outcome.size), "n_finished": int(done.sum()), "success_rate": self.success_rate,
            "outcomes": dict(self.counts),
            "fuel_mean": float(self.fuel_used[ok].mean()) if ok.any() else float("nan"),
            "steps_mean": float(self.n_steps[done].mean()) if done.any() else float("nan"),
            "landing_error_mean": float(np.linalg.norm(self.final_state[ok][:, 2:4], axis=1).mean()) if ok.any() else float("nan"),
        }


class BatchedMonteCarlo:
    """Closed loop of B scenarios around a batched controller.

    controller: `GPMPC` / `NominalMPC3DoF` (uses `solve_batch`, target x' = x with v = 0 is NOT re-derived per step:
    the fixed target [m0, 0...] of monte_carlo.py:428-429 is used, as for controllers with `.step`) or
    `OSQPRTIMPC` / `FastRTI3DoF` (uses `initialize_batch` + `step_batch`).
    drag_cda: C_D*A of the "true" plant's drag residual (0 = the plant is the nominal model)."""

    def __init__(self, dynamics, controller, sim_config: Optional[SimulationConfig] = None, drag_cda: float = 0.0,
                 respawn: bool = False):
        self.dynamics = dynamics
        self.controller = controller
        self.cfg = sim_config or SimulationConfig()
        self.drag_cda = float(drag_cda)
        self.respawn = respawn
        self.max_steps = int(round(self.cfg.max_time / self.cfg.dt))
        self._rti = hasattr(controller, "step_batch")

    def reset(self, first: int, count: int, seed: int = 42):
        torch = ge.torch_()
        X0, XT = sample_initial_conditions(first, count, seed, self.cfg)
        p = self.dynamics.params
        g_vec = getattr(p, "g_vec", None)
        if g_vec is None:
            g_vec = p.g_I
        self._dyn = me.make_dyn(p.alpha, np.asarray(g_vec, dtype=np.float64), self.cfg.dt)
        self.B = count
        self.x0 = ge.to_dev(X0)
        self.xt = ge.to_dev(XT)
        self.x = self.x0.clone()
        self.m0 = self.x0[:, 0].contiguous()
        i32 = dict(dtype=torch.int32, device="cuda")
        self.t_step = torch.zeros((count,), **i32)
        self.active = torch.ones((count,), **i32)
        self.outcome = torch.zeros((count,), **i32)
        self.fail_steps = torch.zeros((count,), **i32)
        self.x_final = torch.zeros((count, 7), dtype=torch.float64, device="cuda")
        self.counts = torch.zeros((8,), dtype=torch.int64, device="cuda")
        self.total_solves = 0
        if self._rti:
            self.controller.initialize_batch(self.x0, self.xt)
        elif hasattr(self.controller, "reset_warm_start"):
            self.controller.reset_warm_start()
        return self

    def step(self):
        """One control step for every active scenario; returns the controller's batch solution."""
        torch = ge.torch_()
        if self._rti:
            sol = self.controller.step_batch(self.x, active=self.active)
        else:
            sol = self.controller.solve_batch(self.x, self.xt, active=self.active)
        self.fail_steps += (~sol.success).to(torch.int32) * self.active
        u0 = sol.u0.contiguous()
        ge.call("gpmpc_mc_step_3dof", me.C.byref(self._dyn), self.B, self.drag_cda, self.max_steps, ge.ptr(self.x),
                ge.ptr(u0), ge.ptr(self.m0), ge.ptr(self.t_step), ge.ptr(self.active), ge.ptr(self.outcome),
                ge.ptr(self.x_final), ge.ptr(self.counts), ge.stream())
        if self.respawn:
            done = self.active == 0
            self._respawn(done)
        return sol

    def _respawn(self, done):
        """Restart finished scenarios from their initial condition (keeps the batch full for throughput runs)."""
        if hasattr(self.controller, "reset_scenarios"):
            self.controller.reset_scenarios(done, self.x0)
        self.x = ge.torch_().where(done[:, None], self.x0, self.x).contiguous()
        self.t_step.masked_fill_(done, 0)
        self.active.fill_(1)

    def n_active(self) -> int:
        return int(self.active.sum().item())

    def run(self, first: int, count: int, seed: int = 42, max_steps: Optional[int] = None, progress=None):
        """Simulate scenarios [first, first+count) to termination -> BatchMonteCarloResults."""
        self.reset(first, count, seed)
        limit = self.max_steps if max_steps is None else max_steps
        for k in range(limit):
            self.step()
            if (k % 10 == 9 or k == limit - 1) and self.n_active() == 0:
                break
            if progress is not None:
                progress(k)
        return self.results()

    def results(self) -> BatchMonteCarloResults:
        outcome = self.outcome.cpu().numpy()
        xf = self.x_final.cpu().numpy()
        running = outcome == 0
        xf[running] = self.x.cpu().numpy()[running]
        m0 = self.m0.cpu().numpy()
        cnt = self.counts.cpu().numpy()
        return BatchMonteCarloResults(
            outcome=outcome, final_state=xf, n_steps=self.t_step.cpu().numpy(), fuel_used=m0 - xf[:, 0],
            solve_fail_steps=self.fail_steps.cpu().numpy(),
            counts={OUTCOME_NAMES[c]: int(cnt[c]) for c in range(1, 7)})

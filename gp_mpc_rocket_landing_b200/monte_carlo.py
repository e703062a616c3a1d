"""Batched Monte Carlo driver: the reference's `MonteCarloSimulator.run/run_single` loop
(src/experiments/monte_carlo.py:401-636) turned inside-out -- all B scenarios advance one control step per
`controller.solve_batch` / `step_batch`, with per-scenario termination masks kept on the device.

Loop semantics (all evaluated on the device, `gpmpc_mc_advance_3dof`):

* `target_mode="per_step"` -- what the reference does for a controller that has `.solve` and no `.step` (GPMPC,
  NominalMPC3DoF; monte_carlo.py:495-499): every step the target is re-derived from the current state,
  x_target = x with v = 0 and altitude = max(0.5, altitude - 2).  `"fixed"` keeps the target [m0, 0, ...] of :428-429
  (what a controller with `.step` sees through `initialize`).  Default: "per_step" for `.solve` controllers.
* `on_solve_failure="divergence"` -- the reference's branch :507-510: a solve with `success == False` ends the run as
  DIVERGENCE.  `"continue"` only counts the flag (what the reference does for `.step` controllers, :492-494).  Default:
  "divergence" for `.solve` controllers, "continue" for RTI controllers.  NB: `GPMPC.success` is the SQP step-size test
  (gp_mpc.py:337-345, 1e-4), which a real-time iteration (max_sqp_iter = 1) almost never meets, so throughput runs use
  "continue" -- DESIGN.md section 11 has the outcome tables for both.
* termination rules :460-488 on every state x_0 .. x_{max_steps-1}, TIMEOUT by the for/else (:539-542).

Sharding: scenarios are independent, so rank r of W simulates the block [r*B/W, (r+1)*B/W) of the scenario
index range; initial conditions are generated per index from `seed + i` (monte_carlo.py:606-616), so the results do
not depend on W.  The only collectives are the stats gather and the hyper-parameter broadcast (distributed.py).
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Any, Dict, Optional

import numpy as np

from . import _lib
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

    controller: `GPMPC` / `NominalMPC3DoF` (uses `solve_batch`) or `OSQPRTIMPC` / `FastRTI3DoF` (uses `initialize_batch` +
    `step_batch`).  target_mode / on_solve_failure: see the module docstring (None = the reference's behaviour for the
    controller's kind).  drag_cda: C_D*A of the "true" plant's drag residual (0 = the plant is the nominal model).
    safety_filter: a `safety.PredictiveSafetyFilter` applied to the first control before the plant step.
    respawn: a finished scenario restarts from its initial condition with a fresh controller state, which keeps the
    batch full (throughput runs); the outcome counters keep accumulating."""

    def __init__(self, dynamics, controller, sim_config: Optional[SimulationConfig] = None, drag_cda: float = 0.0,
                 respawn: bool = False, target_mode: Optional[str] = None, on_solve_failure: Optional[str] = None,
                 safety_filter=None):
        self.dynamics = dynamics
        self.controller = controller
        self.cfg = sim_config or SimulationConfig()
        self.drag_cda = float(drag_cda)
        self.respawn = respawn
        self.safety_filter = safety_filter     # monte_carlo.py:519-525: u <- safety_filter.filter(x, u).u_safe
        self.last_filter = None
        self.max_steps = int(self.cfg.max_time / self.cfg.dt)    # monte_carlo.py:447
        self._rti = hasattr(controller, "step_batch")
        self.target_mode = target_mode or ("fixed" if self._rti else "per_step")
        self.on_solve_failure = on_solve_failure or ("continue" if self._rti else "divergence")
        assert self.target_mode in ("fixed", "per_step") and self.on_solve_failure in ("divergence", "continue")
        if self._rti and self.target_mode == "per_step":
            raise ValueError("RTI controllers take their target in initialize_batch: target_mode must be 'fixed'")

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
        a = _lib.McStep3()
        a.B, a.N, a.dyn, a.drag_cda = count, int(getattr(self.controller.config, "N", 0)), self._dyn, self.drag_cda
        a.max_steps = self.max_steps
        a.fail_policy = 1 if self.on_solve_failure == "divergence" else 0
        a.respawn = 1 if self.respawn else 0
        for name in ("x", "m0", "t_step", "active", "outcome", "fail_steps", "x_final", "counts"):
            setattr(a, name, ge.ptr(getattr(self, name)))
        a.x_init = ge.ptr(self.x0)
        self._args = a
        # the reference applies the termination rules to x0 before the first control (monte_carlo.py:455-488)
        a.classify_only = 1
        ge.call("gpmpc_mc_advance_3dof", _lib.C.byref(a), ge.stream())
        a.classify_only = 0
        return self

    def step(self):
        """One control step for every active scenario; returns the controller's batch solution (device views)."""
        torch = ge.torch_()
        a = self._args
        if self._rti:
            sol = self.controller.step_batch(self.x, active=self.active)
            self._u0 = sol.u0.contiguous()
            self._ok = sol.success.to(torch.int32)
            a.u, a.ldu, a.success = ge.ptr(self._u0), self._u0.stride(0), ge.ptr(self._ok)
            a.qp_iters = None
            a.has_warm = a.y = a.rho = None
        else:
            sol = self.controller.solve_batch(self.x, self.xt, active=self.active,
                                              per_step_target=self.target_mode == "per_step")
            U = sol.U_opt                    # (B,N,3) view of the controller state: u0 = U[:,0]
            a.u, a.ldu, a.success = U.data_ptr(), U.stride(0), ge.ptr(sol.success)
            a.qp_iters = ge.ptr(sol.qp_iterations)
            st = getattr(self.controller, "_batch", None)
            if self.respawn and st is not None:   # controller state cleared on the device for respawned scenarios
                a.has_warm, a.y, a.rho, a.m = ge.ptr(st.has_warm), ge.ptr(st.duals), ge.ptr(st.ws.rho), st.duals.shape[1]
            else:
                a.has_warm = a.y = a.rho = None
        if self.safety_filter is not None:     # filter the first control on the device (safety/safety_filter.py:184-235)
            u_nom = sol.u0 if self._rti else sol.U_opt[:, 0]
            self.last_filter = self.safety_filter.filter_batch(self.x, u_nom, active=self.active)
            a.u, a.ldu = ge.ptr(self.last_filter.u_safe), 3
        ge.call("gpmpc_mc_advance_3dof", _lib.C.byref(a), ge.stream())
        return sol

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

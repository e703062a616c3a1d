"""GPU parity of the device Monte Carlo loop: `gpmpc_mc_advance_3dof` (plant step, termination rules,
check_landing, TIMEOUT, DIVERGENCE on a failed solve, respawn), `gpmpc_reset_scenarios`, the per-step target of the
fused solve prologue, and the composed closed loop on the MEASURED configuration of bench.py against the oracle.
Pinned by tests/golden/mc_loop.npz, written by the reference's own `MonteCarloSimulator.run_single`
(monte_carlo.py:401-583) around deterministic stub controllers."""
import ctypes as C

import numpy as np
import pytest

from conftest import load_golden

pytestmark = pytest.mark.gpu


def _args(B, dyn, x, m0, x0, t, act, out, fails, xf, counts, max_steps, policy, respawn=0, drag=0.0):
    from gp_mpc_rocket_landing_b200 import _lib
    from gp_mpc_rocket_landing_b200.gp import engine as ge
    a = _lib.McStep3()
    a.B, a.N, a.m, a.dyn, a.drag_cda = B, 20, 354, dyn, drag
    a.max_steps, a.fail_policy, a.respawn, a.classify_only = max_steps, policy, respawn, 0
    a.x, a.m0, a.x_init, a.t_step, a.active = ge.ptr(x), ge.ptr(m0), ge.ptr(x0), ge.ptr(t), ge.ptr(act)
    a.outcome, a.fail_steps, a.x_final, a.counts = ge.ptr(out), ge.ptr(fails), ge.ptr(xf), ge.ptr(counts)
    return a


def _state(X0):
    import torch
    from gp_mpc_rocket_landing_b200.gp import engine as ge
    B = X0.shape[0]
    i32 = dict(dtype=torch.int32, device="cuda")
    x0 = ge.to_dev(X0)
    return dict(x0=x0, x=x0.clone(), m0=x0[:, 0].contiguous(), t=torch.zeros(B, **i32), act=torch.ones(B, **i32),
                out=torch.zeros(B, **i32), fails=torch.zeros(B, **i32),
                xf=torch.zeros((B, 7), dtype=torch.float64, device="cuda"),
                counts=torch.zeros(8, dtype=torch.int64, device="cuda"))


@pytest.mark.parametrize("kind,target_mode,policy", [("solve", "per_step", 1), ("step", "fixed", 0)])
def test_mc_advance_reproduces_reference_run_single(kind, target_mode, policy):
    """32 seeds x up to 200 steps: the device loop (classification of x0, then advance after every control) with the
    controls of the same stub law the reference's loop was driven with -> outcome class, final state, step count, fuel,
    all equal to the reference's `run_single` (final state to the last bit: same Euler arithmetic, -fmad=false)."""
    from gp_mpc_rocket_landing_b200 import _lib
    from gp_mpc_rocket_landing_b200.gp import engine as ge
    from gp_mpc_rocket_landing_b200.mpc import engine as me
    from oracle import mpc as ompc
    g = load_golden("mc_loop.npz")
    X0 = g["x0"]
    B = X0.shape[0]
    dyn = me.make_dyn(ompc.Rocket3DoFParams().alpha, np.array([-1.0, 0.0, 0.0]), 0.1)
    s = _state(X0)
    a = _args(B, dyn, s["x"], s["m0"], s["x0"], s["t"], s["act"], s["out"], s["fails"], s["xf"], s["counts"], 200, policy)
    a.classify_only = 1
    ge.call("gpmpc_mc_advance_3dof", C.byref(a), ge.stream())
    a.classify_only = 0
    gains = [ompc.pd_test_gains(i) for i in range(B)]
    XT = np.zeros((B, 7)); XT[:, 0] = X0[:, 0]
    for k in range(200):
        xh = s["x"].cpu().numpy()
        act = s["act"].cpu().numpy()
        if not act.any():
            break
        u = np.zeros((B, 3)); ok = np.ones(B, dtype=np.int32)
        for b in range(B):
            if not act[b]:
                continue
            if target_mode == "per_step":
                xt = xh[b].copy(); xt[4:7] = 0; xt[1] = max(0.5, xh[b, 1] - 2.0)
            else:
                xt = XT[b]
            u[b] = ompc.pd_test_law(xh[b], xt, gains[b][0])
            ok[b] = 0 if k == gains[b][1] else 1
        ud, okd = ge.to_dev(u), ge.to_dev(ok, s["t"].dtype)
        a.u, a.ldu, a.success = ge.ptr(ud), 3, ge.ptr(okd)
        ge.call("gpmpc_mc_advance_3dof", C.byref(a), ge.stream())
    out = s["out"].cpu().numpy()
    np.testing.assert_array_equal(out, g[kind + "_outcome"])
    np.testing.assert_array_equal(s["t"].cpu().numpy(), g[kind + "_nsteps"])
    np.testing.assert_array_equal(s["xf"].cpu().numpy(), g[kind + "_final"])
    np.testing.assert_array_equal(X0[:, 0] - s["xf"].cpu().numpy()[:, 0], g[kind + "_fuel"])
    np.testing.assert_array_equal(s["counts"].cpu().numpy()[1:7], np.bincount(g[kind + "_outcome"], minlength=7)[1:])
    if kind == "solve":
        assert set(np.unique(out)) == {1, 2, 3, 4, 5, 6}
        assert int(s["fails"].sum().item()) == 4          # the four runs whose solve "failed" at step 10


def test_mc_advance_edge_states_timeout_and_respawn():
    """Termination rules and check_landing on hand-built edge states + the reference-generated landing fixture
    (monte_carlo.py:460-488, :54-104), TIMEOUT by the for/else without classifying the last state (:539-542), and
    respawn (fresh controller state: has_warm, duals, rho cleared; counters keep counting)."""
    import torch
    from gp_mpc_rocket_landing_b200.gp import engine as ge
    from gp_mpc_rocket_landing_b200.mpc import engine as me
    from oracle import mpc as ompc
    g = load_golden("monte_carlo.npz")
    ok = np.array([2.0 * 0.8, 0.5, 1.0, -1.0, -1.0, 0.2, -0.2])     # a successful touchdown state (m0 = 2)
    edge = [ok]
    for idx, val in ((1, -1e-9), (0, 1.01), (0, 1.0100001), (1, 0.999999), (1, 1.0), (2, 5.0000001), (3, -5.1), (4, -2.0000001),
                     (4, -4.999), (4, -5.0), (5, 1.01), (6, -1.01), (0, 0.09), (2, np.nan), (3, 1.1e6), (6, -np.inf),
                     (1, -0.5), (4, 2.5)):
        e = ok.copy(); e[idx] = val; edge.append(e)
    edge.append(np.array([0.05 * 2.0 + 1e-9, 0.5, 0, 0, -1, 0, 0]))      # fuel margin: mass fraction used > 0.95 -> m <= 1.01 first
    S = np.vstack([np.array(edge), g["states"]])
    B = S.shape[0]
    m0 = np.full(B, 2.0)
    want = np.array([ompc.termination(S[b], 2.0) for b in range(B)])
    wg = want[len(edge):]
    assert np.all(g["landing_ok"][wg == 1]) and not np.any(g["landing_ok"][wg == 4])   # the reference's check_landing
    assert set(np.unique(want)) >= {0, 1, 2, 3, 4, 6}
    dyn = me.make_dyn(ompc.Rocket3DoFParams().alpha, np.array([-1.0, 0.0, 0.0]), 0.1)
    s = _state(S)
    s["m0"] = ge.to_dev(m0)
    a = _args(B, dyn, s["x"], s["m0"], s["x0"], s["t"], s["act"], s["out"], s["fails"], s["xf"], s["counts"], 300, 1)
    a.classify_only = 1
    ge.call("gpmpc_mc_advance_3dof", C.byref(a), ge.stream())
    np.testing.assert_array_equal(s["out"].cpu().numpy(), want)
    np.testing.assert_array_equal(s["act"].cpu().numpy(), (want == 0).astype(np.int32))
    np.testing.assert_array_equal(s["x"].cpu().numpy(), S)          # classification does not move the state
    xf = s["xf"].cpu().numpy()
    fin = want != 0
    np.testing.assert_array_equal(xf[fin], S[fin])
    # --- plant step with drag + TIMEOUT: t_step = max_steps - 1 -> the next state is not classified
    X0 = np.tile(np.array([1.8, 1.2, 0.5, -0.5, -1.0, 0.1, -0.1]), (4, 1))       # one step above the landing band
    U = np.tile(np.array([1.7, 0.1, -0.1]), (4, 1))
    s = _state(X0)
    s["t"][:] = torch.tensor([0, 298, 299, 5], dtype=torch.int32)
    a = _args(4, dyn, s["x"], s["m0"], s["x0"], s["t"], s["act"], s["out"], s["fails"], s["xf"], s["counts"], 300, 1, drag=0.25)
    ud = ge.to_dev(U)
    okd = ge.to_dev(np.array([1, 1, 1, 0], dtype=np.int32), torch.int32)
    a.u, a.ldu, a.success = ge.ptr(ud), 3, ge.ptr(okd)
    ge.call("gpmpc_mc_advance_3dof", C.byref(a), ge.stream())
    p = ompc.Rocket3DoFParams()
    xn = ompc.step_3dof(X0[0], U[0], 0.1, p)
    rho = 1.225 * np.exp(-X0[0, 1] / 8500.0); sp = np.sqrt(X0[0, 4] ** 2 + X0[0, 5] ** 2 + X0[0, 6] ** 2)
    xn[4:7] = xn[4:7] + 0.1 * (-0.5 * rho * 0.25 * sp * X0[0, 4:7] / X0[0, 0])
    assert ompc.termination(xn, 1.8) == 0
    x = s["x"].cpu().numpy()
    np.testing.assert_array_equal(x[0], xn)
    np.testing.assert_array_equal(s["out"].cpu().numpy(), [0, 0, 5, 6])        # running, running (t=299), TIMEOUT (t=300), DIVERGENCE
    np.testing.assert_array_equal(s["t"].cpu().numpy(), [1, 299, 300, 5])     # the failed solve takes no plant step
    np.testing.assert_array_equal(x[3], X0[3])
    np.testing.assert_array_equal(s["fails"].cpu().numpy(), [0, 0, 0, 1])
    # --- respawn: crash -> restart from x_init with a fresh controller state
    X0 = np.tile(np.array([1.8, 0.05, 0.0, 0.0, -8.0, 0.0, 0.0]), (3, 1)); X0[1, 1] = 20.0
    s = _state(X0)
    xinit = ge.to_dev(np.tile(np.array([2.0, 30.0, 1.0, 2.0, -3.0, 0.0, 0.0]), (3, 1)))
    a = _args(3, dyn, s["x"], s["m0"], xinit, s["t"], s["act"], s["out"], s["fails"], s["xf"], s["counts"], 300, 0, respawn=1)
    hw = torch.ones(3, dtype=torch.int32, device="cuda")
    y = torch.ones((3, 354), dtype=torch.float64, device="cuda")
    rho_ = torch.full((3,), 0.7, dtype=torch.float64, device="cuda")
    a.has_warm, a.y, a.rho = ge.ptr(hw), ge.ptr(y), ge.ptr(rho_)
    ud = ge.to_dev(np.tile(np.array([0.3, 0.0, 0.0]), (3, 1)))
    a.u, a.ldu, a.success = ge.ptr(ud), 3, None
    ge.call("gpmpc_mc_advance_3dof", C.byref(a), ge.stream())
    np.testing.assert_array_equal(s["out"].cpu().numpy(), [2, 0, 2])
    np.testing.assert_array_equal(s["act"].cpu().numpy(), [1, 1, 1])
    np.testing.assert_array_equal(s["x"].cpu().numpy()[[0, 2]], xinit.cpu().numpy()[[0, 2]])
    np.testing.assert_array_equal(s["t"].cpu().numpy(), [0, 1, 0])
    np.testing.assert_array_equal(hw.cpu().numpy(), [0, 1, 0])
    np.testing.assert_array_equal(rho_.cpu().numpy(), [0.0, 0.7, 0.0])
    assert float(y[0].abs().sum()) == 0.0 and float(y[1].sum()) == 354.0 and float(y[2].abs().sum()) == 0.0
    assert int(s["counts"][2].item()) == 2


def test_reset_scenarios_and_per_step_target():
    """`GPMPC.reset_scenarios` (host and device masks) and the per-step target the solve prologue derives on the device
    (monte_carlo.py:497-499) against the targets the reference's loop handed to its controller (golden)."""
    import torch
    from gp_mpc_rocket_landing_b200.dynamics import Rocket3DoFDynamics
    from gp_mpc_rocket_landing_b200.mpc import GPMPC, GPMPCConfig
    g = load_golden("mc_loop.npz")
    X0 = g["x0"]
    B = X0.shape[0]
    ctl = GPMPC(Rocket3DoFDynamics(), None, GPMPCConfig(N=20, dt=0.1, max_sqp_iter=1, qp_max_iter=200))
    sol = ctl.solve_batch(X0, per_step_target=True)
    st = ctl._batch
    xr = st.x_ref.cpu().numpy()
    for k in (0, 7, 20):
        np.testing.assert_array_equal(xr[:, k], g["solve_targets"][:, 0])     # first target of every run = f(x0)
    assert sol.u0.shape == (B, 3) and np.all(st.has_warm.cpu().numpy() == 1)
    assert float(st.ws.y.abs().sum()) > 0 and np.all(st.ws.rho.cpu().numpy() > 0)
    mask = np.zeros(B, dtype=bool); mask[[0, 5, B - 1]] = True
    y_before = st.ws.y.clone()
    ctl.reset_scenarios(mask)
    hw = st.has_warm.cpu().numpy()
    np.testing.assert_array_equal(hw, (~mask).astype(np.int32))
    assert float(st.ws.y[mask.nonzero()[0]].abs().sum()) == 0.0
    assert torch.equal(st.ws.y[(~mask).nonzero()[0]], y_before[(~mask).nonzero()[0]])
    np.testing.assert_array_equal(st.ws.rho.cpu().numpy()[mask], 0.0)
    assert np.all(st.ws.rho.cpu().numpy()[~mask] > 0)
    ctl.reset_scenarios(torch.ones(B, dtype=torch.int32, device="cuda"))
    assert int(st.has_warm.sum().item()) == 0 and float(st.ws.y.abs().sum()) == 0.0
    # a reset scenario solves exactly like a fresh controller
    fresh = GPMPC(Rocket3DoFDynamics(), None, GPMPCConfig(N=20, dt=0.1, max_sqp_iter=1, qp_max_iter=200))
    s1 = fresh.solve_batch(X0, per_step_target=True)
    s2 = ctl.solve_batch(X0, per_step_target=True)
    np.testing.assert_array_equal(s1.U_opt, s2.U_opt)
    np.testing.assert_array_equal(s1.qp_iterations, s2.qp_iterations)


def test_closed_loop_on_the_measured_configuration_matches_oracle():
    """bench.py's exact problem -- 1000-point exact GP at the default noise 1e-4 (cond 9e6), B = 4096 scenarios, N = 20,
    S = 1, explicit-inverse variance path, drag plant, per-step targets, respawn on -- stepped 3 times through
    `BatchedMonteCarlo.step`, with 18 scenarios spread over the batch (first / last, neighbours of the 128-thread CTA
    boundaries and of the two-scenarios-per-warp / 147-CTA rollout tiling) compared with `GPMPCOracle` + the plant step:
    QP status, ADMM iteration count, u0 within the OSQP tolerance, trajectories, covariances, next state."""
    import bench
    from oracle import gp as ogp, mpc as ompc
    from gp_mpc_rocket_landing_b200 import workloads

    class A:
        batch, sqp_iters, qp_max_iter = 4096, 1, 1000
    dyn, gp, ctl, mc = bench.build_problem(A, 0)
    assert ctl.config.variance_via_inverse and mc.respawn and mc.target_mode == "per_step"
    idx = [0, 1, 2, 3, 127, 128, 255, 256, 1023, 1024, 2047, 2048, 2049, 2939, 2940, 4093, 4094, 4095]
    X0, _ = workloads.sample_initial_conditions(0, 4096)
    X, U, D = workloads.residual_dataset(bench.N_TRAIN, seed=0)
    o = ogp.Simple3DoFGPOracle(noise_variance=bench.NOISE)
    o.add_data(X, U, D); o.fit()
    ors = {b: ompc.GPMPCOracle(o.predict, N=20, dt=0.1, settings=ompc.OSQPSettings(max_iter=1000), max_sqp_iter=1)
           for b in idx}
    p = ompc.Rocket3DoFParams()
    x = {b: X0[b].copy() for b in idx}
    for step in range(3):
        xd = mc.x.cpu().numpy()
        sol = mc.step()
        st = ctl._batch
        u0 = sol.u0.cpu().numpy(); Xo = sol.X_opt.cpu().numpy(); Uo = sol.U_opt.cpu().numpy()
        qs = sol.qp_status.cpu().numpy(); qi = sol.qp_iterations.cpu().numpy(); covs = sol.covariances.cpu().numpy()
        ok = sol.success.cpu().numpy(); cost = sol.cost.cpu().numpy()
        xn = mc.x.cpu().numpy()
        for b in idx:
            np.testing.assert_allclose(xd[b], x[b], rtol=0, atol=5e-6 * (step + 1))   # both loops follow the same states
            xt = x[b].copy(); xt[4:7] = 0; xt[1] = max(0.5, x[b][1] - 2.0)
            r = ors[b].solve(xd[b], xt)       # the oracle solves from the DEVICE state: per-step parity, no drift
            assert int(qs[b]) == r.qp_status[-1] == 1, (step, b)
            assert int(qi[b]) == sum(r.qp_iters), (step, b, qi[b], r.qp_iters)
            assert bool(ok[b]) == r.success
            assert np.max(np.abs(u0[b] - r.U_opt[0])) < 1e-4 + 1e-4 * np.max(np.abs(r.U_opt[0]))   # OSQP eps_abs / eps_rel
            np.testing.assert_allclose(Uo[b], r.U_opt, rtol=0, atol=2e-5)
            np.testing.assert_allclose(Xo[b], r.X_opt, rtol=0, atol=2e-5)
            assert cost[b] == pytest.approx(r.cost, rel=1e-5)
            np.testing.assert_allclose(covs[b], r.covariances, rtol=1e-5, atol=1e-16)
            xo = ompc.step_3dof(xd[b], u0[b], 0.1, p)
            rho = 1.225 * np.exp(-xd[b][1] / 8500.0); sp = np.sqrt(xd[b][4] ** 2 + xd[b][5] ** 2 + xd[b][6] ** 2)
            xo[4:7] = xo[4:7] + 0.1 * (-0.5 * rho * 0.25 * sp * xd[b][4:7] / xd[b][0])
            np.testing.assert_array_equal(xn[b], xo)            # plant step + drag: bit-equal given the same u0
            x[b] = ompc.step_3dof(x[b], r.U_opt[0], 0.1, p)
            x[b][4:7] = x[b][4:7] + 0.1 * (-0.5 * rho * 0.25 * sp * xd[b][4:7] / xd[b][0])
    cnt = mc.counts.cpu().numpy()
    assert int(cnt[1:7].sum()) == 0 and int(mc.active.sum().item()) == 4096     # nobody lands within 3 steps
    assert int(cnt[0]) == 3 * 4096 and int(cnt[7]) > 3 * 4096 * 25                # solves consumed, ADMM iterations
    assert ctl.last_kernel_launches == 7     # prologue, rollout, variance GEMM, var+cov, assemble, ADMM, update


@pytest.mark.parametrize("mode", ["per_step", "fixed"])
def test_closed_loop_outcomes_match_the_cpu_oracle_on_64_seeds(mode):
    """the device loop run to termination on the first 64 scenarios of the measured workload (bench.py's
    GP and controller, drag plant, no respawn) against tests/golden/mc_outcomes_64.npz -- the CPU oracle's
    `run_single` + `GPMPCOracle` on the same seeds (oracle/gen_outcomes.py) -- for both target modes of monte_carlo.py:492-510:
    the same outcome class and the same termination step for every scenario, hence the same histogram, final states to
    5e-3 (observed 1.7e-3: a hundred closed-loop steps of two solvers that agree to the ADMM tolerance 1e-4 per step)."""
    import bench
    from gp_mpc_rocket_landing_b200.monte_carlo import BatchedMonteCarlo
    g = load_golden("mc_outcomes_64.npz")

    class A:
        batch, sqp_iters, qp_max_iter = 64, 1, 1000
    dyn, gp, ctl, _ = bench.build_problem(A, 0)
    mc = BatchedMonteCarlo(dyn, ctl, drag_cda=0.25, respawn=False, target_mode=mode, on_solve_failure="continue")
    res = mc.run(first=0, count=64, seed=42)
    want = g[mode + "_outcome"]
    assert np.bincount(res.outcome, minlength=7).tolist() == np.bincount(want, minlength=7).tolist()
    np.testing.assert_array_equal(res.outcome, want)
    np.testing.assert_array_equal(res.n_steps, g[mode + "_steps"])
    np.testing.assert_allclose(res.final_state, g[mode + "_x_final"], rtol=0, atol=5e-3)
    np.testing.assert_array_equal(res.solve_fail_steps, g[mode + "_failed_solves"])
    # ADMM work of the whole run: the same iteration counts summed over every solve of every scenario (within 2 %: a QP
    # whose residual sits on the tolerance at a check may stop one check -- 25 iterations -- earlier or later)
    it = int(mc.counts.cpu().numpy()[7])
    assert abs(it - int(g[mode + "_admm_iters"].sum())) <= 0.02 * it


def test_reset_scenarios_with_a_host_mask_equals_the_device_mask():
    """`GPMPC.reset_scenarios` with a numpy mask (the path bench.py's end-to-end loop takes: one pinned H2D copy, one launch)
    against the same reset with a device mask: warm-start flags, duals and rho of the masked scenarios are cleared, the
    others untouched, and the next solve is identical either way."""
    import torch
    import bench
    from gp_mpc_rocket_landing_b200 import workloads

    class A:
        batch, sqp_iters, qp_max_iter = 96, 1, 1000
    X0, _ = workloads.sample_initial_conditions(0, 96)
    mask = np.zeros(96, dtype=bool); mask[[0, 5, 31, 32, 95]] = True
    outs = []
    for host in (True, False):
        dyn, gp, ctl, _ = bench.build_problem(A, 0)
        x = torch.as_tensor(X0, device="cuda")
        ctl.solve_batch(x, per_step_target=True)
        st = ctl._batch
        assert int(st.has_warm.sum().item()) == 96 and float(st.ws.rho.min().item()) > 0.0
        ctl.reset_scenarios(mask if host else torch.as_tensor(mask, device="cuda"))
        hw = st.has_warm.cpu().numpy(); rho = st.ws.rho.cpu().numpy(); y = st.ws.y.cpu().numpy()
        assert np.array_equal(hw == 0, mask) and np.array_equal(rho == 0.0, mask)
        assert not y[mask].any() and np.abs(y[~mask]).max() > 0
        sol = ctl.solve_batch(x, per_step_target=True)
        outs.append((sol.U_opt.cpu().numpy().copy(), sol.qp_iterations.cpu().numpy().copy()))
    np.testing.assert_array_equal(outs[0][0], outs[1][0])
    np.testing.assert_array_equal(outs[0][1], outs[1][1])

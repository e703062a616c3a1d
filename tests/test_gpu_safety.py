"""GPU parity of the predictive safety filter: `gpmpc_safety_filter_3dof` through
`safety.PredictiveSafetyFilter.filter / filter_batch` against the reference's own filter (golden fixture written by
src/safety/safety_filter.py's classes on a 14-state embedding) and against the numpy restatement."""
import numpy as np
import pytest

from conftest import load_golden

pytestmark = pytest.mark.gpu


class CP:
    T_min, T_max, theta_max, gamma_gs = 0.3, 5.0, 60.0, 30.0


@pytest.mark.parametrize("tag", ["a25", "a100"])
def test_safety_filter_matches_reference(tag):
    import torch
    from gp_mpc_rocket_landing_b200.dynamics import Rocket3DoFDynamics
    from gp_mpc_rocket_landing_b200.safety import PredictiveSafetyFilter, SafetyFilterConfig
    from oracle import safety as osf
    g = load_golden("safety_filter.npz")
    sf = PredictiveSafetyFilter(Rocket3DoFDynamics(), SafetyFilterConfig(N=10, dt=0.1), CP())
    with pytest.raises(RuntimeError):
        sf.filter(g[f"{tag}_X"][0], g[f"{tag}_U"][0])                  # safety_filter.py:200-201
    sf.initialize(g[f"{tag}_x_eq"], g[f"{tag}_u_eq"], backup_alpha=float(g[f"{tag}_alpha"]), K=g[f"{tag}_K"], P=g[f"{tag}_P"])
    X, U = g[f"{tag}_X"], g[f"{tag}_U"]
    r = sf.filter_batch(X, U)
    np.testing.assert_array_equal(r.is_modified, g[f"{tag}_modified"])
    np.testing.assert_array_equal(r.is_safe, g[f"{tag}_safe"])
    # 50 gradient steps with forward differences (eps 1e-4) amplify the last-bit differences of the backup law's dot
    # products (BLAS order in the reference) by ~1e4 per step: 1e-8 relative on the filtered control
    np.testing.assert_allclose(r.u_safe, g[f"{tag}_u_safe"], rtol=1e-8, atol=1e-10)
    np.testing.assert_allclose(r.safety_margin, g[f"{tag}_margin"], rtol=1e-7, atol=1e-8)
    np.testing.assert_allclose(r.backup_value, g[f"{tag}_backup_value"], rtol=1e-7, atol=1e-8)
    keep = ~g[f"{tag}_modified"]
    np.testing.assert_array_equal(r.u_safe[keep], U[keep])                 # safe controls pass through untouched
    # restatement, case by case
    s = osf.SafetySetup(K=g[f"{tag}_K"], P=g[f"{tag}_P"], x_eq=g[f"{tag}_x_eq"], u_eq=g[f"{tag}_u_eq"], alpha=float(g[f"{tag}_alpha"]))
    for i in (0, 7, 25, 32, 47):
        u, m, ok, margin, bv = osf.safety_filter(s, X[i], U[i])
        np.testing.assert_allclose(r.u_safe[i], u, rtol=1e-8, atol=1e-10)
        assert bool(r.is_modified[i]) == m and bool(r.is_safe[i]) == ok
    # single-state API, device tensors, a strided view of a plan as the nominal control, the active mask
    one = sf.filter(X[2], U[2])
    assert one.is_modified == bool(g[f"{tag}_modified"][2]) and one.u_safe.shape == (3,)
    np.testing.assert_array_equal(one.u_safe, r.u_safe[2])
    plan = torch.zeros((X.shape[0], 20, 3), dtype=torch.float64, device="cuda")
    plan[:, 0] = torch.from_numpy(U).cuda()
    act = torch.ones(X.shape[0], dtype=torch.int32, device="cuda"); act[5] = 0
    rd = sf.filter_batch(torch.from_numpy(X).cuda(), plan[:, 0], active=act)
    got = rd.u_safe.cpu().numpy()
    m = np.ones(X.shape[0], bool); m[5] = False
    np.testing.assert_array_equal(got[m], r.u_safe[m])
    # (reference quirk, reproduced: when the terminal value already satisfies the margin the gradient loop stops before its
    #  first projection, so a nominal control that only violates the thrust annulus comes back unchanged and "unsafe")
    T = np.linalg.norm(r.u_safe, axis=1)
    outside = (T < 0.3) | (T > 5.0 + 1e-12)
    assert np.array_equal(r.u_safe[outside], U[outside]) and not r.is_safe[outside].any()


def test_backup_controller_and_monte_carlo_hook():
    """LQRBackupController on the 7-state model (DARE or the reference's PD fall-back) and the filter inside the batched
    Monte Carlo loop (monte_carlo.py:519-525): the applied control is the filtered one."""
    import torch
    from gp_mpc_rocket_landing_b200.dynamics import Rocket3DoFDynamics
    from gp_mpc_rocket_landing_b200.monte_carlo import BatchedMonteCarlo
    from gp_mpc_rocket_landing_b200.mpc import GPMPC, GPMPCConfig
    from gp_mpc_rocket_landing_b200.safety import LQRBackupController, PredictiveSafetyFilter, SafetyFilterConfig
    from oracle import mpc as ompc
    dyn = Rocket3DoFDynamics()
    bk = LQRBackupController(dyn)
    x_eq = np.array([1.8, 5.0, 0, 0, 0, 0, 0.0])
    bk.compute_gains(x_eq)
    assert bk.K.shape == (3, 7) and bk.P.shape == (7, 7) and np.allclose(bk.u_eq, [1.8, 0, 0])
    assert np.allclose(bk.get_control(x_eq), bk.u_eq)
    sf = PredictiveSafetyFilter(dyn, SafetyFilterConfig(N=10, dt=0.1), type("CP", (), dict(T_min=0.3, T_max=5.0, gamma_gs=30.0))())
    sf.initialize(x_eq, backup_alpha=400.0)
    ctl = GPMPC(dyn, None, GPMPCConfig(N=20, dt=0.1, max_sqp_iter=1, qp_max_iter=400))
    mc = BatchedMonteCarlo(dyn, ctl, on_solve_failure="continue", safety_filter=sf)
    mc.reset(0, 16)
    x_before = mc.x.cpu().numpy()
    sol = mc.step()
    u0 = sol.U_opt[:, 0].cpu().numpy()
    res = sf.filter_batch(x_before, u0)
    p = ompc.Rocket3DoFParams()
    x_after = mc.x.cpu().numpy()
    for b in range(16):
        np.testing.assert_array_equal(x_after[b], ompc.step_3dof(x_before[b], res.u_safe[b], 0.1, p))
    np.testing.assert_array_equal(mc.last_filter.u_safe.cpu().numpy(), res.u_safe)

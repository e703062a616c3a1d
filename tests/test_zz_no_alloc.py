"""GPU: the allocation-free entry point `gpmpc_admm_solve_noalloc` (static device scratch; no counterpart in the reference)
against the existing `gpmpc_admm_solve` on the same QPs, through the C-ABI,
on two calls in a row that share the static array (different batch sizes and horizons, so the second call lands on the first
call's leftovers), plus the size limit.  Collected last on purpose (file name)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _batch(nx, N, B, seed):
    from test_gpu_mpc import _random_stage_qp
    rng = np.random.default_rng(seed)
    probs = [_random_stage_qp(nx, 3, N, rng) for _ in range(B)]
    AB = np.stack([p[0] for p in probs]); beq = np.stack([p[1] for p in probs]); q = np.stack([p[3] for p in probs])
    return AB, beq, q, probs[0][2], probs[0][4], probs[0][5]      # Pdiag and the box are shared by the batch


def _solve(entry, N, B, data, variant):
    import torch
    from gp_mpc_rocket_landing_b200 import _lib
    from gp_mpc_rocket_landing_b200.gp import engine as ge
    from gp_mpc_rocket_landing_b200.mpc import engine as me
    AB, beq, q, Pd, lb, ub = (ge.to_dev(a) for a in data)
    ws = me.AdmmWorkspace(B, N)
    S = me.AdmmSettings.reference_defaults(max_iter=2000, kernel_variant=variant)
    S.bound_stride = 0
    act = torch.ones(B, dtype=torch.int32, device="cuda")
    act[B - 1] = 0                                             # an inactive scenario stays untouched
    out = []
    for _ in range(2):                                         # cold, then warm-started from its own x / y / rho
        head = (7, 3, N, B) if entry == "gpmpc_admm_solve" else (N, B)
        ge.call(entry, *head, ge.ptr(AB), ge.ptr(beq), ge.ptr(q), ge.ptr(Pd), ge.ptr(lb), ge.ptr(ub), _lib.C.byref(S),
                ge.ptr(act), ge.ptr(ws.x), ge.ptr(ws.y), ge.ptr(ws.rho), ge.ptr(ws.status), ge.ptr(ws.iters), ge.ptr(ws.obj),
                ge.ptr(ws.pri), ge.ptr(ws.dua), ge.stream())
        S.warm_x = S.warm_y = S.rho_from_input = 1
        torch.cuda.synchronize()
        out.append({k: getattr(ws, k).cpu().numpy().copy() for k in ("x", "y", "rho", "status", "iters", "obj")})
    return out


def test_noalloc_entry_point_matches_admm_solve_on_two_calls_in_a_row():
    # call 1: 6 QPs at N = 20 (the path's shape); call 2: 9 QPs at N = 11 -- another layout over the same static array
    for N, B, seed in ((20, 6, 11), (11, 9, 12)):
        data = _batch(7, N, B, seed)
        got = _solve("gpmpc_admm_solve_noalloc", N, B, data, 0)
        want = _solve("gpmpc_admm_solve", N, B, data, 2)       # the same stage kernel with its stream-ordered scratch
        for g, w in zip(got, want):
            np.testing.assert_array_equal(g["status"], w["status"])
            np.testing.assert_array_equal(g["iters"], w["iters"])
            # random QPs that share scenario 0's box: most are solved, a few may end primal infeasible (-3) -- on both sides
            assert (g["status"][:-1] == 1).sum() >= (B - 1) // 2 + 1 and g["status"][-1] == 0 and g["iters"][:-1].min() > 0
            for k in ("x", "y", "rho", "obj"):                 # 1e-6: the tolerance of test_admm_generic_stage_qp
                assert np.max(np.abs(g[k] - w[k])) <= 1e-6 * max(1.0, np.max(np.abs(w[k]))), k
            assert not g["x"][-1].any()


def test_noalloc_entry_point_rejects_sizes_beyond_the_static_array():
    from gp_mpc_rocket_landing_b200 import _lib
    from gp_mpc_rocket_landing_b200.gp import engine as ge
    from gp_mpc_rocket_landing_b200.mpc import engine as me
    S = me.AdmmSettings.reference_defaults()
    nul = [None] * 6
    with pytest.raises(_lib.GpmpcError, match="GPMPC_NOALLOC_MAX_DOUBLES"):      # 1025 x 561 doubles > 574464
        ge.call("gpmpc_admm_solve_noalloc", 20, 1025, *nul, _lib.C.byref(S), None, *([None] * 8), ge.stream())
    # exactly at the limit the size check passes (B = 0 work is not launched here: only the arithmetic of the limit is probed)
    assert 1024 * (2 * 207 + 147) == 574464

"""GPU: out-of-bounds canaries for the kernels added in round 2 (compute-sanitizer is not available on the GPU pool): every
output buffer is the interior of a larger allocation whose margins hold a sentinel; after the call the margins must be
untouched and the interior fully written where the contract says so."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
G = 4096          # sentinel doubles / ints on each side
SENT = -7.25e300


def _guarded(shape, dtype=None, fill=SENT):
    import torch
    n = int(np.prod(shape))
    dtype = dtype or torch.float64
    buf = torch.full((n + 2 * G,), fill if dtype == torch.float64 else -123456789, dtype=dtype, device="cuda")
    return buf, buf[G:G + n].view(*shape)


def _intact(buf, n):
    import torch
    v = SENT if buf.dtype == torch.float64 else -123456789
    return bool((buf[:G] == v).all().item()) and bool((buf[G + n:] == v).all().item())


def test_cone_augmentation_and_admm_14_20_stay_inside_their_buffers():
    import torch
    from gp_mpc_rocket_landing_b200 import _lib
    from gp_mpc_rocket_landing_b200.gp import engine as ge
    from gp_mpc_rocket_landing_b200.mpc import engine as me
    from oracle import mpc as ompc
    from test_cone_rows import _problem
    cone = ompc.ConeRows(T_max=3.0)
    B, N = 7, 20                      # odd batch: the last CTA of every launch is partially filled
    probs = [_problem(s, N) for s in range(B)]
    AB = np.stack([np.stack([pr[5][7 + k * 7:14 + k * 7, k * 10:(k + 1) * 10] for k in range(N)]) for pr in probs])
    beq = np.stack([pr[6][:7 * (N + 1)] for pr in probs]); q = np.stack([pr[4] for pr in probs])
    X = np.stack([pr[1] for pr in probs]); U = np.stack([pr[2] for pr in probs])
    n_c, m_c = ompc.cone_dims(N)
    bufs = {}
    for name, shape in (("ABc", (B, N, 14, 17)), ("beqc", (B, 14 * (N + 1))), ("qc", (B, n_c)), ("zc", (B, n_c)),
                        ("y", (B, m_c)), ("z", (B, 10 * N + 7)), ("rho", (B,)), ("obj", (B,)), ("pri", (B,)), ("dua", (B,))):
        bufs[name] = _guarded(shape)
    ibufs = {name: _guarded((B,), torch.int32) for name in ("status", "iters")}
    c = _lib.Cone3(1, cone.T_max, cone.facet_b, cone.half_tan)
    d = [ge.to_dev(a) for a in (AB, beq, q, X, U)]
    ge.call("gpmpc_qp_augment_cone_3dof", _lib.C.byref(c), B, N, *[ge.ptr(t) for t in d], ge.ptr(bufs["ABc"][1]),
            ge.ptr(bufs["beqc"][1]), ge.ptr(bufs["qc"][1]), ge.ptr(bufs["zc"][1]), None, ge.stream())
    for name in ("ABc", "beqc", "qc", "zc"):
        buf, view = bufs[name]
        assert _intact(buf, view.numel()), name
        assert not bool((view == SENT).any().item()), name          # every entry of the outputs is written
    Pd, lb, ub = (ge.to_dev(a) for a in ompc.cone_shared_data(N, cone))
    lb = torch.clamp(lb, min=-1e30); ub = torch.clamp(ub, max=1e30)
    S = me.AdmmSettings.reference_defaults(max_iter=2000, warm_x=1)
    S.bound_stride = 0
    bufs["y"][1].zero_(); bufs["rho"][1].zero_()
    for variant in (0, 2):            # third and second generation at the (14, 20) shape
        S.kernel_variant = variant
        x = bufs["zc"][1].clone()
        xb, xv = _guarded((B, n_c)); xv.copy_(x)
        ge.call("gpmpc_admm_solve", 14, 3, N, B, ge.ptr(bufs["ABc"][1]), ge.ptr(bufs["beqc"][1]), ge.ptr(bufs["qc"][1]),
                ge.ptr(Pd), ge.ptr(lb), ge.ptr(ub), _lib.C.byref(S), None, ge.ptr(xv), ge.ptr(bufs["y"][1]),
                ge.ptr(bufs["rho"][1]), ge.ptr(ibufs["status"][1]), ge.ptr(ibufs["iters"][1]), ge.ptr(bufs["obj"][1]),
                ge.ptr(bufs["pri"][1]), ge.ptr(bufs["dua"][1]), ge.stream())
        torch.cuda.synchronize()
        assert _intact(xb, xv.numel()), variant
        for name in ("y", "rho", "obj", "pri", "dua"):
            assert _intact(bufs[name][0], bufs[name][1].numel()), (variant, name)
        for name in ("status", "iters"):
            assert _intact(ibufs[name][0], B), (variant, name)
        assert (ibufs["status"][1] == 1).all() and torch.isfinite(xv).all()
    ge.call("gpmpc_qp_extract_cone_3dof", B, N, ge.ptr(xv), ge.ptr(bufs["z"][1]), None, ge.stream())
    assert _intact(bufs["z"][0], bufs["z"][1].numel()) and not bool((bufs["z"][1] == SENT).any().item())


def test_variance_gemm_with_strip_skipping_stays_inside_its_work_buffer():
    import torch
    from gp_mpc_rocket_landing_b200.gp import engine as ge
    n, n_pad, Q = 200, 256, 384
    g = torch.Generator(device="cuda").manual_seed(1)
    L = torch.tril(torch.rand((n_pad, n_pad), dtype=torch.float64, device="cuda", generator=g))
    L[n:, :] = 0; L[n:, n:] = torch.eye(n_pad - n, dtype=torch.float64, device="cuda")
    Ks = torch.rand((Q, n_pad), dtype=torch.float64, device="cuda", generator=g); Ks[:, n:] = 0
    buf, work = _guarded((n_pad // 64, Q))
    for nn in (0, n):
        work.fill_(SENT)
        ge.call("gpmpc_linv_sumsq_parts_n", ge.ptr(L), nn, n_pad, L.stride(0), ge.ptr(Ks), Q, Ks.stride(0), ge.ptr(work), ge.stream())
        torch.cuda.synchronize()
        assert _intact(buf, work.numel()) and not bool((work == SENT).any().item())
        want = ((Ks @ L.T) ** 2).sum(1)
        torch.testing.assert_close(work.sum(0), want, rtol=1e-12, atol=0)


def test_fused_step_is_independent_of_the_batch_size():
    """Ragged batches (100 and 257 scenarios: neither a multiple of the CTA sizes, the two-scenarios-per-warp rollout tiling
    or the 128-row padding of the K* rows): the first 100 scenarios evolve bit-identically in both, over 4 closed-loop steps
    with respawn bookkeeping on -- an index or padding bug in any of the eight kernels of the step would show here."""
    import bench
    out = []
    for B in (100, 257):
        class A:
            batch, sqp_iters, qp_max_iter = B, 1, 1000
        dyn, gp, ctl, mc = bench.build_problem(A, 0)
        xs, us, its = [], [], []
        for _ in range(4):
            sol = mc.step()
            xs.append(mc.x[:100].cpu().numpy().copy()); us.append(sol.U_opt[:100].cpu().numpy().copy())
            its.append(sol.qp_iterations[:100].cpu().numpy().copy())
        out.append((xs, us, its, mc.counts.cpu().numpy().copy()))
    for a, b in zip(out[0][:3], out[1][:3]):
        for p, q in zip(a, b):
            np.testing.assert_array_equal(p, q)
    assert int(out[0][3][0]) == 400 and int(out[1][3][0]) == 4 * 257

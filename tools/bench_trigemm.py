"""GPU measurement (not part of the product): the triangular variance GEMM (gpmpc_linv_sumsq_parts_n) alone, at the bench
shape (n = 1000, n_pad = 1024, 81 920 query rows), with and without the strip-wise DMMA skipping
(GPMPC_NO_TRI_SKIP=1 switches it off for the process)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from gp_mpc_rocket_landing_b200.gp import engine as ge

n, n_pad, Q = 1000, 1024, int(os.environ.get("Q", 81920))
g = torch.Generator(device="cuda").manual_seed(0)
L = torch.tril(torch.rand((n_pad, n_pad), dtype=torch.float64, device="cuda", generator=g))
L[n:, :] = 0; L[n:, n:] = torch.eye(n_pad - n, dtype=torch.float64, device="cuda")
Ks = torch.rand((Q, n_pad), dtype=torch.float64, device="cuda", generator=g); Ks[:, n:] = 0
work = torch.zeros((n_pad // 64) * Q, dtype=torch.float64, device="cuda")
out = {}
for name, nn in (("n_unknown", 0), ("n_1000", n), ("n_512_timing_only", 512), ("n_64_timing_only", 64)):
    work.zero_()
    for _ in range(3):
        ge.call("gpmpc_linv_sumsq_parts_n", ge.ptr(L), nn, n_pad, L.stride(0), ge.ptr(Ks), Q, Ks.stride(0), ge.ptr(work), ge.stream())
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        ge.call("gpmpc_linv_sumsq_parts_n", ge.ptr(L), nn, n_pad, L.stride(0), ge.ptr(Ks), Q, Ks.stride(0), ge.ptr(work), ge.stream())
    e1.record(); torch.cuda.synchronize()
    out[name + "_ms"] = e0.elapsed_time(e1) / 10
    out[name + "_checksum"] = float(work.view(n_pad // 64, Q).sum(0)[:8].sum().item())
ref = ((Ks[:8] @ L.T) ** 2).sum().item()
out["torch_reference_checksum"] = ref
out["skip"] = os.environ.get("GPMPC_NO_TRI_SKIP") is None
out["Q"] = Q
out["n_1000_tflops_executed_padded"] = 1024.0 * 1024 * Q / (out["n_1000_ms"] * 1e-3) / 1e12
print(json.dumps(out))

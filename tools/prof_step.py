"""GPU profiling target (not part of the product): a few closed-loop steps of the bench workload."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench

class A: pass
a = A(); a.batch = int(os.environ.get("B", 4096)); a.sqp_iters = int(os.environ.get("S", 1)); a.qp_max_iter = 1000
a.admm_variant = int(os.environ.get("ADMM_VARIANT", 0))
dyn, gp, ctl, mc = bench.build_problem(a, 0)
for k in range(int(os.environ.get("STEPS", 4))):
    mc.step()
torch.cuda.synchronize()
if os.environ.get("PROFILE_API"):  # ncu --profile-from-start off: capture exactly one more closed-loop step
    torch.cuda.profiler.start()
    mc.step()
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()
print("done")

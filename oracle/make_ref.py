"""TEST INFRASTRUCTURE ONLY.  Recipe for `oracle/_ref/`: the reference's own GP package, taken from where it lies under
/root/reference (read-only, present only in the build container) into `oracle/_ref/src/gp/` so that it travels to the GPU
box with the snapshot (`oracle/_ref/` is git-ignored, not gpurun-ignored: nothing of the reference enters the history).
The reference is pure Python, so "building" it is a file copy of `src/gp/*.py` and of the five `src/mpc` modules behind
`GPMPC.solve`; `oracle/ref_gp.py` loads them.

    python oracle/make_ref.py            # also run by __graft_entry__.build() when /root/reference exists
"""
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = "/root/reference/src/gp"
DST = os.path.join(HERE, "_ref", "src", "gp")
# the controller side that imports without third-party code once `casadi` is an empty stub module (oracle/ref_gp.py): the
# reference's own GPMPC.solve loop, its uncertainty propagator and the dataclasses they use
MPC_SRC = "/root/reference/src/mpc"
MPC_DST = os.path.join(HERE, "_ref", "src", "mpc")
MPC_FILES = ("gp_mpc.py", "nominal_mpc.py", "constraints.py", "cost_functions.py", "uncertainty_prop.py")


def make(verbose: bool = True) -> bool:
    if not os.path.isdir(SRC):
        if verbose:
            print(f"oracle/make_ref.py: {SRC} not present (GPU box?): keeping whatever oracle/_ref already holds")
        return os.path.isdir(DST)
    os.makedirs(DST, exist_ok=True)
    n = 0
    for f in sorted(os.listdir(SRC)):
        if f.endswith(".py"):
            shutil.copyfile(os.path.join(SRC, f), os.path.join(DST, f))
            n += 1
    os.makedirs(MPC_DST, exist_ok=True)
    for f in MPC_FILES:
        shutil.copyfile(os.path.join(MPC_SRC, f), os.path.join(MPC_DST, f))
        n += 1
    if verbose:
        print(f"oracle/make_ref.py: {n} files -> {os.path.dirname(DST)}")
    return True


if __name__ == "__main__":
    sys.exit(0 if make() else 1)

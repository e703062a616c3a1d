"""TEST INFRASTRUCTURE ONLY.  Loader for the reference's own `src.gp` package out of `oracle/_ref/` (see oracle/make_ref.py).
Used by bench.py's CPU legs so that the GP part of the timed CPU loop is the reference's code, and by the CPU tests that
check the numpy restatement against it when no golden file covers a case.  Returns None when `oracle/_ref` is absent."""
import importlib.util
import os
import sys
import types

_ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "src")
_NAME = "_gpmpc_ref_src"          # private package name: does not shadow a real `src` on sys.path


def load():
    """-> the reference's `src.gp` module (its __init__ and sub-modules, unmodified) or None."""
    gp_dir = os.path.join(_ROOT, "gp")
    if not os.path.isfile(os.path.join(gp_dir, "__init__.py")):
        return None
    name = _NAME + ".gp"
    if name in sys.modules:
        return sys.modules[name]
    pkg = types.ModuleType(_NAME)
    pkg.__path__ = [_ROOT]       # bypasses src/__init__.py, which drags in the un-vendored `simdyn`
    sys.modules[_NAME] = pkg
    spec = importlib.util.spec_from_file_location(name, os.path.join(gp_dir, "__init__.py"),
                                                  submodule_search_locations=[gp_dir])
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    old = sys.dont_write_bytecode
    sys.dont_write_bytecode = True
    try:
        spec.loader.exec_module(mod)
    finally:
        sys.dont_write_bytecode = old
    return mod


def load_gpmpc():
    """-> the reference's `src.mpc.gp_mpc` module (GPMPC, GPMPCConfig; unmodified) out of oracle/_ref, or None.  The module
    and its siblings do `import casadi` at the top (nominal_mpc.py:30-36 raises without it); CasADi is absent, so an EMPTY stub
    module stands in -- everything except `GPMPC._solve_qp` runs without touching it, and the callers replace that one method."""
    mpc_dir = os.path.join(_ROOT, "mpc")
    if not os.path.isfile(os.path.join(mpc_dir, "gp_mpc.py")) or load() is None:
        return None
    name = _NAME + ".mpc"
    if name + ".gp_mpc" in sys.modules:
        return sys.modules[name + ".gp_mpc"]
    if importlib.util.find_spec("casadi") is None:
        sys.modules["casadi"] = types.ModuleType("casadi")
    pkg = types.ModuleType(name)
    pkg.__path__ = [mpc_dir]          # bypasses src/mpc/__init__.py (osqp, rti_mpc, ...)
    sys.modules[name] = pkg
    old = sys.dont_write_bytecode
    sys.dont_write_bytecode = True
    try:
        return importlib.import_module(name + ".gp_mpc")
    finally:
        sys.dont_write_bytecode = old

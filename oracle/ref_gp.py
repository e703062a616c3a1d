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

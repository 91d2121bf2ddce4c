"""Imports the package directory `3d-genome-cycle_b200/` under the name gc_b200.

    from gcpkg import gc_b200
    sys = gc_b200.System(n_beads, n_replicas)
"""
import importlib.util
import sys
from pathlib import Path

_ROOT = Path(__file__).resolve().parent
_PKG = _ROOT / "3d-genome-cycle_b200"


def _load():
    if "gc_b200" in sys.modules:
        return sys.modules["gc_b200"]
    spec = importlib.util.spec_from_file_location("gc_b200", _PKG / "__init__.py",
                                                  submodule_search_locations=[str(_PKG)])
    mod = importlib.util.module_from_spec(spec)
    sys.modules["gc_b200"] = mod
    spec.loader.exec_module(mod)
    return mod


gc_b200 = _load()

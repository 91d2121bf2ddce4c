"""B200-native force-evaluation + Brownian-integration hot path of 3d-genome-cycle.

The directory name is not a Python identifier; import it through the repo-root
helper:  `from gcpkg import gc_b200`  (modules: api, synth, stages, ensemble, build).
"""
from . import api, build, cycle, decomp, ensemble, prepare, stages, store, synth, transitions  # noqa: F401
from .api import GcError, Observables, System, load_library  # noqa: F401

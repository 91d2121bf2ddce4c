"""In-tree build of libgc_b200.so for sm_100a (nvcc cross-compiles without a GPU)."""
from __future__ import annotations

import os
import subprocess
from pathlib import Path

PKG_DIR = Path(__file__).resolve().parent
CSRC = PKG_DIR / "csrc"
LIB = PKG_DIR / "libgc_b200.so"
SOURCES = ["gc_api.cu", "gc_obs.cu"]
HEADERS = ["gc_types.cuh", "gc_math.cuh", "gc_kernels.cuh", "gc_pairwalk.cuh", "gc_step.cuh", "gc_query.cuh", "gc_tile.cuh", "gc_cells.cuh", "gc_contact.cuh", "gc_decomp.cuh", "gc_transition.cuh", "gc_peak.cuh"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "-shared"]


def _stale():
    if not LIB.exists():
        return True
    t = LIB.stat().st_mtime
    deps = [CSRC / s for s in SOURCES + HEADERS] + [PKG_DIR.parent / "include" / "gc_b200.h"]
    return any(d.stat().st_mtime > t for d in deps)


def build(force=False, verbose=False):
    """Compiles csrc/*.cu into libgc_b200.so.  Returns the library path."""
    if not force and not _stale():
        return LIB
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + \
          ["-o", str(LIB)] + [str(CSRC / s) for s in SOURCES]
    subprocess.run(cmd, check=True, cwd=str(CSRC))
    return LIB

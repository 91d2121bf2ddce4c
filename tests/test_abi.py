"""CPU checks of the drop-in boundary: the C-ABI library builds for sm_100a,
loads, and exports every symbol include/gc_b200.h declares.  No compute calls."""
import ctypes
import re
from pathlib import Path

import pytest

from gcpkg import gc_b200

ROOT = Path(__file__).resolve().parent.parent


@pytest.fixture(scope="module")
def lib():
    gc_b200.build.build()
    return gc_b200.api.load_library()


def declared_symbols():
    text = (ROOT / "include" / "gc_b200.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(gc_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_are_exported(lib):
    names = declared_symbols()
    assert len(names) >= 35
    for n in names:
        assert hasattr(lib, n), f"{n} declared in gc_b200.h but not exported"
    bound = {n for n, _, _ in gc_b200.api.SYMBOLS}
    assert bound == set(names), (bound ^ set(names))


def test_struct_layouts_match_header():
    A = gc_b200.api
    assert ctypes.sizeof(A.BondParams) == 48
    assert ctypes.sizeof(A.WallParams) == 3 * 8 + 7 * 8 + 8 + 8 + 3 * 8
    assert ctypes.sizeof(A.Schedule) == 40
    assert ctypes.sizeof(A.Context) == 8 * 14
    assert ctypes.sizeof(A.RunConfig) == 8 * 4 + 8 + 8


def test_version_and_null_handling(lib):
    assert b"sm_100a" in lib.gc_version()
    assert lib.gc_destroy(None) < 0
    assert lib.gc_last_error(None) == b"null context"


def test_no_cpu_fallback(lib):
    """Without a CUDA device gc_create must fail (there is no CPU path)."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    h = ctypes.c_void_p()
    rc = lib.gc_create(10, 1, 0, ctypes.byref(h))
    assert rc == -2 and not h
    with pytest.raises(gc_b200.GcError):
        gc_b200.System(10)


def test_product_does_not_reference_oracle():
    """The product path must not import, link or execute anything under oracle/."""
    pkg = ROOT / "3d-genome-cycle_b200"
    for f in list(pkg.rglob("*.py")) + list(pkg.rglob("*.cu*")) + list(pkg.rglob("*.hpp")) + \
            list(pkg.rglob("*.cpp")):
        t = f.read_text()
        assert "oracle_py" not in t and "gco_" not in t and "libgc_ref" not in t and \
            "libgc_port" not in t, f

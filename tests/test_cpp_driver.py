"""The C++ host side: include/gc_md.hpp (md:: mirror over the C-ABI) and the
interphase stage driver re-hosted on it (examples/interphase_driver.cpp, structured
like SIM/stage_interphase/simulation_driver*.cpp)."""
import struct
import subprocess
from pathlib import Path

import numpy as np
import pytest

from gcpkg import gc_b200
from helpers import small_interphase, stages

ROOT = Path(__file__).resolve().parent.parent
SRC = ROOT / "examples" / "interphase_driver.cpp"
EXE = ROOT / "examples" / "interphase_driver"


def build_driver():
    lib = gc_b200.build.build()
    deps = [SRC, ROOT / "include" / "gc_md.hpp", ROOT / "include" / "gc_b200.h"]
    if EXE.exists() and all(EXE.stat().st_mtime >= d.stat().st_mtime for d in deps):
        return EXE
    subprocess.run(["g++", "-std=c++20", "-O2", "-Wall", "-o", str(EXE), str(SRC),
                    f"-L{lib.parent}", "-lgc_b200", "-Wl,-rpath," + str(lib.parent)], check=True)
    return EXE


def write_design(path, d, x):
    with open(path, "wb") as f:
        f.write(b"GCD1")
        f.write(struct.pack("<iii", d.n, len(d.chains), len(d.nucleolar_bonds)))
        for a in (d.mobility, d.a, d.b):
            f.write(np.ascontiguousarray(a, dtype="<f8").tobytes())
        f.write(np.ascontiguousarray(d.chains, dtype="<i4").tobytes())
        f.write(np.ascontiguousarray(d.nucleolar_bonds, dtype="<i4").tobytes())
        f.write(np.ascontiguousarray(x, dtype="<f8").tobytes())


def read_output(path):
    raw = Path(path).read_bytes()
    assert raw[:4] == b"GCO1"
    n, m = struct.unpack_from("<ii", raw, 4)
    off = 12
    pos = np.frombuffer(raw, dtype="<f8", count=3 * n, offset=off).reshape(n, 3)
    off += 24 * n
    tail = np.frombuffer(raw, dtype="<f8", count=7, offset=off)
    off += 56
    contacts = np.frombuffer(raw, dtype="<u4", count=3 * m, offset=off).reshape(m, 3)
    return pos, tail, contacts


def test_driver_builds_and_fails_loudly_without_gpu(tmp_path):
    exe = build_driver()
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    d, cfg, x = small_interphase(scale=0.01, radius=0.5)
    write_design(tmp_path / "in.bin", d, x)
    p = subprocess.run([str(exe), str(tmp_path / "in.bin"), str(tmp_path / "out.bin"), "1", "1", "1"],
                       capture_output=True, text=True)
    assert p.returncode == 1                      # std::exception -> "error: ..." + exit 1,
    assert "error:" in p.stderr                   # as SIM/stage_prepare/main.cpp:64-70 does
    assert not (tmp_path / "out.bin").exists()


@pytest.mark.gpu
def test_cpp_driver_matches_python_driver(tmp_path):
    """Same calls through two host languages -> bit-identical device results."""
    exe = build_driver()
    d, cfg, x = small_interphase(scale=0.02, radius=0.6)
    relax, steps, seed = 60, 100, 77
    write_design(tmp_path / "in.bin", d, x)
    p = subprocess.run([str(exe), str(tmp_path / "in.bin"), str(tmp_path / "out.bin"), str(relax),
                        str(steps), str(seed), "50"], capture_output=True, text=True)
    assert p.returncode == 0, p.stderr
    assert p.stderr.count("[relaxation]") == 2 and p.stderr.count("[interphase]") == 2
    pos, tail, contacts = read_output(tmp_path / "out.bin")

    g = gc_b200.System(d.n, 1, 0)
    stages.setup_interphase(g, d, cfg)
    g.set_positions(x)
    g.set_seeds([seed])
    stages.begin_relaxation(g, cfg)
    g.run(relax, cfg.timestep, spacestep=cfg.relaxation_spacestep)
    stages.begin_interphase(g, cfg)
    g.run(steps, cfg.timestep)
    np.testing.assert_array_equal(pos, g.get_positions())
    np.testing.assert_array_equal(contacts, g.contacts_fetch())
    c = g.get_context()
    assert tail[0] == c.time and tail[4] == c.core_scale and tail[5] == c.bond_scale
    assert list(tail[1:4]) == list(c.semiaxes)
    assert tail[6] == pytest.approx(g.compute_energy() / d.n, rel=1e-12)
    assert len(contacts) > 0

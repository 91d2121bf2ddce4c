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


def build_driver(name="interphase_driver"):
    lib = gc_b200.build.build()
    src, exe = ROOT / "examples" / f"{name}.cpp", ROOT / "examples" / name
    deps = [src, ROOT / "examples" / "mitotic_common.hpp", ROOT / "include" / "gc_md.hpp",
            ROOT / "include" / "gc_b200.h"]
    if exe.exists() and all(exe.stat().st_mtime >= d.stat().st_mtime for d in deps):
        return exe
    subprocess.run(["g++", "-std=c++20", "-O2", "-Wall", "-o", str(exe), str(src),
                    f"-L{lib.parent}", "-lgc_b200", "-Wl,-rpath," + str(lib.parent)], check=True)
    return exe


def write_mitotic_design(path, md, x):
    with open(path, "wb") as f:
        f.write(b"GCM1")
        f.write(struct.pack("<iii", md.n, len(md.chains), len(md.sisters)))
        f.write(np.ascontiguousarray(md.chains, dtype="<i4").tobytes())
        f.write(np.ascontiguousarray(md.kinetochores, dtype="<i4").tobytes())
        f.write(np.ascontiguousarray(md.sisters, dtype="<i4").tobytes())
        f.write(np.ascontiguousarray(md.poles, dtype="<f8").tobytes())
        f.write(np.ascontiguousarray(x, dtype="<f8").tobytes())


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


def _mitotic_inputs(sisters):
    synth = gc_b200.synth
    d, _, _ = small_interphase(scale=0.3, radius=1.0)
    mcfg = synth.MitoticConfig()
    md = synth.make_mitotic_design(d, mcfg, sisters=sisters)
    x = synth.quantize_f32(synth.rod_positions(md, mcfg, seed=3)).astype(np.float64)
    return md, mcfg, x


def test_mitotic_drivers_build_and_fail_loudly_without_gpu(tmp_path):
    import torch
    exes = [build_driver("anatelophase_driver"), build_driver("prometaphase_driver")]
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    md, mcfg, x = _mitotic_inputs(sisters=False)
    write_mitotic_design(tmp_path / "in.bin", md, x)
    for exe, tail in zip(exes, (["1", "1", "1", "2"], ["1", "1"])):
        p = subprocess.run([str(exe), str(tmp_path / "in.bin"), str(tmp_path / "out.bin")] + tail,
                           capture_output=True, text=True)
        assert p.returncode == 1 and "error:" in p.stderr
        assert not (tmp_path / "out.bin").exists()


@pytest.mark.gpu
def test_cpp_anatelophase_driver_matches_python_driver(tmp_path):
    """Dragging stage, then the runtime swap dragging -> packing and anaphase -> telophase
    bonded terms (disable()/enable() of registered force fields): bit-identical to the same
    calls from stages.py."""
    exe = build_driver("anatelophase_driver")
    md, mcfg, x = _mitotic_inputs(sisters=False)
    n_ana, n_telo, s_ana, s_telo = 150, 90, 11, 12
    write_mitotic_design(tmp_path / "in.bin", md, x)
    p = subprocess.run([str(exe), str(tmp_path / "in.bin"), str(tmp_path / "out.bin"), str(n_ana),
                        str(n_telo), str(s_ana), str(s_telo), "50"], capture_output=True, text=True)
    assert p.returncode == 0, p.stderr
    assert p.stderr.count("[anaphase]") == 3 and p.stderr.count("[telophase]") == 2
    pos, tail, _ = read_output(tmp_path / "out.bin")

    g = gc_b200.System(md.n, 1, 0)
    t = stages.setup_anatelophase(g, md, mcfg)
    g.set_positions(x)
    stages.enter_anaphase(g, t)
    g.set_seeds([s_ana])
    g.run(n_ana, mcfg.timestep)
    stages.enter_telophase(g, t)
    g.set_seeds([s_telo])
    g.run(n_telo, mcfg.timestep)
    np.testing.assert_array_equal(pos, g.get_positions())
    assert tail[0] == g.get_context().time
    assert tail[6] == pytest.approx(g.compute_energy() / md.n, rel=1e-12)
    assert np.abs(pos - x).max() > 0.05          # the stage did move the chromosomes
    g.close()


@pytest.mark.gpu
def test_cpp_prometaphase_driver_matches_python_driver(tmp_path):
    """Sister bonds, one k-fibre force field per pole and polar ejection with f != 0."""
    import dataclasses
    exe = build_driver("prometaphase_driver")
    md, mcfg, x = _mitotic_inputs(sisters=True)
    mcfg = dataclasses.replace(mcfg, polar_ejection_force=0.8, polar_ejection_cross_section=0.25)
    steps, seed = 160, 21
    write_mitotic_design(tmp_path / "in.bin", md, x)
    p = subprocess.run([str(exe), str(tmp_path / "in.bin"), str(tmp_path / "out.bin"), str(steps),
                        str(seed), "40", "0.8", "0.25"], capture_output=True, text=True)
    assert p.returncode == 0, p.stderr
    assert p.stderr.count("[prometaphase]") == 4
    pos, tail, _ = read_output(tmp_path / "out.bin")

    g = gc_b200.System(md.n, 1, 0)
    stages.setup_prometaphase(g, md, mcfg)
    g.set_positions(x)
    g.set_seeds([seed])
    g.run(steps, mcfg.timestep)
    np.testing.assert_array_equal(pos, g.get_positions())
    assert tail[6] == pytest.approx(g.compute_energy() / md.n, rel=1e-12)
    g.close()

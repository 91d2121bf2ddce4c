"""Topology preparation library (3d-genome-cycle_b200/prepare.py) against the rules of the
reference's `prepare init` (SIM/stage_prepare/prepare.cpp, chains.cpp) and, for the seed
arithmetic, against libstdc++ itself."""
import subprocess
import sys

import numpy as np
import pytest

from gcpkg import gc_b200

synth, P = gc_b200.synth, gc_b200.prepare

PROBE = r'''
#include <array>
#include <cstdint>
#include <cstdio>
#include <random>
int main(int argc, char **argv) {
    for (int k = 1; k < argc; k++) {
        std::uint32_t master = std::uint32_t(std::stoul(argv[k]));
        std::seed_seq seq = { master };                       // prepare.cpp:549-553
        std::array<std::uint32_t, 3> v;
        seq.generate(v.begin(), v.end());
        std::printf("%u %u %u", v[0], v[1], v[2]);
        for (int s = 0; s < 3; s++) {                         // drivers: std::mt19937_64 _random(design.seed)
            std::mt19937_64 r(v[s]);
            unsigned long long first = r(), second = r();     // (argument evaluation order is unspecified)
            std::printf(" %llu %llu", first, second);
        }
        std::printf("\n");
    }
}
'''


def test_seed_chain_matches_libstdcxx(tmp_path):
    src = tmp_path / "probe.cpp"
    src.write_text(PROBE)
    exe = tmp_path / "probe"
    subprocess.run(["g++", "-std=c++17", "-O1", "-o", str(exe), str(src)], check=True)
    masters = [0, 1, 1000, 1001, 1063, 20250001, 2 ** 32 - 1]
    out = subprocess.run([str(exe)] + [str(m) for m in masters], check=True, capture_output=True, text=True).stdout
    for m, line in zip(masters, out.strip().splitlines()):
        f = [int(t) for t in line.split()]
        s = P.stage_seeds(m)
        assert [s["anaphase"], s["interphase"], s["prometaphase"]] == f[:3], m
        r = P.run_seeds(m)
        assert [r["anaphase"], r["telophase"], r["relaxation"], r["interphase"]] == f[3:7], m
        assert r["prometaphase"] == f[7], m
    # long generate() calls take the other parameter branches of the algorithm
    assert len(P.seed_seq_generate([1, 2, 3], 700)) == 700 and len(set(P.seed_seq_generate([5], 40))) > 30


def test_mt19937_64_known_answer():
    """ISO C++ [rand.predef]: the 10000th invocation of a default-constructed mt19937_64."""
    r = P.MT19937_64()
    for _ in range(9999):
        r()
    assert r() == 9981545732273789042


def test_tsv_round_trip_and_prepare_reproduce_the_synthetic_design():
    cfg, mcfg = synth.InterphaseConfig(), synth.MitoticConfig()
    d = synth.make_diploid_design(scale=1.0, cfg=cfg)
    text = P.format_chains_tsv(P.chains_from_design(d))
    chains = P.parse_chains_tsv(text)
    assert [c.name for c in chains] == d.names and sum(len(c.beads) for c in chains) == int(d.chains[-1, 1])
    p = P.prepare(chains, cfg, mcfg, master_seed=1000)
    i = p.interphase
    assert i.n == d.n
    np.testing.assert_array_equal(i.chains, d.chains)
    np.testing.assert_array_equal(i.a, d.a)
    np.testing.assert_array_equal(i.b, d.b)
    np.testing.assert_array_equal(i.mobility, d.mobility)
    np.testing.assert_array_equal(i.nucleolar_bonds, d.nucleolar_bonds)
    np.testing.assert_array_equal(i.centromeres, d.centromeres)
    assert p.particle_types.count("nucleolus") == len(d.nucleolar_bonds) == 30
    assert i.seed == P.stage_seeds(1000)["interphase"]
    # coarse chains: floor(len / 100) beads, kinetochore at the centromere midpoint
    ref = synth.make_mitotic_design(d, mcfg)
    np.testing.assert_array_equal(p.anatelophase.chains, ref.chains)
    np.testing.assert_array_equal(p.anatelophase.kinetochores, ref.kinetochores)
    assert p.mitotic_types.count("kinetochore") == 46
    # prometaphase: reference layout, target c at 2 * start_c and its copy right behind it
    pro = p.prometaphase
    assert pro.n == 2 * ref.n and p.prometaphase_chain_names[:2] == ["chr1:a", "chr1:a-copy"]
    for c, (s, e) in enumerate(ref.chains):
        t, sis = pro.sisters[c]
        assert (t, sis) == (2 * c, 2 * c + 1)
        assert tuple(pro.chains[t]) == (2 * s, 2 * s + (e - s)) and tuple(pro.chains[sis]) == (2 * s + (e - s), 2 * e)
        assert pro.kinetochores[t] - pro.chains[t][0] == ref.kinetochores[c] - s == pro.kinetochores[sis] - pro.chains[sis][0]
    np.testing.assert_array_equal(pro.poles, [[0, -5, 0], [0, 5, 0]])


def test_tag_rules_and_errors():
    assert P.check_tag("cen,A", "A") and P.check_tag("anor", "anor") and not P.check_tag("anor", "nor")
    assert not P.check_tag("AB", "A") and not P.check_tag("", "A")
    text = "chain\tstart\tend\ta\tb\ttags\nc1\t0\t100\t1\t0\tA\nc1\t100\t200\t0\t1\tbnor,B\nc2\t0\t100\t0.5\t0.5\t\n"
    chains = P.parse_chains_tsv(text)
    assert [c.name for c in chains] == ["c1", "c2"] and [len(c.beads) for c in chains] == [2, 1]
    mc = synth.MitoticConfig(coarse_graining=1)
    p = P.prepare(chains, mcfg=mc)
    assert p.particle_types == ["A", "bnor", "unknown"]       # bnor wins over B; no tag -> unknown
    assert len(p.nor_indices) == 0 and p.interphase.n == 3
    with pytest.raises(ValueError, match="header"):
        P.parse_chains_tsv("c1\t0\t100\t1\t0\tA\n")
    with pytest.raises(ValueError, match="expected 6 fields"):
        P.parse_chains_tsv("chain\tstart\tend\ta\tb\ttags\nc1\t0\t100\t1\t0\n")
    with pytest.raises(ValueError, match="line 2"):
        P.parse_chains_tsv("chain\tstart\tend\ta\tb\ttags\nc1\tx\t100\t1\t0\tA\n")
    with pytest.raises(ValueError, match="kinetochore"):
        P.prepare(chains, mcfg=synth.MitoticConfig(coarse_graining=100))   # chains shorter than one coarse bead

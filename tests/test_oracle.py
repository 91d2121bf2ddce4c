"""CPU tests that PIN THE ORACLE (no GPU).

1. Known-answer vectors copied from the reference's own test-suite
   (micromd tests, MDT/ = submodules/github.com/snsinfu/micromd/tests/) are
   checked against both oracles.
2. The port (FP64 restatement) is checked against the compiled reference
   (micromd unmodified) on seeded systems of every stage.
3. The restated integrator loop with injectable noise is proven bit-identical
   to unmodified md::simulate_brownian_dynamics.
"""
import math

import numpy as np
import pytest

import oracle_py as O
from helpers import build_pair, rel_force_err, small_interphase, stages, synth

KINDS = [k for k in ("port", "reference", "strict") if O.available(k) or k == "port"]


@pytest.fixture(scope="module", autouse=True)
def _build():
    O.build()


def approx(x, rel=1e-5):
    return pytest.approx(x, rel=rel)


# ---------------------------------------------------------------------------
# 1. golden vectors of the reference's own tests
# ---------------------------------------------------------------------------
@pytest.mark.parametrize("kind", KINDS)
def test_softcore_golden(kind):
    # MDT/potential/test_softcore_potential.cc:55-76 (energy = diameter = 1, r = 0.5)
    table = {21: (0.75, 1.0), 22: (0.5625, 1.5), 33: (0.669921875, 1.72265625),
             44: (0.7724761962890625, 1.64794921875), 55: (0.8532151877880096, 1.3761535286903381)}
    for pq, (e_ref, f_ref) in table.items():
        e, f = O.potential_probe(8, 1.0, 1.0, pq, [0.5, 0, 0], kind)
        assert e == approx(e_ref, 1e-12)
        assert f[0] == approx(f_ref, 1e-12)
    # the two instantiations the drivers use, from the closed form
    eps, sig, r = 2.5, 0.3, np.array([0.1, -0.05, 0.12])
    u2 = r.dot(r) / sig ** 2
    e, f = O.potential_probe(0, eps, sig, 0, r, kind)
    assert e == approx(eps * (1 - u2) ** 3, 1e-12)
    np.testing.assert_allclose(f, 6 * eps / sig ** 2 * (1 - u2) ** 2 * r, rtol=1e-12)
    e, f = O.potential_probe(1, eps, sig, 0, r, kind)
    assert e == approx(eps * (1 - u2 ** 4) ** 3, 1e-12)
    np.testing.assert_allclose(f, 24 * eps / sig ** 2 * (1 - u2 ** 4) ** 2 * u2 ** 3 * r, rtol=1e-12)
    # zero beyond the diameter (test_softcore_potential.cc:40-53)
    e, f = O.potential_probe(0, eps, sig, 0, 10 * r, kind)
    assert e == 0 and not f.any()


@pytest.mark.parametrize("kind", KINDS)
def test_softwell_golden(kind):
    # MDT/potential/test_softwell_potential.cc:74-89
    e, f = O.potential_probe(2, 1.23, 4.56, 0, [1.2, 3.4, 5.6], kind)
    assert e == approx(-0.11485401188866758, 1e-9)
    np.testing.assert_allclose(f, [-0.016901052221443922, -0.047886314627424445,
                                   -0.07887157703340497], rtol=1e-9)
    # energy at the decay distance is -eps / 2 (same file, 47-54)
    e, _ = O.potential_probe(2, 1.23, 4.56, 0, [4.56, 0, 0], kind)
    assert e == approx(-1.23 / 2, 1e-12)
    # cutoff is exclusive: r^2 >= cut^2 -> 0 (MD/potential/cutoff_potential.hpp:45-48)
    e, f = O.potential_probe(7, 0.3, 0.2, 0.4, [0.4, 0, 0], kind)
    assert e == 0 and not f.any()
    e, f = O.potential_probe(7, 0.3, 0.2, 0.4, [0.39999, 0, 0], kind)
    assert e < 0 and f[0] < 0


@pytest.mark.parametrize("kind", KINDS)
def test_spring_family_golden(kind):
    # MDT/potential/test_semispring_potential.cc:56-73 and siblings
    k, b, r = 1.23, 4.56, np.array([-4.0, 5.0, -6.0])
    rn = np.linalg.norm(r)
    for pot in (3, 5):
        e, f = O.potential_probe(pot, k, b, 0, r, kind)
        assert e == approx(k / 2 * (rn - b) ** 2, 1e-12)
        np.testing.assert_allclose(f, -k * (rn - b) * r / rn, rtol=1e-12)
    short = np.array([1.0, 1.0, 1.0])
    e, f = O.potential_probe(3, k, b, 0, short, kind)      # semispring inside: zero
    assert e == 0 and not f.any()
    e, f = O.potential_probe(5, k, b, 0, short, kind)      # spring inside: pushes out
    assert e > 0 and f[0] > 0
    e, f = O.potential_probe(4, k, 0, 0, r, kind)          # harmonic
    assert e == approx(0.5 * k * rn ** 2, 1e-12)
    np.testing.assert_allclose(f, -k * r, rtol=1e-12)
    # force_flux reduces to f b^2 / r^2 far away (force_flux_potential.hpp:5-8)
    e, f = O.potential_probe(6, 2.0, 0.1, 0, [50.0, 0, 0], kind)
    assert f[0] == approx(2.0 * 0.01 / 2500.0, 1e-4)


@pytest.mark.parametrize("kind", KINDS)
def test_cosine_bending_golden(kind):
    # MDT/potential/test_cosine_bending_potential.cc:34-56
    e, f = O.bending_probe(1.23, [1, 2, 3], [4, 3, 2], kind)
    assert e == approx(1.23 * 0.205933, 1e-5)
    ref = [[0.174411, 0.0436028, -0.0872055], [-0.248085, -0.0225532, 0.202978],
           [0.0736736, -0.0210496, -0.115773]]
    np.testing.assert_allclose(f, ref, rtol=2e-5)
    e, f = O.bending_probe(1.23, [0, 0, 0], [4, 3, 2], kind)   # zero-length arm (24-33)
    assert e == 0 and not f.any()


@pytest.mark.parametrize("kind", KINDS)
def test_ellipsoid_evaluate_point_golden(kind):
    # MDT/forcefield/test_ellipsoid_surface_forcefield.cc:31-95 (unit sphere)
    one = [1, 1, 1]
    assert O.ellipsoid_probe(one, [0, 0, 0], kind)["undefined"]
    ev = O.ellipsoid_probe(one, [1, 0, 0], kind)
    assert not ev["undefined"] and not ev["delta"].any() and ev["implicit"] == 0
    ev = O.ellipsoid_probe(one, [0.99, 0, 0], kind)
    assert ev["delta"][0] == pytest.approx(-0.01, rel=0.1)
    np.testing.assert_allclose(ev["strain"], [-0.0101520] * 3, rtol=1e-5)
    assert ev["implicit"] == approx(-0.0199)
    ev = O.ellipsoid_probe(one, [1.01, 0, 0], kind)
    np.testing.assert_allclose(ev["strain"], [0.00985198] * 3, rtol=1e-5)
    assert ev["implicit"] == approx(0.0201)


@pytest.mark.parametrize("kind", KINDS)
def test_neighbor_search_vs_bruteforce(kind):
    # MDT/misc/test_neighbor_searcher.cc:54-127: 100 normal points, r = 0.3,
    # result must equal the brute-force set (no duplicates, no misses)
    rng = np.random.default_rng(0)
    for n, r in ((100, 0.3), (1000, 0.1), (2, 5.0), (1, 1.0)):
        x = rng.normal(size=(n, 3)) if n != 1000 else rng.random((n, 3))
        s = O.OracleSystem(n, kind)
        s.set_bead_attrs(np.ones(n))
        s.set_positions(x)
        got = s.neighbor_pairs(r)
        i, j = np.triu_indices(n, 1)
        d2 = ((x[i] - x[j]) ** 2).sum(1)
        want = np.stack([i, j], 1)[d2 <= r * r]
        assert len(got) == len(want)
        assert len(np.unique(got, axis=0)) == len(got)
        np.testing.assert_array_equal(got, want)


# ---------------------------------------------------------------------------
# 2. port vs compiled reference
# ---------------------------------------------------------------------------
needs_ref = pytest.mark.skipif(not O.available("reference"),
                               reason="oracle/_ref not built (no /root/reference here)")


def _pair(kind_a, kind_b, setup, design, cfg_or_none, x):
    out = []
    for k in (kind_a, kind_b):
        s = O.OracleSystem(design.n, k)
        terms = setup(s)
        s.set_positions(x)
        out.append((s, terms))
    return out


@needs_ref
def test_port_matches_reference_interphase_forces():
    d, cfg, x = small_interphase(scale=0.04, radius=1.2)
    # push some beads outside the wall so the outward branch is exercised
    x = x.copy()
    x[::97] *= 1.75
    (p, terms), (r, _) = [build_pair(lambda n, k=k: O.OracleSystem(n, k), d, cfg, x)
                          for k in ("port", "reference")]
    for s in (p, r):
        s.set_context(core_scale=0.73, bond_scale=0.61, semiaxes=(2.0, 1.8, 2.1))
    for name, t in terms.items():
        Fp, Fr = p.compute_force(1 << t), r.compute_force(1 << t)
        assert rel_force_err(Fp, Fr) < 1e-12, name
        Ep, Er = p.compute_energy(1 << t), r.compute_energy(1 << t)
        assert Ep == pytest.approx(Er, rel=1e-11, abs=1e-9), name
    assert rel_force_err(p.compute_force(), r.compute_force()) < 1e-12
    cp, cr = p.get_context(), r.get_context()
    np.testing.assert_allclose(list(cp.axial_reaction), list(cr.axial_reaction), rtol=1e-10)
    assert cp.reaction_force == pytest.approx(cr.reaction_force, rel=1e-10)
    assert abs(cr.reaction_force) > 0
    np.testing.assert_array_equal(p.neighbor_pairs(0.2), r.neighbor_pairs(0.2))
    # the reference's Verlet list is a superset of the exact set at dcut
    # (MDT/forcefield/detail/test_neighbor_list.cc:65-69) and equals the exact
    # set at 1.5 dcut
    dcut = 0.3 * 0.73
    exact = {tuple(q) for q in r.neighbor_pairs(dcut)}
    verlet = {tuple(q) for q in r.verlet_pairs(terms["repulsion"])}
    assert exact <= verlet
    np.testing.assert_array_equal(r.verlet_pairs(terms["repulsion"]),
                                  p.verlet_pairs(terms["repulsion"]))


@needs_ref
def test_port_matches_reference_mitotic_forces():
    d = synth.make_diploid_design(scale=0.3)
    mcfg = synth.MitoticConfig(polar_ejection_force=0.7, polar_ejection_cross_section=0.04,
                               kfiber_length_anaphase=0.2, kfiber_length_prometaphase=0.3,
                               telophase_bond_spring_multiplier=0.5,
                               telophase_bending_energy_multiplier=2.0)
    rng = np.random.default_rng(2)
    # anatelophase
    md = synth.make_mitotic_design(d, mcfg)
    x = synth.rod_positions(md, mcfg) + rng.normal(size=(md.n, 3)) * 0.05
    sysms = []
    for k in ("port", "reference"):
        s = O.OracleSystem(md.n, k)
        t = stages.setup_anatelophase(s, md, mcfg)
        s.set_positions(x)
        sysms.append(s)
    p, r = sysms
    stages.enter_anaphase(p, t), stages.enter_anaphase(r, t)
    assert rel_force_err(p.compute_force(), r.compute_force()) < 1e-12
    assert p.compute_energy() == pytest.approx(r.compute_energy(), rel=1e-11)
    stages.enter_telophase(p, t), stages.enter_telophase(r, t)
    assert rel_force_err(p.compute_force(), r.compute_force()) < 1e-12
    assert p.compute_energy() == pytest.approx(r.compute_energy(), rel=1e-11)
    # prometaphase
    md = synth.make_mitotic_design(d, mcfg, sisters=True)
    x = rng.normal(size=(md.n, 3)) * 1.5
    sysms = []
    for k in ("port", "reference"):
        s = O.OracleSystem(md.n, k)
        t = stages.setup_prometaphase(s, md, mcfg)
        s.set_positions(x)
        sysms.append(s)
    p, r = sysms
    for name, tid in t.items():
        assert rel_force_err(p.compute_force(1 << tid), r.compute_force(1 << tid)) < 1e-12, name
        assert p.compute_energy(1 << tid) == pytest.approx(r.compute_energy(1 << tid), rel=1e-11,
                                                           abs=1e-12), name


@needs_ref
def test_port_matches_reference_trajectory_and_contacts():
    d, cfg, x = small_interphase(scale=0.02, radius=0.6)
    steps = 40
    z = np.random.default_rng(7).normal(size=(steps, d.n, 3))
    res = []
    for k in ("port", "reference"):
        s, _ = build_pair(lambda n, k=k: O.OracleSystem(n, k), d, cfg, x)
        stages.interphase_callback0(s, cfg)
        out = s.run(steps, cfg.timestep, noise=z, interphase_callback=True, want_traj=True,
                    want_ctx=True)
        res.append((s, out))
    (p, op), (r, orr) = res
    np.testing.assert_allclose(op["traj"], orr["traj"], rtol=0, atol=1e-11)
    np.testing.assert_allclose(op["ctx"], orr["ctx"], rtol=1e-12, atol=1e-15)
    cp, cr = p.contacts_fetch(), r.contacts_fetch()
    assert len(cr) > 0
    np.testing.assert_array_equal(cp, cr)
    # wall moved and the schedule advanced
    assert orr["ctx"][-1, 4] > cfg.core_scale_init
    assert orr["ctx"][-1, 1] != cfg.wall_semiaxes_init[0]


@needs_ref
def test_port_matches_reference_adaptive_relaxation():
    d, cfg, x = small_interphase(scale=0.02, radius=0.6)
    steps = 15
    z = np.random.default_rng(8).normal(size=(steps, d.n, 3))
    outs = []
    for k in ("port", "reference"):
        s, _ = build_pair(lambda n, k=k: O.OracleSystem(n, k), d, cfg, x)
        outs.append(s.run(steps, cfg.timestep, spacestep=cfg.relaxation_spacestep, noise=z,
                          want_traj=True, want_ctx=True))
    np.testing.assert_allclose(outs[0]["traj"], outs[1]["traj"], rtol=0, atol=1e-11)
    np.testing.assert_allclose(outs[0]["ctx"][:, 6], outs[1]["ctx"][:, 6], rtol=1e-10)
    assert (outs[1]["ctx"][:, 6] < cfg.timestep).all()      # dt was actually adapted


# ---------------------------------------------------------------------------
# 3. restated integrator == md::simulate_brownian_dynamics
# ---------------------------------------------------------------------------
@needs_ref
@pytest.mark.parametrize("spacestep", [0.0, 0.001])
def test_restated_stepper_matches_reference(spacestep):
    """The 15-line simulate_step restatement with the noise source as a
    parameter, fed sfc64 + ziggurat draws in the same order, is bit-identical
    to unmodified md::simulate_brownian_dynamics."""
    d, cfg, x = small_interphase(scale=0.02, radius=0.6)
    steps, seed = 25, 12345
    res = []
    for stepper in (0, 1):
        s, _ = build_pair(lambda n: O.OracleSystem(n, "reference"), d, cfg, x)
        stages.interphase_callback0(s, cfg)
        out = s.run(steps, cfg.timestep, spacestep=spacestep, seed=seed, interphase_callback=True,
                    want_traj=True, stepper=stepper)
        res.append(out["traj"])
    assert np.array_equal(res[0], res[1])
    # and injecting the same draws as explicit noise gives the same trajectory
    z = O.reference_normals(seed, steps * d.n * 3).reshape(steps, d.n, 3)
    s, _ = build_pair(lambda n: O.OracleSystem(n, "reference"), d, cfg, x)
    stages.interphase_callback0(s, cfg)
    out = s.run(steps, cfg.timestep, spacestep=spacestep, noise=z, interphase_callback=True,
                want_traj=True)
    assert np.array_equal(res[0], out["traj"])


@needs_ref
def test_reference_normals_are_standard():
    z = O.reference_normals(1, 200000)
    assert abs(z.mean()) < 0.01 and abs(z.std() - 1) < 0.01

"""GPU parity tests: the CUDA path through the C-ABI vs the CPU oracle on the
same seeded inputs.  Bars (BASELINE.json north_star):
  * neighbor pair sets and contact triples: bit-exact
  * forces / positions under identical injected noise: <= 1e-5 relative
    (FP32 device vs FP64 reference); force norm: max_i|dF_i| / max_i|F_i|,
    position norm: max|dx| / nucleus size (2.0)
"""
import numpy as np
import pytest

import oracle_py as O
from gcpkg import gc_b200
from helpers import build_pair, rel_force_err, small_interphase, stages, synth

pytestmark = pytest.mark.gpu

FORCE_TOL = 1e-5
POS_TOL = 1e-5 * 2.0


def gpu(n, r=1):
    return gc_b200.System(n, r, 0)


def oracle(kind):
    return lambda n: O.OracleSystem(n, kind)


# ---------------------------------------------------------------------------
# neighbor search
# ---------------------------------------------------------------------------
@pytest.mark.parametrize("n,radius,dist", [(100, 0.3, "normal"), (1000, 0.1, "uniform"),
                                           (5000, 0.15, "normal"), (2, 5.0, "normal"),
                                           (1, 1.0, "normal")])
def test_pair_set_exact_random_points(oracle_kind, n, radius, dist):
    """MDT/misc/test_neighbor_searcher.cc:54-127 shapes; FP32-representable input."""
    rng = np.random.default_rng(n)
    x = rng.normal(size=(n, 3)) if dist == "normal" else rng.random((n, 3))
    x = x.astype(np.float32).astype(np.float64)
    g, o = gpu(n), O.OracleSystem(n, oracle_kind)
    for s in (g, o):
        s.set_bead_attrs(np.ones(n))
        s.set_positions(x)
    np.testing.assert_array_equal(g.neighbor_pairs(radius), o.neighbor_pairs(radius))


def test_pair_set_exact_interphase_state(oracle_kind):
    d, cfg, x = small_interphase()
    g, o = gpu(d.n), O.OracleSystem(d.n, oracle_kind)
    for s in (g, o):
        stages.setup_interphase(s, d, cfg)
        s.set_positions(x)
    for radius in (0.15, 0.225, 0.12):     # dcut, Verlet radius, contact distance at s = 0.5
        a, b = g.neighbor_pairs(radius), o.neighbor_pairs(radius)
        assert len(b) > d.n
        np.testing.assert_array_equal(a, b)


def test_pair_set_borderline_pairs(oracle_kind):
    """Pairs placed exactly at / one ulp around the cutoff (inclusive <=,
    MD/misc/neighbor_searcher.hpp:90)."""
    r = 0.25
    base = np.array([[0, 0, 0], [r, 0, 0], [0, np.nextafter(np.float32(r), np.float32(1)), 0],
                     [0, 0, np.nextafter(np.float32(r), np.float32(0))], [3, 3, 3],
                     [3 + r, 3, 3]], dtype=np.float32).astype(np.float64)
    g, o = gpu(len(base)), O.OracleSystem(len(base), oracle_kind)
    for s in (g, o):
        s.set_bead_attrs(np.ones(len(base)))
        s.set_positions(base)
    a, b = g.neighbor_pairs(r), o.neighbor_pairs(r)
    np.testing.assert_array_equal(a, b)
    assert [0, 1] in b.tolist() and [0, 2] not in b.tolist() and [0, 3] in b.tolist()


# ---------------------------------------------------------------------------
# forces, energies, wall statistics
# ---------------------------------------------------------------------------
@pytest.mark.parametrize("core,bond,semi", [(0.5, 0.5, (2.0, 2.0, 2.0)),
                                            (0.73, 0.61, (2.0, 1.8, 2.1)),
                                            (1.0, 1.0, (2.3, 2.2, 2.4))])
def test_interphase_forces_per_term(oracle_kind, core, bond, semi):
    d, cfg, x = small_interphase(scale=0.05, radius=1.3)
    x = x.copy()
    x[::53] *= 1.75                     # some beads beyond the wall -> outward branch
    x = x.astype(np.float32).astype(np.float64)
    (g, terms), (o, _) = build_pair(gpu, d, cfg, x), build_pair(oracle(oracle_kind), d, cfg, x)
    for s in (g, o):
        s.set_context(core_scale=core, bond_scale=bond, semiaxes=semi, contact_distance=0.1)
    for name, t in terms.items():
        Fg, Fo = g.compute_force(1 << t), o.compute_force(1 << t)
        if np.abs(Fo).max() == 0:
            assert np.abs(Fg).max() == 0, name
            continue
        assert rel_force_err(Fg, Fo) < FORCE_TOL, name
    assert rel_force_err(g.compute_force(), o.compute_force()) < FORCE_TOL
    cg, co = g.get_context(), o.get_context()
    scale = max(abs(v) for v in co.axial_reaction)
    assert scale > 0
    for k in range(3):
        assert abs(cg.axial_reaction[k] - co.axial_reaction[k]) < 1e-5 * scale
    assert cg.reaction_force == pytest.approx(co.reaction_force, rel=1e-5)
    for name, t in terms.items():
        Eg, Eo = g.compute_energy(1 << t), o.compute_energy(1 << t)
        assert Eg == pytest.approx(Eo, rel=2e-5, abs=1e-6), name
    assert g.compute_energy() == pytest.approx(o.compute_energy(), rel=2e-5)


def test_mitotic_forces(oracle_kind):
    d = synth.make_diploid_design(scale=0.3)
    mcfg = synth.MitoticConfig(polar_ejection_force=0.7, polar_ejection_cross_section=0.04,
                               kfiber_length_anaphase=0.2, kfiber_length_prometaphase=0.3,
                               telophase_bond_spring_multiplier=0.5,
                               telophase_bending_energy_multiplier=2.0)
    rng = np.random.default_rng(2)
    md = synth.make_mitotic_design(d, mcfg)
    x = (synth.rod_positions(md, mcfg) + rng.normal(size=(md.n, 3)) * 0.05)
    x = x.astype(np.float32).astype(np.float64)
    g, o = gpu(md.n), O.OracleSystem(md.n, oracle_kind)
    for s in (g, o):
        t = stages.setup_anatelophase(s, md, mcfg)
        s.set_positions(x)
    for enter in (stages.enter_anaphase, stages.enter_telophase):
        enter(g, t), enter(o, t)
        assert rel_force_err(g.compute_force(), o.compute_force()) < FORCE_TOL
        assert g.compute_energy() == pytest.approx(o.compute_energy(), rel=2e-5)

    md = synth.make_mitotic_design(d, mcfg, sisters=True)
    x = (rng.normal(size=(md.n, 3)) * 1.5).astype(np.float32).astype(np.float64)
    g, o = gpu(md.n), O.OracleSystem(md.n, oracle_kind)
    for s in (g, o):
        t = stages.setup_prometaphase(s, md, mcfg)
        s.set_positions(x)
    for name, tid in t.items():
        Fo = o.compute_force(1 << tid)
        assert rel_force_err(g.compute_force(1 << tid), Fo) < FORCE_TOL, name
        assert g.compute_energy(1 << tid) == pytest.approx(o.compute_energy(1 << tid), rel=2e-5,
                                                           abs=1e-6), name


def test_full_size_diploid_cell(oracle_kind):
    """BASELINE.json configs[0] at its FULL size (46 chains, 60,672 beads): pair sets and contact
    triples bit-exact, forces / energy / wall statistics within 1e-5 against the oracle, plus
    size-independent properties (Newton's third law for the pair and bonded terms, bit
    reproducibility, identical replicas from identical seeds)."""
    cfg = synth.InterphaseConfig()
    d = synth.make_diploid_design(scale=1.0, cfg=cfg)
    assert d.n > 60000 and len(d.chains) == 46
    x = synth.globule_positions(d, seed=1, radius=1.5).astype(np.float32).astype(np.float64)
    g = gpu(d.n, 2)
    o = O.OracleSystem(d.n, oracle_kind)
    for s in (g, o):
        terms = stages.setup_interphase(s, d, cfg)
        s.set_positions(x)
        s.set_context(core_scale=0.5, bond_scale=0.5, semiaxes=(1.6, 1.5, 1.55),
                      contact_distance=cfg.contactmap_distance * 0.5)
    # neighbour search: list radius and contact distance
    for radius in (0.15, 0.12):
        a, b = g.neighbor_pairs(radius), o.neighbor_pairs(radius)
        assert len(b) > 5 * d.n
        np.testing.assert_array_equal(a, b)
    # forces per term and in total, energy, wall statistics
    for name, t in terms.items():
        Fg, Fo = g.compute_force(1 << t), o.compute_force(1 << t)
        if np.abs(Fo).max() == 0:
            assert np.abs(Fg).max() == 0, name
            continue
        assert rel_force_err(Fg, Fo) < FORCE_TOL, name
        if name in ("repulsion", "chain", "loops", "nor_bonds", "droplet"):
            # internal forces cancel: |sum F| << sum |F|
            assert np.abs(Fg.sum(axis=0)).max() < 2e-6 * np.abs(Fg).sum(), name
    Fg, Fo = g.compute_force(), o.compute_force()
    assert rel_force_err(Fg, Fo) < FORCE_TOL
    assert g.compute_energy() == pytest.approx(o.compute_energy(), rel=2e-5)
    cg, co = g.get_context(), o.get_context()
    assert cg.reaction_force == pytest.approx(co.reaction_force, rel=1e-5)
    # one contact update at the contact distance in force == the pair set at that distance
    g.contacts_update()
    o.contacts_update()
    trip = g.contacts_fetch()
    np.testing.assert_array_equal(trip, o.contacts_fetch())
    np.testing.assert_array_equal(trip[:, :2], o.neighbor_pairs(cfg.contactmap_distance * 0.5))
    assert np.all(trip[:, 2] == 1)
    g.contacts_clear()
    # 60 steps (plain launches, then graph replays): both replicas carry the same seed and
    # must stay bit-identical; a second context reproduces them bit for bit
    stages.interphase_callback0(g, cfg)
    g.set_seeds([5, 5])
    g.run(60, cfg.timestep)
    x0, x1 = g.get_positions_f32(0), g.get_positions_f32(1)
    assert np.array_equal(x0, x1)
    assert np.isfinite(x0).all() and np.abs(x0 - x).max() > 1e-3
    g.close()
    h = gpu(d.n, 1)
    stages.setup_interphase(h, d, cfg)
    h.set_positions(x)
    h.set_context(core_scale=0.5, bond_scale=0.5, semiaxes=(1.6, 1.5, 1.55),
                  contact_distance=cfg.contactmap_distance * 0.5)
    stages.interphase_callback0(h, cfg)
    h.set_seeds([5])
    h.run(60, cfg.timestep)
    assert np.array_equal(h.get_positions_f32(0), x0)
    h.close()


# ---------------------------------------------------------------------------
# trajectories under injected noise
# ---------------------------------------------------------------------------
def test_interphase_per_step_parity_resynced(oracle_kind):
    """The north-star bar is PER-STEP parity under identical noise: every step
    starts from the oracle's own (FP32-rounded) state, so chaotic divergence
    cannot accumulate.  40 consecutive steps along the oracle trajectory."""
    d, cfg, x = small_interphase(scale=0.02, radius=0.6)
    steps = 40
    z = np.random.default_rng(11).normal(size=(steps, d.n, 3))
    (g, _), (o, _) = build_pair(gpu, d, cfg, x, contacts=False), \
        build_pair(oracle(oracle_kind), d, cfg, x, contacts=False)
    for s in (g, o):
        stages.interphase_callback0(s, cfg)
    worst = 0.0
    for k in range(steps):
        xo = o.get_positions().astype(np.float32).astype(np.float64)
        co = o.get_context()
        for s in (g, o):
            s.set_positions(xo)
        g.set_context(time=co.time, semiaxes=list(co.semiaxes), core_scale=co.core_scale,
                      bond_scale=co.bond_scale, contact_distance=0.0)
        # W_{n-1} is zero after a restart on both sides
        g.run(1, cfg.timestep, noise=z[k:k + 1])
        o.run(1, cfg.timestep, noise=z[k:k + 1], interphase_callback=True, step_offset=k)
        worst = max(worst, np.abs(g.get_positions() - o.get_positions()).max())
        np.testing.assert_allclose(list(g.get_context().semiaxes), list(o.get_context().semiaxes),
                                   rtol=1e-9)
    assert worst < 1e-6                                 # 20x inside the 1e-5 * 2.0 bar


def test_interphase_trajectory_injected_noise(oracle_kind):
    """50 fused steps without re-syncing: schedule + wall ODE + contact map on
    device.  (Longer windows diverge chaotically: the error doubles every ~8
    steps in this dense state.)"""
    d, cfg, x = small_interphase(scale=0.02, radius=0.6)
    steps = 50
    z = np.random.default_rng(7).normal(size=(steps, d.n, 3))
    (g, _), (o, _) = build_pair(gpu, d, cfg, x), build_pair(oracle(oracle_kind), d, cfg, x)
    for s in (g, o):
        stages.interphase_callback0(s, cfg)
    g.run(steps, cfg.timestep, noise=z)
    out = o.run(steps, cfg.timestep, noise=z, interphase_callback=True, want_ctx=True)
    assert np.abs(g.get_positions() - o.get_positions()).max() < POS_TOL
    cg, co = g.get_context(), o.get_context()
    assert cg.step == steps
    assert cg.time == pytest.approx(co.time, rel=1e-12)
    assert cg.core_scale == pytest.approx(co.core_scale, rel=1e-12)
    assert cg.bond_scale == pytest.approx(co.bond_scale, rel=1e-12)
    np.testing.assert_allclose(list(cg.semiaxes), list(co.semiaxes), rtol=1e-7)
    assert abs(co.semiaxes[0] - cfg.wall_semiaxes_init[0]) > 1e-4       # the wall did move
    # contact maps: same pairs except where a pair sits within FP32 noise of
    # the contact distance after 100 steps
    a = {(i, j): v for i, j, v in g.contacts_fetch().tolist()}
    b = {(i, j): v for i, j, v in o.contacts_fetch().tolist()}
    assert len(b) > 0
    diff = sum(abs(a.get(k, 0) - b.get(k, 0)) for k in set(a) | set(b))
    assert diff <= max(2, int(2e-3 * sum(b.values())))


def test_relaxation_adaptive_timestep(oracle_kind):
    d, cfg, x = small_interphase(scale=0.02, radius=0.6)
    steps = 30
    z = np.random.default_rng(8).normal(size=(steps, d.n, 3))
    # one step: dt is a pure function of the forces -> tight
    (g, _), (o, _) = build_pair(gpu, d, cfg, x), build_pair(oracle(oracle_kind), d, cfg, x)
    g.run(1, cfg.timestep, spacestep=cfg.relaxation_spacestep, noise=z[:1])
    out = o.run(1, cfg.timestep, spacestep=cfg.relaxation_spacestep, noise=z[:1], want_ctx=True)
    assert g.get_context().last_timestep == pytest.approx(out["ctx"][0, 6], rel=1e-5)
    assert out["ctx"][0, 6] < cfg.timestep
    # 30 steps: positions within tolerance; dt follows (it is set by the single
    # most-stressed bead, so it amplifies position differences)
    (g, _), (o, _) = build_pair(gpu, d, cfg, x), build_pair(oracle(oracle_kind), d, cfg, x)
    g.run(steps, cfg.timestep, spacestep=cfg.relaxation_spacestep, noise=z)
    out = o.run(steps, cfg.timestep, spacestep=cfg.relaxation_spacestep, noise=z, want_ctx=True)
    assert np.abs(g.get_positions() - o.get_positions()).max() < POS_TOL
    assert g.get_context().last_timestep == pytest.approx(out["ctx"][-1, 6], rel=2e-2)


def test_anatelophase_trajectory(oracle_kind):
    d = synth.make_diploid_design(scale=0.3)
    mcfg = synth.MitoticConfig()
    md = synth.make_mitotic_design(d, mcfg)
    x = synth.rod_positions(md, mcfg)
    g, o = gpu(md.n), O.OracleSystem(md.n, oracle_kind)
    for s in (g, o):
        t = stages.setup_anatelophase(s, md, mcfg)
        s.set_positions(x)
    rng = np.random.default_rng(3)
    for enter, steps in ((stages.enter_anaphase, 60), (stages.enter_telophase, 40)):
        enter(g, t), enter(o, t)
        z = rng.normal(size=(steps, md.n, 3))
        g.run(steps, mcfg.timestep, noise=z)
        o.run(steps, mcfg.timestep, noise=z)
        assert np.abs(g.get_positions() - o.get_positions()).max() < 1e-5 * 5.0


def test_contact_update_exact(oracle_kind):
    d, cfg, x = small_interphase()
    (g, _), (o, _) = build_pair(gpu, d, cfg, x), build_pair(oracle(oracle_kind), d, cfg, x)
    for s in (g, o):
        s.set_context(contact_distance=0.12)
        s.contacts_update()
        s.contacts_update()
    a, b = g.contacts_fetch(), o.contacts_fetch()
    assert len(b) > d.n and (b[:, 2] == 2).all()
    np.testing.assert_array_equal(a, b)
    g.contacts_clear()
    assert len(g.contacts_fetch()) == 0


# ---------------------------------------------------------------------------
# replicas, skin, determinism, Philox
# ---------------------------------------------------------------------------
def test_replicas_are_independent_and_consistent(oracle_kind):
    d, cfg, x = small_interphase(scale=0.02, radius=0.6)
    R, steps = 3, 20
    z = np.random.default_rng(9).normal(size=(R, steps, d.n, 3))
    g = gpu(d.n, R)
    stages.setup_interphase(g, d, cfg)
    g.set_positions(x)
    stages.interphase_callback0(g, cfg)
    g.run(steps, cfg.timestep, noise=z)
    for r in range(R):
        o, _ = build_pair(oracle(oracle_kind), d, cfg, x)
        stages.interphase_callback0(o, cfg)
        o.run(steps, cfg.timestep, noise=z[r], interphase_callback=True)
        assert np.abs(g.get_positions(r) - o.get_positions()).max() < POS_TOL


@pytest.mark.parametrize("skin", [0.1, 0.2, 0.5])
def test_skin_keeps_every_pair_and_the_trajectory(oracle_kind, skin):
    """With a skin the cells are reused across steps (rebuild only when
    dcut(t) + 2 max displacement exceeds the cell edge, the criterion of
    MD/forcefield/detail/neighbor_list.hpp:129-158).  The stale cells must
    still yield EXACTLY the pair set at the current cutoff, and the trajectory
    must stay within tolerance of the oracle."""
    d, cfg, x = small_interphase(scale=0.02, radius=0.6)
    steps = 30     # trajectories of this dense system diverge chaotically (x2 per ~8 steps)
    z = np.random.default_rng(10).normal(size=(steps, d.n, 3))
    g, _ = build_pair(gpu, d, cfg, x)
    o, _ = build_pair(oracle(oracle_kind), d, cfg, x)
    g.set_skin(skin)
    for s_ in (g, o):
        stages.interphase_callback0(s_, cfg)
    g.run(steps, cfg.timestep, noise=z)
    o.run(steps, cfg.timestep, noise=z, interphase_callback=True)
    c = g.get_context()
    if skin >= 0.2:
        assert c.rebuilds < steps + 1                  # cells were actually reused
    grid = g.debug_grid()
    rnow = cfg.a_core_diameter * c.core_scale
    assert rnow + 2 * grid["max_disp"] <= grid["rc_valid"] or grid["need_rebuild"] == 1
    if grid["need_rebuild"] == 0:
        stale = g.neighbor_pairs(-rnow)                # through the cells as they are
        fresh = g.neighbor_pairs(rnow)                 # re-binned
        assert len(fresh) > d.n
        np.testing.assert_array_equal(stale, fresh)
    assert np.abs(g.get_positions() - o.get_positions()).max() < POS_TOL


def test_run_is_deterministic_and_chunkable():
    d, cfg, x = small_interphase(scale=0.02, radius=0.6)
    out = []
    for chunks in ((40,), (40,), (10, 25, 5)):
        g, _ = build_pair(gpu, d, cfg, x)
        g.set_seeds([1234])
        stages.interphase_callback0(g, cfg)
        first = True
        for c in chunks:
            g.run(c, cfg.timestep, restart=first)
            first = False
        out.append((g.get_positions_f32(), g.contacts_fetch()))
    assert np.array_equal(out[0][0], out[1][0])          # bit-reproducible
    assert np.array_equal(out[0][0], out[2][0])          # chunking changes nothing
    assert np.array_equal(out[0][1], out[2][1])


def _with_env(name, value, fn):
    import os
    old = os.environ.get(name)
    if value is None:
        os.environ.pop(name, None)
    else:
        os.environ[name] = value
    try:
        return fn()
    finally:
        if old is None:
            os.environ.pop(name, None)
        else:
            os.environ[name] = old


@pytest.mark.parametrize("replicas", [1, 3])
def test_small_system_path_is_bit_identical_to_the_four_kernel_rebuild(replicas):
    """<= 4,096 beads: one block per replica rebuilds the cell list (k_rebuild_small) and, in
    fixed-dt runs, does so in the step epilogue (k_finalize_rebuild: 3 launches per step
    instead of 7).  Same sorted order => identical bits; GC_NO_SMALL=1 selects the
    four-kernel path at gc_create."""
    d = synth.make_diploid_design(scale=0.3)
    mcfg = synth.MitoticConfig()
    md = synth.make_mitotic_design(d, mcfg, sisters=True)
    x = (np.random.default_rng(5).normal(size=(md.n, 3)) * 1.5).astype(np.float32).astype(np.float64)

    def prometaphase():
        g = gpu(md.n, replicas)
        stages.setup_prometaphase(g, md, mcfg)
        g.set_positions(x)
        g.set_seeds(np.arange(7, 7 + replicas, dtype=np.uint64))
        f0 = g.compute_force()
        pairs = g.neighbor_pairs(0.3)
        launches0 = g.launch_count()
        g.run(300, mcfg.timestep)                 # plain launches, then graph replays
        per_step = (g.launch_count() - launches0) / 300
        g.run(37, mcfg.timestep, restart=False)   # ragged chunk
        out = (f0, pairs, np.stack([g.get_positions_f32(r) for r in range(replicas)]),
               g.compute_energy(), per_step)
        g.close()
        return out

    small = _with_env("GC_NO_SMALL", None, prometaphase)
    plain = _with_env("GC_NO_SMALL", "1", prometaphase)
    for a, b in zip(small[:4], plain[:4]):
        assert np.array_equal(a, b)
    # launches per step and replica group (the graph has one branch per replica here)
    assert small[4] / replicas < 3.3 and plain[4] > 2.2 * small[4]

    # interphase force set (wall ODE, schedule, contact map every 20 steps), adaptive and fixed dt
    di, cfg, xi = small_interphase(scale=0.02, radius=0.6)

    def interphase():
        g, _ = build_pair(gpu, di, cfg, xi)
        g.set_seeds([99])
        g.run(30, cfg.timestep, spacestep=cfg.relaxation_spacestep)
        stages.interphase_callback0(g, cfg)
        g.run(130, cfg.timestep)
        out = (g.get_positions_f32(), g.contacts_fetch(), np.array(g.get_context().semiaxes))
        g.close()
        return out

    assert di.n <= 4096
    small = _with_env("GC_NO_SMALL", None, interphase)
    plain = _with_env("GC_NO_SMALL", "1", interphase)
    for a, b in zip(small, plain):
        assert np.array_equal(a, b)


@pytest.mark.parametrize("scale,replicas,skin", [(0.02, 5, 0.0), (0.1, 3, 0.0), (0.1, 11, 0.0),
                                                 (0.05, 4, 0.25)])
def test_replica_groups_of_the_step_graph_change_nothing(scale, replicas, skin):
    """The step graph runs the replicas as parallel branches (one capture stream per group,
    kernel arguments windowed to the group's replicas).  GC_GROUPS=1 is the single-branch
    graph: positions, contact triples and wall state must be bit-identical."""
    d, cfg, x = small_interphase(scale=scale, radius=0.6 if scale < 0.05 else 1.0)

    def run():
        g = gpu(d.n, replicas)
        stages.setup_interphase(g, d, cfg)
        g.set_positions(x)
        g.set_skin(skin)                  # > 0: cells reused while displacements allow
        g.set_seeds(np.arange(11, 11 + replicas, dtype=np.uint64))
        stages.interphase_callback0(g, cfg)
        g.run(150, cfg.timestep)
        g.run(90, cfg.timestep, restart=False)
        out = []
        for r in range(replicas):
            out += [g.get_positions_f32(r), g.contacts_fetch(r), np.array(g.get_context(r).semiaxes)]
        g.close()
        return out

    grouped = _with_env("GC_GROUPS", None, run)
    single = _with_env("GC_GROUPS", "1", run)
    assert len(grouped[1]) > 0
    for a, b in zip(grouped, single):
        assert np.array_equal(a, b)
    # replicas differ from one another (different seeds), i.e. the windows are distinct
    assert not np.array_equal(grouped[0], grouped[3])


def test_philox_stream_is_standard_normal():
    n = 200000
    g = gpu(n)
    g.set_bead_attrs(np.ones(n))
    g.set_seeds([42])
    z1, z2 = g.debug_noise(1), g.debug_noise(2)
    assert not np.array_equal(z1, z2)
    z = np.concatenate([z1.ravel(), z2.ravel()]).astype(np.float64)
    assert abs(z.mean()) < 5e-3 and abs(z.std() - 1) < 5e-3
    assert abs((z ** 3).mean()) < 2e-2 and abs((z ** 4).mean() - 3) < 5e-2
    from scipy import stats
    assert stats.kstest(z[:100000], "norm").pvalue > 1e-3
    # components are uncorrelated
    c = np.corrcoef(z1.T.astype(np.float64))
    assert np.abs(c - np.eye(3)).max() < 1e-2


def test_brownian_free_diffusion_statistics():
    """Free particles: <dx^2> per axis = 2 kT mu t  (BAOAB-limit averaging of
    consecutive noises keeps the long-time diffusion constant)."""
    n, steps, dt, mu = 100000, 50, 1e-3, 0.7
    g = gpu(n)
    g.set_bead_attrs(np.full(n, mu))
    x0 = np.zeros((n, 3))
    g.set_positions(x0)
    g.set_seeds([7])
    g.run(steps, dt)
    x = g.get_positions()
    # x_n = W_n / 2 + sum_{k<n} W_k  =>  variance = (n - 3/4) * 2 kT mu dt
    want = 2 * mu * dt * (steps - 0.75)
    assert x.var(axis=0) == pytest.approx(want, rel=0.03)
    assert np.abs(x.mean(axis=0)).max() < 4 * np.sqrt(want / n)


def test_error_reporting():
    g = gpu(10)
    with pytest.raises(gc_b200.GcError):
        g.run(1, 1e-3)                          # attrs not set
    g.set_bead_attrs(np.ones(10))
    x = np.zeros((10, 3))
    x[3, 1] = np.nan
    g.set_positions(x)
    with pytest.raises(gc_b200.GcError) as e:
        g.run(1, 1e-3)
    assert e.value.code == -4
    with pytest.raises(gc_b200.GcError):
        g.add_bonds([[0, 99]])


# ---------------------------------------------------------------------------
# trajectory output path
# ---------------------------------------------------------------------------
def test_quantised_frames_match_the_store_rounding():
    """gc_get_frame = simulation_store::save_positions' rounding to 16 significant bits
    (SIM/common/simulation_store.cpp:22-32,192-216, restated in synth.quantize_f32)."""
    rng = np.random.default_rng(5)
    n, R = 1000, 3
    g = gpu(n, R)
    g.set_bead_attrs(np.ones(n))
    xs = []
    for r in range(R):
        x = (rng.normal(size=(n, 3)) * 10.0 ** rng.integers(-6, 4, size=(n, 1))).astype(np.float32)
        x[0] = [0.0, -0.0, 1.0]
        x[1] = [np.float32(1.9999999), np.float32(65535.5 / 65536), -3.0e-30]   # carry, tie, tiny
        xs.append(x)
        g.set_positions(x.astype(np.float64), r)
    for bits in (16, 8, 23):
        all_frames = g.get_frame(-1, bits)
        assert all_frames.shape == (R, n, 3) and all_frames.dtype == np.float32
        for r in range(R):
            want = synth.quantize_f32(xs[r].astype(np.float64), bits)
            np.testing.assert_array_equal(all_frames[r], want)
            np.testing.assert_array_equal(g.get_frame(r, bits), want)
    g.close()


# ---------------------------------------------------------------------------
# parity at BENCHMARK size (SURVEY 8d "correctness run"; VERDICT r1 "parity gaps")
# ---------------------------------------------------------------------------
def _relaxed_full_cell():
    """The committed t = 0 interphase state bench.py starts from (60,672 beads)."""
    from pathlib import Path
    z = np.load(Path(__file__).resolve().parent / "golden" / "interphase_relaxed_60672.npz")
    cfg = synth.InterphaseConfig()
    d = synth.make_diploid_design(scale=1.0, cfg=cfg)
    x = z["positions"].astype(np.float32).astype(np.float64)
    assert x.shape == (d.n, 3)
    return d, cfg, x


def test_full_size_per_step_parity_resynced(oracle_kind):
    """20 consecutive interphase steps of the 60,672-bead benchmark cell under identical
    injected noise, every step restarted from the oracle's own FP32-rounded state: positions,
    wall semiaxes after every step, and the contact triples of the update at step 20."""
    d, cfg, x = _relaxed_full_cell()
    steps = 20
    rng = np.random.default_rng(2025)
    g, o = gpu(d.n), O.OracleSystem(d.n, oracle_kind)
    for s in (g, o):
        stages.setup_interphase(s, d, cfg)
        s.set_positions(x)
        stages.interphase_callback0(s, cfg)
    worst = 0.0
    for k in range(steps):
        z = rng.normal(size=(1, d.n, 3))
        xo = o.get_positions().astype(np.float32).astype(np.float64)
        co = o.get_context()
        for s in (g, o):
            s.set_positions(xo)
        g.set_context(time=co.time, semiaxes=list(co.semiaxes), core_scale=co.core_scale,
                      bond_scale=co.bond_scale, contact_distance=co.contact_distance)
        g.contacts_clear()
        o.contacts_clear()
        g.run(1, cfg.timestep, noise=z)                  # restart: W_{n-1} = 0 on both sides
        o.run(1, cfg.timestep, noise=z, interphase_callback=True, step_offset=k)
        worst = max(worst, np.abs(g.get_positions() - o.get_positions()).max())
        np.testing.assert_allclose(list(g.get_context().semiaxes), list(o.get_context().semiaxes),
                                   rtol=1e-9)
    assert worst < 2e-6, worst                           # bar: 1e-5 x nucleus size 2.0
    # step_offset + 1 = 20 is a contact-map step on the oracle; the device contact update is
    # the same callback applied to the same FP32 state
    trip_o = o.contacts_fetch()
    assert len(trip_o) > d.n
    g.set_positions(o.get_positions().astype(np.float32).astype(np.float64))
    o.set_positions(o.get_positions().astype(np.float32).astype(np.float64))
    co = o.get_context()
    g.set_context(time=co.time, semiaxes=list(co.semiaxes), core_scale=co.core_scale,
                  bond_scale=co.bond_scale, contact_distance=co.contact_distance)
    g.contacts_clear()
    o.contacts_clear()
    g.contacts_update()
    o.contacts_update()
    np.testing.assert_array_equal(g.contacts_fetch(), o.contacts_fetch())


def test_prometaphase_trajectory(oracle_kind):
    """60 steps of the prometaphase stage (sister chromatids, both k-fibre force fields,
    polar ejection with f != 0) under identical injected noise."""
    d = synth.make_diploid_design(scale=0.3)
    mcfg = synth.MitoticConfig(polar_ejection_force=0.7, polar_ejection_cross_section=0.04,
                               kfiber_length_prometaphase=0.3)
    md = synth.make_mitotic_design(d, mcfg, sisters=True)
    rng = np.random.default_rng(12)
    x = (rng.normal(size=(md.n, 3)) * 1.2).astype(np.float32).astype(np.float64)
    g, o = gpu(md.n), O.OracleSystem(md.n, oracle_kind)
    for s in (g, o):
        t = stages.setup_prometaphase(s, md, mcfg)
        s.set_positions(x)
    for name in ("kfiber_target", "kfiber_sister", "ejection_target", "ejection_sister"):
        assert np.abs(o.compute_force(1 << t[name])).max() > 0, name     # the terms are live
    steps = 60
    z = rng.normal(size=(steps, md.n, 3))
    g.run(steps, mcfg.timestep, noise=z)
    o.run(steps, mcfg.timestep, noise=z)
    assert np.abs(g.get_positions() - o.get_positions()).max() < 1e-5 * 5.0
    assert g.compute_energy() == pytest.approx(o.compute_energy(), rel=1e-4)


def test_force_only_100k_polymer(oracle_kind):
    """BASELINE configs[3]: nonbonded + bonded + wall evaluation of a 100k-bead polymer in an
    ellipsoid (the state bench.py's force_100k block times): total force <= 1e-5, pair set at
    the cutoff bit-exact."""
    semi = (5.0, 4.0, 3.0)
    cfg = synth.InterphaseConfig(wall_semiaxes_init=semi, nucleolus_droplet_energy=0.0)
    d, x = synth.make_polymer(100000, seed=4, semiaxes=semi)
    x = x.astype(np.float32).astype(np.float64)
    g, o = gpu(d.n), O.OracleSystem(d.n, oracle_kind)
    for s in (g, o):
        stages.setup_interphase(s, d, cfg, wall_ode=False, schedule=False, contacts=False)
        s.set_context(core_scale=1.0, bond_scale=1.0)
        s.set_positions(x)
    Fg, Fo = g.compute_force(), o.compute_force()
    assert rel_force_err(Fg, Fo) < FORCE_TOL
    assert g.compute_energy() == pytest.approx(o.compute_energy(), rel=2e-5)
    dcut = max(cfg.a_core_diameter, cfg.b_core_diameter)
    a, b = g.neighbor_pairs(dcut), o.neighbor_pairs(dcut)
    assert len(b) > 10 * d.n
    np.testing.assert_array_equal(a, b)


def test_batched_host_transfers_and_async_frames():
    """gc_set/get_positions_all_f32 (pageable and page-locked buffers), gc_get_contexts,
    gc_compute_energy_all and the asynchronous frame == their per-replica counterparts."""
    d, cfg, x = small_interphase(scale=0.02, radius=0.6)
    R = 3
    g = gpu(d.n, R)
    stages.setup_interphase(g, d, cfg)
    rng = np.random.default_rng(4)
    xs = (x[None] + rng.normal(size=(R, d.n, 3)) * 1e-3).astype(np.float32)
    g.set_positions_all_f32(xs)                               # pageable source
    for r in range(R):
        np.testing.assert_array_equal(g.get_positions_f32(r), xs[r])
    pinned = g.host_buffer()
    pinned[:] = xs[::-1]
    g.set_positions_all_f32(pinned)                           # page-locked source, in place
    out = g.get_positions_all_f32()
    np.testing.assert_array_equal(out, xs[::-1])
    np.testing.assert_array_equal(g.get_positions_all_f32(g.host_buffer()), xs[::-1])
    stages.begin_interphase(g, cfg)
    g.run(40, cfg.timestep)
    ctxs = g.get_contexts()
    e_all = g.compute_energy_all()
    for r in range(R):
        c = g.get_context(r)
        assert ctxs[r].step == c.step == 40 and ctxs[r].time == c.time
        assert list(ctxs[r].semiaxes) == list(c.semiaxes)
        assert e_all[r] == g.compute_energy(replica=r)
    # asynchronous frame: started, more steps queued behind it, then collected
    want = g.get_frame(-1)
    g.frame_begin(-1)
    with pytest.raises(gc_b200.api.GcError):
        g.frame_begin(-1)                                     # one frame in flight
    g.run(100, cfg.timestep, restart=False)
    np.testing.assert_array_equal(g.frame_end(), want)
    assert not np.array_equal(g.get_frame(-1), want)


def test_set_step_resumes_a_run_bit_for_bit():
    """A saved (positions, context, step) is a mid-stage checkpoint: resuming it in a fresh
    context continues the trajectory bit for bit (Philox counter = step, epoch = run)."""
    d, cfg, x = small_interphase(scale=0.02, radius=0.6)
    g = gpu(d.n)
    stages.setup_interphase(g, d, cfg)
    g.set_positions(x)
    g.set_seeds([77])
    stages.begin_interphase(g, cfg)
    g.run(80, cfg.timestep)
    xa, ca = g.get_positions_f32(0), g.get_context(0)
    g.run(80, cfg.timestep, restart=False)
    xb = g.get_positions_f32(0)
    h = gpu(d.n)
    stages.setup_interphase(h, d, cfg)
    h.set_seeds([77])
    h.set_positions_f32(xa)
    h.run(0, cfg.timestep)                                    # restart: epoch 1, as in g
    h.set_context(time=ca.time, semiaxes=list(ca.semiaxes), core_scale=ca.core_scale,
                  bond_scale=ca.bond_scale, contact_distance=ca.contact_distance)
    h.set_step(80)
    h.run(80, cfg.timestep, restart=False)
    np.testing.assert_array_equal(h.get_positions_f32(0), xb)
    assert h.get_context(0).step == 160



# ---------------------------------------------------------------------------
# edge cases (shapes of MDT/forcefield/test_neighbor_pairwise_forcefield.cc:37-94 and
# MDT/forcefield/test_bonded_*: empty and tiny systems, ragged chains, crowded cells)
# ---------------------------------------------------------------------------
def _plain_system(make, n, chains, x, eps=2.0, sigma=0.3):
    s = make(n)
    s.set_bead_attrs(np.full(n, 0.1))
    s.add_pair_softcore(eps, sigma)
    bonds = stages.chain_bonds(chains)
    if len(bonds):
        s.add_bonds(bonds, gc_b200.api.GC_BOND_SEMISPRING, K_a=1000.0, l_a=0.3)
    tri = stages.chain_triples(chains)
    if len(tri):
        s.add_triples(tri, 1.0)
    s.set_positions(x)
    return s


@pytest.mark.parametrize("case", ["one_bead", "two_coincident", "line_at_cutoff", "ragged_chains",
                                  "crowded_cell"])
def test_edge_case_forces_and_steps(oracle_kind, case):
    rng = np.random.default_rng(11)
    f32 = lambda a: np.asarray(a, dtype=np.float32).astype(np.float64)
    if case == "one_bead":                       # no pair, no bond: zero force, pure diffusion
        n, chains, x = 1, [(0, 1)], f32([[0.1, -0.2, 0.3]])
    elif case == "two_coincident":               # r = 0: the soft core's force factor is finite, F = 0
        n, chains, x = 2, [(0, 1), (1, 2)], f32([[0.5, 0.5, 0.5], [0.5, 0.5, 0.5]])
    elif case == "line_at_cutoff":               # neighbours exactly at sigma and one ulp either side
        s = np.float32(0.3)
        xs = np.cumsum([0.0, s, np.nextafter(s, np.float32(1)), np.nextafter(s, np.float32(0)), 0.5 * s])
        n, chains, x = 5, [(0, 5)], f32(np.stack([xs, np.zeros(5), np.zeros(5)], axis=1))
    elif case == "ragged_chains":                # chains of 1, 2, 3 and 40 beads: bonds / triples per length
        chains = [(0, 1), (1, 3), (3, 6), (6, 46)]
        n = 46
        x = f32(rng.normal(size=(n, 3)) * 0.4)
    else:                                        # 700 beads inside ONE cell: 44 home groups, tile overflow passes
        n, chains = 700, [(0, 700)]
        x = f32(rng.random((n, 3)) * 0.28)
    g, o = _plain_system(gpu, n, chains, x), _plain_system(oracle(oracle_kind), n, chains, x)
    Fg, Fo = g.compute_force(), o.compute_force()
    assert np.all(np.isfinite(Fg))
    if np.abs(Fo).max() > 0:
        assert rel_force_err(Fg, Fo) <= 1e-5
    else:
        assert np.abs(Fg).max() == 0.0
    assert g.compute_energy() == pytest.approx(o.compute_energy(), rel=2e-5, abs=1e-9)
    if case == "two_coincident":
        assert np.abs(Fg).max() == 0.0
    # a few fused steps under identical injected noise (zero steps first: a no-op)
    x0 = g.get_positions()
    g.run(0, 1e-5)
    np.testing.assert_array_equal(g.get_positions(), x0)
    steps = 8
    noise = rng.normal(size=(1, steps, n, 3))
    g.run(steps, 1e-5, noise=noise)
    o.run(steps, 1e-5, noise=noise)
    scale = max(1.0, np.abs(o.get_positions()).max())
    assert np.abs(g.get_positions() - o.get_positions()).max() <= 1e-5 * scale
    g.close()

"""Host-side inter-stage transitions and the full-cycle runner."""
import numpy as np
import pytest

from gcpkg import gc_b200

synth, T = gc_b200.synth, gc_b200.transitions


def test_resample_chain_reproduces_cubics_exactly():
    """A not-a-knot spline through >= 4 knots reproduces any cubic, so resampling a
    cubic curve must land on the curve (also outside the first/last knot)."""
    n, m = 9, 57
    t = (0.5 + np.arange(n)) / n
    tn = (0.5 + np.arange(m)) / m
    cub = lambda u: np.stack([1 + 2 * u - u ** 2 + 3 * u ** 3, u ** 3 - u, 4 * u ** 2 - 2], axis=1)  # noqa: E731
    np.testing.assert_allclose(T.resample_chain(cub(t), m), cub(tn), atol=1e-12)
    line = T.resample_chain(cub(t[:2]), 5)
    assert np.allclose(np.diff(line, axis=0), np.diff(line, axis=0)[0])
    assert T.resample_chain(cub(t[:1]), 3).shape == (3, 3)


def test_transition_round_trip_shapes_and_centroids():
    d = synth.make_diploid_design(scale=0.3)
    mcfg = synth.MitoticConfig()
    ma, mp = synth.make_mitotic_design(d, mcfg), synth.make_mitotic_design(d, mcfg, sisters=True)
    rng = np.random.default_rng(0)
    x_telo = rng.normal(size=(ma.n, 3))
    xi = T.transition_interphase(ma, d, x_telo)
    assert xi.shape == (d.n, 3) and np.isfinite(xi).all()
    for nor, nuc in d.nucleolar_bonds:
        np.testing.assert_array_equal(xi[nuc], xi[nor])
    xp = T.transition_prometaphase(d, mp, mcfg, xi)
    assert xp.shape == (mp.n, 3)
    s0, e0 = d.chains[0]
    np.testing.assert_allclose(xp[mp.chains[mp.sisters[0, 0]][0]], xi[s0:s0 + 100].mean(axis=0))
    sep = xp[mp.chains[mp.sisters[0, 1]][0]] - xp[mp.chains[mp.sisters[0, 0]][0]]
    np.testing.assert_allclose(sep, [0.0, -mcfg.sister_separation, 0.0], atol=1e-12)
    xa = T.transition_cycle(mp, ma, mcfg, xp)
    np.testing.assert_allclose(xa[:3], xp[:3] - np.asarray(mcfg.spindle_axis))


def test_transitions_match_the_reference_code(oracle_kind):
    """transitions.py against the reference's own spline header (cxx-spline, compiled
    unmodified into oracle/_ref) and its restated centroid / store-rounding loops."""
    import oracle_py as O
    rng = np.random.default_rng(5)
    # three knots: the reference's not-a-knot system is singular (h0 - h1 = 0 on both closing
    # rows) and its spline returns NaN; real chromosomes have >= 4 coarse beads.  The product
    # falls back to the parabola through the three points there (as scipy does).
    three = rng.normal(size=(3, 3))
    assert np.isnan(O.resample_chain(three, 9, oracle_kind)).all()
    assert np.isfinite(T.resample_chain(three, 9)).all()
    for n, m in ((2, 7), (4, 11), (13, 1320), (25, 2489), (6, 5)):
        chain = rng.normal(size=(n, 3)).cumsum(axis=0)
        np.testing.assert_allclose(T.resample_chain(chain, m), O.resample_chain(chain, m, oracle_kind),
                                   rtol=0, atol=2e-11 * max(1.0, np.abs(chain).max()))
    # whole transition on the full diploid design with prepare's coarse lengths (floor(len / cg))
    d = synth.make_diploid_design(scale=1.0)
    mcfg = synth.MitoticConfig()
    ma, mp = synth.make_mitotic_design(d, mcfg), synth.make_mitotic_design(d, mcfg, sisters=True)
    x_telo = rng.normal(size=(ma.n, 3))
    xi = T.transition_interphase(ma, d, x_telo)
    assert min(me - ms for ms, me in ma.chains) >= 4
    for (ms, me), (s, e) in zip(ma.chains, d.chains):
        np.testing.assert_allclose(xi[s:e], O.resample_chain(x_telo[ms:me], e - s, oracle_kind), atol=1e-10)
    xp = T.transition_prometaphase(d, mp, mcfg, xi)
    axis = np.asarray(mcfg.spindle_axis, dtype=float)
    disp = -mcfg.sister_separation * axis / np.linalg.norm(axis)
    for c, (s, e) in enumerate(d.chains):
        ti, si = mp.sisters[c]
        (ts, te), (ss, se) = mp.chains[ti], mp.chains[si]
        if s + mcfg.coarse_graining * (te - ts) > d.n:
            continue                       # window past the array end: undefined in the reference
        tgt, sis = O.coarse_grain(xi, s, e - s, mcfg.coarse_graining, te - ts, disp, oracle_kind)
        np.testing.assert_allclose(xp[ts:te], tgt, atol=1e-13)
        np.testing.assert_allclose(xp[ss:se], sis, atol=1e-13)
    # the stored-frame rounding (host restatement used by the fixtures)
    v = np.concatenate([rng.normal(size=2000) * 3.0, [0.0, -0.0, 1.0, -1.5, 1e-30, 123456.789, 2.0 ** -20]])
    np.testing.assert_array_equal(synth.quantize_f32(v), O.quantize_f32(v, 16, oracle_kind))
    np.testing.assert_array_equal(synth.quantize_f32(v, bits=12), O.quantize_f32(v, 12, oracle_kind))


@pytest.mark.gpu
def test_full_cycle_runs_every_stage():
    d = synth.make_diploid_design(scale=0.05)
    rep = gc_b200.cycle.run_cell_cycle(design=d, scale_steps=2e-4, seed=3)
    names = [r["stage"] for r in rep["stages"]]
    assert names == ["anaphase", "telophase", "relaxation", "interphase", "prometaphase"]
    assert all(np.isfinite(r["energy_per_bead"]) for r in rep["stages"])
    assert rep["next_anaphase_start"].shape[0] == rep["beads_anaphase"]
    assert rep["interphase_end"]["core_scale"] > 0.5 and rep["frames_saved"] >= 0


@pytest.mark.gpu
@pytest.mark.parametrize("bits", [0, 16])
def test_device_transitions_match_the_reference_code(oracle_kind, bits):
    """gc_transition_* (FP64 kernels, positions never leave the device) against the
    reference's spline header and its centroid / displacement loops; bits = 16 adds the
    store's rounding on both sides of every transition, as the file hand-off does."""
    import oracle_py as O
    d = synth.make_diploid_design(scale=1.0)
    mcfg = synth.MitoticConfig()
    ma, mp = synth.make_mitotic_design(d, mcfg), synth.make_mitotic_design(d, mcfg, sisters=True)
    rng = np.random.default_rng(17)
    q = (lambda v: O.quantize_f32(v, bits, oracle_kind).astype(np.float64)) if bits else \
        (lambda v: np.asarray(v, dtype=np.float32).astype(np.float64))

    def close(have, want, what):
        # FP64 on both sides, FP32 storage: equal up to one unit of the stored precision
        tol = 2.0 ** -(bits if bits else 23) * np.maximum(np.abs(want), 1e-30) * 1.01
        bad = np.abs(have - want) > tol
        assert not bad.any(), (what, int(bad.sum()), float(np.abs(have - want).max()))
        if bits:                                            # and almost always bit-equal
            assert (have != want).mean() < 2e-3, (what, float((have != want).mean()))

    ga, gi = gc_b200.System(ma.n, 1, 0), gc_b200.System(d.n, 1, 0)
    x_telo = (rng.normal(size=(ma.n, 3)) * 0.8).astype(np.float32).astype(np.float64)
    ga.set_positions(x_telo)
    gi.transition_interphase_from(ga, ma.chains, d.chains, d.nucleolar_bonds, bits=bits)
    xi = gi.get_positions()
    src = q(x_telo)
    want = np.zeros((d.n, 3))
    for (ms, me), (s, e) in zip(ma.chains, d.chains):
        want[s:e] = O.resample_chain(src[ms:me], e - s, oracle_kind)
    for nor, nuc in d.nucleolar_bonds:
        want[nuc] = want[nor]
    close(xi, q(want), "interphase")

    gp = gc_b200.System(mp.n, 1, 0)
    axis = np.asarray(mcfg.spindle_axis, dtype=float)
    disp = -mcfg.sister_separation * axis / np.linalg.norm(axis)
    gp.transition_prometaphase_from(gi, d.chains, mp.chains[mp.sisters[:, 0]], mp.chains[mp.sisters[:, 1]],
                                    mcfg.coarse_graining, disp, bits=bits)
    xp = gp.get_positions()
    src = q(xi)
    want = np.zeros((mp.n, 3))
    for c, (s, e) in enumerate(d.chains):
        ti, si = mp.sisters[c]
        (ts, te), (ss, se) = mp.chains[ti], mp.chains[si]
        want[ts:te], want[ss:se] = O.coarse_grain(src, s, e - s, mcfg.coarse_graining, te - ts, disp, oracle_kind)
    close(xp, q(want), "prometaphase")

    gn = gc_b200.System(ma.n, 1, 0)
    gn.transition_cycle_from(gp, mp.chains[mp.sisters[:, 0]], ma.chains, -axis, bits=bits)
    xa = gn.get_positions()
    src = q(xp)
    want = np.zeros((ma.n, 3))
    for c, (s, e) in enumerate(ma.chains):
        ms, _ = mp.chains[mp.sisters[c, 0]]
        want[s:e] = src[ms:ms + (e - s)] - axis
    close(xa, q(want), "cycle")
    for g in (ga, gi, gp, gn):
        g.close()


@pytest.mark.gpu
def test_cycle_device_transitions_equal_host_transitions():
    """The cycle runner with gc_transition_* and with transitions.py (scipy spline, numpy
    centroids) starts every stage from the same stored frame up to one unit of the 16-bit
    rounding, so stage energies agree closely; a 3-knot chain uses the parabola on both."""
    d = synth.make_diploid_design(scale=0.05)
    a = gc_b200.cycle.run_cell_cycle(design=d, scale_steps=1e-4, seed=3)
    b = gc_b200.cycle.run_cell_cycle(design=d, scale_steps=1e-4, seed=3, host_transitions=True)
    assert np.isfinite(a["next_anaphase_start"]).all()
    # anaphase / telophase are identical runs; later stages start from frames that may differ
    # by one rounding unit in a few coordinates
    assert a["stages"][0]["energy_per_bead"] == b["stages"][0]["energy_per_bead"]
    for ra, rb in zip(a["stages"][2:], b["stages"][2:]):
        assert ra["energy_per_bead"] == pytest.approx(rb["energy_per_bead"], rel=5e-2), ra["stage"]

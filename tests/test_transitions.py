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


@pytest.mark.gpu
def test_full_cycle_runs_every_stage():
    d = synth.make_diploid_design(scale=0.05)
    rep = gc_b200.cycle.run_cell_cycle(design=d, scale_steps=2e-4, seed=3)
    names = [r["stage"] for r in rep["stages"]]
    assert names == ["anaphase", "telophase", "relaxation", "interphase", "prometaphase"]
    assert all(np.isfinite(r["energy_per_bead"]) for r in rep["stages"])
    assert rep["next_anaphase_start"].shape[0] == rep["beads_anaphase"]
    assert rep["interphase_end"]["core_scale"] > 0.5 and rep["frames_saved"] >= 0

"""Contact-map observables: the numpy oracle against vectors produced by
the reference's own analysis modules, and the device implementation against both."""
from pathlib import Path

import numpy as np
import pytest

import analysis_oracle as A
from gcpkg import gc_b200

GOLD = np.load(Path(__file__).resolve().parent / "golden" / "analysis_vectors.npz")
MODES = {"default": (False, False), "mask_intra": (True, False), "covariance": (False, True)}


def _close(have, want, rtol, atol=0.0):
    have, want = np.asarray(have, dtype=np.float64), np.asarray(want, dtype=np.float64)
    np.testing.assert_array_equal(np.isnan(have), np.isnan(want))
    ok = ~np.isnan(want)
    np.testing.assert_allclose(have[ok], want[ok], rtol=rtol, atol=atol)


@pytest.mark.parametrize("mode", list(MODES))
def test_oracle_matches_the_reference_vectors(mode):
    m, ranges = GOLD["matrix"], GOLD["chrom_ranges"]
    mi, cov = MODES[mode]
    intra, inter = A.expected_contacts(m, ranges)
    _close(intra, GOLD[f"{mode}_intra"], 1e-6)
    assert inter == pytest.approx(float(GOLD[f"{mode}_inter"]), rel=1e-7)
    _close(A.observed_over_expected(m, ranges, intra, inter, mi), GOLD[f"{mode}_oe"], 1e-6)
    r = A.compartment_pc1(m, ranges, mi, cov)
    assert r["steps"] == int(GOLD[f"{mode}_steps"])
    _close(r["ev1"], GOLD[f"{mode}_ev1"], 1e-4, 1e-6)
    _close(r["pc1"], GOLD[f"{mode}_pc1"], 1e-4, 1e-6)
    assert r["evr"] == pytest.approx(float(GOLD[f"{mode}_evr"]), rel=1e-4)
    _close(A.nci(m, ranges), GOLD["nci"], 1e-6)


# ---------------------------------------------------------------------------
# device implementation (gc_obs.cu)
# ---------------------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("mode", list(MODES))
def test_device_observables_match_the_reference_vectors(mode):
    m, ranges = GOLD["matrix"], GOLD["chrom_ranges"]
    mi, cov = MODES[mode]
    ob = gc_b200.Observables(m.shape[0], ranges)
    ob.set_matrix(m)
    intra, inter = ob.expected()
    _close(intra, GOLD[f"{mode}_intra"], 2e-6)
    assert inter == pytest.approx(float(GOLD[f"{mode}_inter"]), rel=1e-6)
    r = ob.pc1(mask_intra=mi, use_covariance=cov)
    assert abs(r["steps"] - int(GOLD[f"{mode}_steps"])) <= 1
    sign = np.sign(np.nansum(r["ev1"] * GOLD[f"{mode}_ev1"]))
    assert sign != 0
    _close(sign * r["ev1"], GOLD[f"{mode}_ev1"], 2e-3, 2e-4)     # iteration stops at |dv| < 1e-4
    _close(sign * r["pc1"], GOLD[f"{mode}_pc1"], 2e-3, 2e-4)
    assert r["evr"] == pytest.approx(float(GOLD[f"{mode}_evr"]), rel=2e-3)
    _close(ob.nci(), GOLD["nci"], 1e-6)
    ob.close()


@pytest.mark.gpu
def test_device_observables_from_a_simulated_contact_map():
    """Ensemble contact runs reduced on the device -> binned matrix, P(s), PC1, NCI and the
    A/B enrichment, against the numpy oracle on the fetched triples."""
    import torch
    synth, stages, ens = gc_b200.synth, gc_b200.stages, gc_b200.ensemble
    cfg = synth.InterphaseConfig()
    d = synth.make_diploid_design(scale=0.05, cfg=cfg)
    g = gc_b200.System(d.n, 4, 0)
    stages.setup_interphase(g, d, cfg)
    g.set_positions(synth.globule_positions(d, seed=21, radius=0.75))
    g.set_seeds([1, 2, 3, 4])
    stages.begin_interphase(g, cfg)
    g.run(400, cfg.timestep)
    keys, counts = g.contacts_reduce_device(-1, 0)
    ijv = ens.keys_to_ijv(keys, counts).copy()
    assert len(ijv) > d.n
    # bins of 5 beads inside every chain; nucleolar beads are dropped
    bin_size, bead_bin, bead_chain, ranges, nb = 5, np.full(d.n, -1, np.int32), np.full(d.n, -1, np.int32), [], 0
    for c, (s, e) in enumerate(d.chains):
        k = -(-(e - s) // bin_size)
        bead_bin[s:e] = nb + (np.arange(e - s) // bin_size)
        bead_chain[s:e] = c
        ranges.append((nb, nb + k))
        nb += k
    ob = gc_b200.Observables(nb, ranges)
    ob.set_bead_map(bead_bin, bead_chain)
    ob.add_runs(keys, counts)
    want = A.bin_contacts(ijv, bead_bin, nb)
    np.testing.assert_array_equal(ob.matrix(), want)
    ob.clear()
    ob.add_triples(ijv)                                   # the host path builds the same matrix
    np.testing.assert_array_equal(ob.matrix(), want)
    # P(s) at bead resolution == the numpy pass of ensemble.contact_probability (raw sums)
    max_s = 60
    hist = ob.ps_runs(keys, counts, max_s)
    i, j, v = ijv[:, 0].astype(int), ijv[:, 1].astype(int), ijv[:, 2].astype(np.int64)
    cis = (bead_chain[i] == bead_chain[j]) & (bead_chain[i] >= 0) & (j - i <= max_s)
    np.testing.assert_array_equal(hist, np.bincount((j - i)[cis], weights=v[cis], minlength=max_s + 1).astype(np.uint64))
    intra, inter = ob.expected()
    wi, we = A.expected_contacts(want, ranges)
    _close(intra, wi, 1e-5)
    assert inter == pytest.approx(we, rel=1e-5)
    r, w = ob.pc1(), A.compartment_pc1(want, ranges)
    sign = np.sign(np.nansum(r["ev1"] * w["ev1"]))
    _close(sign * r["pc1"], w["pc1"], 5e-3, 5e-4)
    assert r["evr"] == pytest.approx(w["evr"], rel=5e-3)
    _close(ob.nci(), A.nci(want, ranges), 1e-6)
    # A/B enrichment: same-type bins touch more often than cross-type bins
    frac_a = np.zeros(nb)
    np.add.at(frac_a, bead_bin[bead_bin >= 0], (d.a > d.b)[bead_bin >= 0])
    size = np.bincount(bead_bin[bead_bin >= 0], minlength=nb)
    btype = np.where(frac_a > 0.5 * size, 0, 1).astype(np.int32)
    en = ob.type_enrichment(btype, 2, min_separation=3)
    iu, ju = np.triu_indices(nb, 1)
    chrom_of = np.concatenate([np.full(e - s, c) for c, (s, e) in enumerate(ranges)])
    ok = (chrom_of[iu] != chrom_of[ju]) | (ju - iu >= 3)
    for a in range(2):
        for b in range(a, 2):
            sel = ok & (((btype[iu] == a) & (btype[ju] == b)) | ((btype[iu] == b) & (btype[ju] == a)))
            assert en[a, b] == pytest.approx(want[iu[sel], ju[sel]].astype(np.float64).mean(), rel=1e-9)
            assert en[b, a] == en[a, b]
    ob.close()
    g.close()

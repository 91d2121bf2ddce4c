"""CPU restatement of the reference's contact-map analysis.  TEST INFRASTRUCTURE ONLY (may be
imported from tests/ and bench.py's checker legs; never from the product package).

Restates, in numpy, the parts of the reference's Python analysis that the device observables
replace (citations relative to /root/reference):
  estimate_expected_contacts   src/pc1.py:156-183  (cis decay profile = P(s) on the binned map,
                               mean trans contact)
  O/E transform, column standardisation, PC1 / EV1 / explained variance   src/pc1.py:44-101
  power_svd                    src/pkg/pc1/math.py:13-37
  compute_nci                  src/pkg/nci/nci.py:9-35
Pinned: tests/test_observables.py checks every function against tests/golden/
analysis_vectors.npz, which scripts/make_analysis_golden.py produced by running the reference's
own modules.
"""
from __future__ import annotations

import numpy as np


def expected_contacts(matrix, chrom_ranges):
    """src/pc1.py:156-183 -> (intra[max_len], inter)."""
    m = np.asarray(matrix, dtype=np.float32)
    max_sep = max(int(e - s) for s, e in chrom_ranges)
    inter_sum, inter_count = 0.0, 0
    intra_sums = np.zeros(max_sep, dtype=m.dtype)
    intra_counts = np.zeros(max_sep, dtype=m.dtype)
    for s, e in chrom_ranges:
        rows = m[s:e]
        for block in (rows[:, :s], rows[:, e:]):
            ok = np.isfinite(block)
            inter_sum += block[ok].sum()
            inter_count += ok.sum()
        cis = rows[:, s:e]
        for k in range(len(cis)):
            d = np.diag(cis, k)
            ok = np.isfinite(d)
            intra_sums[k] += d[ok].sum()
            intra_counts[k] += ok.sum()
    with np.errstate(invalid="ignore", divide="ignore"):
        return intra_sums / intra_counts, float(inter_sum / inter_count)


def observed_over_expected(matrix, chrom_ranges, intra, inter, mask_intra=False):
    """src/pc1.py:44-58, returned as a new array."""
    data = np.array(matrix, dtype=np.float32, order="C")
    with np.errstate(divide="ignore", invalid="ignore"):
        for s, e in chrom_ranges:
            rows = data[s:e]
            rows[:, :s] /= inter
            rows[:, e:] /= inter
            cis = rows[:, s:e]
            if mask_intra:
                cis[np.isfinite(cis)] = 1
            else:
                n = len(cis)
                i, j = np.meshgrid(np.arange(n), np.arange(n), indexing="ij")
                cis[...] = cis / intra[np.abs(i - j)].astype(np.float32)
    return data


def power_svd(data, tol=1e-4, max_steps=500):
    """src/pkg/pc1/math.py:13-37 -> (vector, steps)."""
    dim = data.shape[1]
    prev = np.ones(dim, dtype=data.dtype) / np.sqrt(dim)
    steps = 0
    while True:
        w = data @ prev
        vec = np.nansum(w[:, None] * data, axis=0)
        vec /= np.linalg.norm(vec)
        steps += 1
        delta = np.abs(vec - prev).max()
        prev = vec
        if delta < tol or steps >= max_steps:
            return vec, steps


def compartment_pc1(matrix, chrom_ranges, mask_intra=False, use_covariance=False, tol=1e-4):
    """src/pc1.py:36-101 -> dict(intra, inter, ev1, pc1, evr, steps)."""
    intra, inter = expected_contacts(matrix, chrom_ranges)
    data = observed_over_expected(matrix, chrom_ranges, intra, inter, mask_intra)
    coverage = np.nansum(data, axis=0)
    sel = coverage > 0
    data = data[:, sel]
    with np.errstate(divide="ignore", invalid="ignore"):
        data -= np.nanmean(data, axis=0)
        if not use_covariance:
            data /= np.nanstd(data, axis=0)
    vec, steps = power_svd(data, tol)
    pc1 = data @ vec
    ev1 = np.full(len(sel), np.nan, dtype=vec.dtype)
    ev1[sel] = vec
    var = np.nansum(np.nanvar(data, axis=0))
    pc1 = pc1 / np.sqrt(var)
    return dict(intra=intra, inter=inter, ev1=ev1, pc1=pc1, evr=float(np.nanvar(pc1)), steps=steps)


def nci(matrix, chrom_ranges):
    """src/pkg/nci/nci.py:9-35 per chromosome, NaN appended at every chromosome's last bin."""
    m = np.asarray(matrix, dtype=np.float32)
    out = []
    with np.errstate(divide="ignore", invalid="ignore"):
        for s, e in chrom_ranges:
            d = np.diag(m)[s:e]
            sub = np.diag(m, 1)[s:e - 1]
            out.append(np.append(sub / np.sqrt(d[1:] * d[:-1]), np.nan))
    return np.concatenate(out)


def bin_contacts(ijv, bead_bin, n_bins):
    """Symmetric dense bin matrix from (i, j, count) triples, as cooler's matrix() of the pixels
    src/cool.py:80-110 ingests (upper-triangle pixels mirrored, the diagonal counted once)."""
    m = np.zeros((n_bins, n_bins), dtype=np.float64)
    bi, bj = bead_bin[ijv[:, 0].astype(np.int64)], bead_bin[ijv[:, 1].astype(np.int64)]
    ok = (bi >= 0) & (bj >= 0)
    bi, bj, v = bi[ok], bj[ok], ijv[ok, 2].astype(np.float64)
    np.add.at(m, (bi, bj), v)
    off = bi != bj
    np.add.at(m, (bj[off], bi[off]), v[off])
    return m.astype(np.float32)

"""Generates tests/golden/analysis_vectors.npz by running the REFERENCE's own analysis code
(src/pc1.py, src/pkg/pc1/math.py, src/pkg/nci/nci.py under /root/reference) on a small seeded
contact matrix.  Those modules import `cooler` and `tqdm` at the top, which are not installed
here; only functions that never touch them are called, so the two names are satisfied by empty
stub modules.  Runs in the build container only (needs /root/reference); the output is
committed so that the tests travel.

    python scripts/make_analysis_golden.py
"""
import sys
import types
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
REF = Path("/root/reference/src")
for name in ("cooler", "tqdm"):
    sys.modules.setdefault(name, types.ModuleType(name))
sys.modules["cooler"].Cooler = type("Cooler", (), {})       # only named in a type annotation
sys.path.insert(0, str(REF))
import pc1 as ref_pc1                      # noqa: E402  (the reference module, unmodified)
from pkg.nci.nci import compute_nci      # noqa: E402
from pkg.pc1.math import power_svd        # noqa: E402


def synthetic_matrix(seed=7):
    """Symmetric integer contact matrix with a P(s) decay, a checkerboard compartment signal,
    five chromosomes of different length and one empty bin (zero coverage)."""
    rng = np.random.default_rng(seed)
    lengths = [70, 55, 40, 64, 31]
    n = sum(lengths)
    ranges, pos = {}, 0
    for k, L in enumerate(lengths):
        ranges[f"chr{k + 1}"] = (pos, pos + L)
        pos += L
    comp = np.sign(np.sin(np.arange(n) * 0.37 + 0.3))
    chrom = np.concatenate([np.full(L, k) for k, L in enumerate(lengths)])
    i, j = np.meshgrid(np.arange(n), np.arange(n), indexing="ij")
    sep = np.abs(i - j)
    lam = np.where(chrom[i] == chrom[j], 40.0 / (1.0 + sep) ** 0.9, 0.6)
    lam = lam * (1.0 + 0.5 * comp[i] * comp[j])
    m = rng.poisson(lam).astype(np.float64)
    m = np.triu(m) + np.triu(m, 1).T
    dead = 97
    m[dead, :] = 0
    m[:, dead] = 0
    return m.astype(np.float32), ranges


def run_reference(matrix, ranges, mask_intra, use_covariance, tol=1e-4):
    """The body of pc1.main between loading the matrix and writing the table, verbatim calls."""
    contact = np.array(matrix, dtype=np.float32, order="C")
    expected = ref_pc1.estimate_expected_contacts(contact, ranges)
    data = contact
    with np.errstate(divide="ignore", invalid="ignore"):
        for patch in ref_pc1.scan_chrom_rows(data, ranges):
            patch.trans_1[...] /= expected.inter
            patch.trans_2[...] /= expected.inter
            if mask_intra:
                patch.cis[np.isfinite(patch.cis)] = 1
            else:
                size = len(patch.cis)
                for k in range(-size + 1, size):
                    diag = np.diagonal(patch.cis, k)
                    diag.setflags(write=True)
                    diag[:] /= expected.intra[abs(k)]
    oe = data.copy()
    coverages = np.nansum(data, axis=0)
    selection = coverages > 0
    data = data[:, selection]
    with np.errstate(divide="ignore", invalid="ignore"):
        data[...] -= np.nanmean(data, axis=0)
        if not use_covariance:
            data[...] /= np.nanstd(data, axis=0)
    steps = 0
    for svd in power_svd(data):
        steps += 1
        if svd.delta < tol or steps >= 500:
            break
    pc = data @ svd.vector
    ev1 = ref_pc1.unselect_vector(svd.vector, selection)
    data_var = np.nansum(np.nanvar(data, axis=0))
    pc[:] /= np.sqrt(data_var)
    evr = np.nanvar(pc)
    return dict(intra=expected.intra, inter=expected.inter, oe=oe, ev1=ev1, pc1=pc, evr=float(evr),
                steps=steps, selection=selection)


m, ranges = synthetic_matrix()
out = {"matrix": m, "chrom_ranges": np.asarray(list(ranges.values()), dtype=np.int32)}
for tag, (mi, cov) in {"default": (False, False), "mask_intra": (True, False), "covariance": (False, True)}.items():
    r = run_reference(m, ranges, mi, cov)
    for k, v in r.items():
        out[f"{tag}_{k}"] = np.asarray(v)
    print(tag, "steps", r["steps"], "evr", r["evr"], "inter", r["inter"])
with np.errstate(divide="ignore", invalid="ignore"):
    out["nci"] = np.concatenate([np.append(compute_nci(m, s, e, chunk_size=16), np.nan) for s, e in ranges.values()])
dst = ROOT / "tests" / "golden" / "analysis_vectors.npz"
np.savez_compressed(dst, **out)
print("->", dst, dst.stat().st_size, "bytes")

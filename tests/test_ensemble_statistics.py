"""Ensemble observables, CUDA path vs the reference with INDEPENDENT noise streams
(BASELINE.json north_star: "ensemble observables (radius of gyration, contact-probability
P(s), compartment contact maps) within stated statistical tolerance").

The device draws Philox + Box-Muller normals, the reference sfc64 + ziggurat, so no
trajectory can agree; what must agree is every ensemble mean.  Both sides run the whole
interphase driver loop (expansion schedule, wall ODE, contact cadence) on a small
diploid-like cell, R replicas each, and every observable's replica mean is compared with

    |mean_gpu - mean_ref| <= Z * sqrt(se_gpu^2 + se_ref^2) + REL * |mean_ref|,   Z = 4.5

(se = standard error over replicas; REL = 0.5 % covers the FP32 / fast-math bias).
Seeds are fixed, so the test is deterministic.
"""
import threading

import numpy as np
import pytest

import oracle_py as O
from gcpkg import gc_b200
from helpers import stages, synth

pytestmark = pytest.mark.gpu

R = 24
CHUNK, BURN_CHUNKS, CHUNKS = 2000, 3, 10
Z, REL = 4.5, 5e-3


def _cell():
    cfg = synth.InterphaseConfig(timestep=1e-4, wall_semiaxes_init=(0.5, 0.5, 0.5),
                                 wall_semiaxes_spring=(2000.0, 2000.0, 2000.0), wall_mobility=3e-2)
    d = synth.make_diploid_design(scale=0.005, cfg=cfg)
    x = synth.globule_positions(d, seed=11, radius=0.35)
    return d, cfg, x


def _rg2(x, chains):
    return np.array([((x[s:e] - x[s:e].mean(axis=0)) ** 2).sum(axis=1).mean() for s, e in chains])


def _observables(d, cfg, frames, contacts, n_updates, semiaxes):
    """One replica -> vector of observables."""
    rg2 = np.mean([_rg2(x, d.chains).mean() for x in frames])
    chain_of = np.full(d.n, -1)
    for c, (s, e) in enumerate(d.chains):
        chain_of[s:e] = c
    i, j, v = contacts[:, 0].astype(int), contacts[:, 1].astype(int), contacts[:, 2].astype(float)
    cis = (chain_of[i] == chain_of[j]) & (chain_of[i] >= 0)
    ps = []
    for s in (1, 2, 3, 4):
        npairs = sum(max(e - b - s, 0) for b, e in d.chains)
        ps.append(v[cis & (j - i == s)].sum() / (npairs * n_updates))
    ta, tb = d.a > d.b, d.b > d.a
    trans = ~cis
    aa = v[trans & ta[i] & ta[j]].sum() / n_updates
    bb = v[trans & tb[i] & tb[j]].sum() / n_updates
    ab = v[trans & ((ta[i] & tb[j]) | (tb[i] & ta[j]))].sum() / n_updates
    return np.array([rg2, *ps, aa, bb, ab, v.sum() / n_updates, np.mean(semiaxes)])


NAMES = ["Rg2", "P(1)", "P(2)", "P(3)", "P(4)", "trans AA", "trans BB", "trans AB", "contacts/frame",
         "semiaxis"]


def _run_gpu(d, cfg, x):
    g = gc_b200.System(d.n, R, 0)
    stages.setup_interphase(g, d, cfg)
    g.set_positions(x)
    g.set_seeds([5000 + r for r in range(R)])
    stages.begin_interphase(g, cfg)
    frames = [[] for _ in range(R)]
    for k in range(CHUNKS):
        g.run(CHUNK, cfg.timestep, restart=(k == 0))
        if k + 1 == BURN_CHUNKS:
            g.contacts_clear()
        if k + 1 > BURN_CHUNKS:
            for r in range(R):
                frames[r].append(g.get_positions(r))
    n_updates = (CHUNKS - BURN_CHUNKS) * CHUNK // cfg.contactmap_update_interval
    out = [_observables(d, cfg, frames[r], g.contacts_fetch(r), n_updates, g.get_context(r).semiaxes)
           for r in range(R)]
    g.close()
    return np.array(out)


def _run_ref(d, cfg, x, kind):
    out = [None] * R
    systems = [O.OracleSystem(d.n, kind) for _ in range(R)]   # (library loading is not thread-safe)

    def work(r):
        o = systems[r]
        stages.setup_interphase(o, d, cfg)
        o.set_positions(x)
        stages.interphase_callback0(o, cfg)
        frames = []
        for k in range(CHUNKS):
            # every call re-seeds the reference's generator: one seed per (replica, chunk)
            o.run(CHUNK, cfg.timestep, interphase_callback=True, step_offset=k * CHUNK, stepper=0,
                  seed=7000 + 100 * r + k)
            if k + 1 == BURN_CHUNKS:
                o.contacts_clear()
            if k + 1 > BURN_CHUNKS:
                frames.append(o.get_positions())
        n_updates = (CHUNKS - BURN_CHUNKS) * CHUNK // cfg.contactmap_update_interval
        out[r] = _observables(d, cfg, frames, o.contacts_fetch(), n_updates, o.get_context().semiaxes)
        o.close()

    threads = [threading.Thread(target=work, args=(r,)) for r in range(R)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    return np.array(out)


def test_ensemble_observables_match_reference(oracle_kind):
    d, cfg, x = _cell()
    G, F = _run_gpu(d, cfg, x), _run_ref(d, cfg, x, oracle_kind)
    mg, mf = G.mean(axis=0), F.mean(axis=0)
    se = np.sqrt(G.var(axis=0, ddof=1) / R + F.var(axis=0, ddof=1) / R)
    report = "\n".join(f"{n:15s} gpu {a:.5g}  ref {b:.5g}  diff {a - b:+.3g}  allowed {Z * s + REL * abs(b):.3g}"
                       for n, a, b, s in zip(NAMES, mg, mf, se))
    print(report)
    assert (mf[:5] > 0).all() and mf[8] > 0, "degenerate observables:\n" + report
    bad = np.abs(mg - mf) > Z * se + REL * np.abs(mf)
    assert not bad.any(), "ensemble means differ:\n" + report


# ---------------------------------------------------------------------------
# The same comparison with the reference's DEFAULT configuration on a ~3,000-bead cell:
# P(s) up to s = 30 and the A/B compartment enrichment of the contact map.
# ---------------------------------------------------------------------------
R2 = 8
CHUNK2, BURN2, CHUNKS2 = 500, 1, 5
SMAX = 30


def _cell_default():
    cfg = synth.InterphaseConfig()                       # struct defaults of simulation_config.hpp:61-115
    d = synth.make_diploid_design(scale=0.05, cfg=cfg)
    x = synth.globule_positions(d, seed=21, radius=0.75)
    return d, cfg, x


def _observables2(d, cfg, frames, contacts, n_updates):
    chain_of = np.full(d.n, -1)
    for c, (s, e) in enumerate(d.chains):
        chain_of[s:e] = c
    i, j, v = contacts[:, 0].astype(int), contacts[:, 1].astype(int), contacts[:, 2].astype(float)
    cis = (chain_of[i] == chain_of[j]) & (chain_of[i] >= 0)
    sep = j - i
    ps = []
    for s in range(1, SMAX + 1):
        npairs = sum(max(e - b - s, 0) for b, e in d.chains)
        ps.append(v[cis & (sep == s)].sum() / (npairs * n_updates))
    # compartment enrichment of non-adjacent contacts (|i - j| > 4 or trans): same-type
    # contact frequency per available pair over cross-type frequency per available pair
    ta, tb = d.a > d.b, d.b > d.a
    far = (~cis) | (sep > 4)
    na, nb = float(ta.sum()), float(tb.sum())
    aa = v[far & ta[i] & ta[j]].sum() / (na * (na - 1) / 2)
    bb = v[far & tb[i] & tb[j]].sum() / (nb * (nb - 1) / 2)
    ab = v[far & ((ta[i] & tb[j]) | (tb[i] & ta[j]))].sum() / (na * nb)
    rg2 = np.mean([_rg2(x, d.chains).mean() for x in frames])
    return np.array([*ps, (aa + bb) / (2 * ab), bb / aa, rg2, v.sum() / n_updates])


NAMES2 = [f"P({s})" for s in range(1, SMAX + 1)] + ["(AA+BB)/2AB", "BB/AA", "Rg2", "contacts/frame"]


def test_ensemble_observables_default_config(oracle_kind):
    d, cfg, x = _cell_default()
    assert d.n >= 3000
    n_updates = (CHUNKS2 - BURN2) * CHUNK2 // cfg.contactmap_update_interval

    g = gc_b200.System(d.n, R2, 0)
    stages.setup_interphase(g, d, cfg)
    g.set_positions(x)
    g.set_seeds([9000 + r for r in range(R2)])
    stages.begin_interphase(g, cfg)
    frames = [[] for _ in range(R2)]
    for k in range(CHUNKS2):
        g.run(CHUNK2, cfg.timestep, restart=(k == 0))
        if k + 1 == BURN2:
            g.contacts_clear()
        if k + 1 > BURN2:
            for r in range(R2):
                frames[r].append(g.get_positions(r))
    G = np.array([_observables2(d, cfg, frames[r], g.contacts_fetch(r), n_updates) for r in range(R2)])
    g.close()

    F = [None] * R2
    systems = [O.OracleSystem(d.n, oracle_kind) for _ in range(R2)]

    def work(r):
        o = systems[r]
        stages.setup_interphase(o, d, cfg)
        o.set_positions(x)
        stages.interphase_callback0(o, cfg)
        fr = []
        for k in range(CHUNKS2):
            o.run(CHUNK2, cfg.timestep, interphase_callback=True, step_offset=k * CHUNK2, stepper=0,
                  seed=11000 + 100 * r + k)
            if k + 1 == BURN2:
                o.contacts_clear()
            if k + 1 > BURN2:
                fr.append(o.get_positions())
        F[r] = _observables2(d, cfg, fr, o.contacts_fetch(), n_updates)
        o.close()

    threads = [threading.Thread(target=work, args=(r,)) for r in range(R2)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    F = np.array(F)
    mg, mf = G.mean(axis=0), F.mean(axis=0)
    se = np.sqrt(G.var(axis=0, ddof=1) / R2 + F.var(axis=0, ddof=1) / R2)
    report = "\n".join(f"{n:15s} gpu {a:.5g}  ref {b:.5g}  diff {a - b:+.3g}  allowed {Z * s + REL * abs(b):.3g}"
                       for n, a, b, s in zip(NAMES2, mg, mf, se))
    print(report)
    assert (mf[:SMAX] > 0).all(), "degenerate P(s):\n" + report
    assert mf[SMAX] > 1.0, "no compartment signal in the reference:\n" + report
    bad = np.abs(mg - mf) > Z * se + REL * np.abs(mf)
    assert not bad.any(), "ensemble means differ:\n" + report

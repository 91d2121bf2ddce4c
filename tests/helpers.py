"""Shared builders for the parity tests (test infrastructure)."""
import numpy as np

from gcpkg import gc_b200

synth = gc_b200.synth
stages = gc_b200.stages


def small_interphase(scale=0.05, radius=0.75, seed=3, **cfg_kw):
    """A small diploid-like system: ~3k beads at scale 0.05."""
    cfg = synth.InterphaseConfig(**cfg_kw)
    d = synth.make_diploid_design(scale=scale, cfg=cfg)
    x = synth.globule_positions(d, seed=seed, radius=radius)
    return d, cfg, x


def build_pair(make_sys, design, cfg, x, **kw):
    s = make_sys(design.n)
    terms = stages.setup_interphase(s, design, cfg, **kw)
    s.set_positions(x)
    return s, terms


def rel_force_err(F, Fref):
    """max_i |dF_i| / max_i |F_i|  (the norm SURVEY.md 8c states)."""
    num = np.linalg.norm(F - Fref, axis=1).max()
    den = np.linalg.norm(Fref, axis=1).max()
    return num / den if den > 0 else num

"""Inter-stage transitions (host side, O(N) one-shot).

Restated from the reference's stage_transition programs (SIM = src/simulation):
  telophase -> relaxation      SIM/stage_transition/transition_interphase.cpp:13-84
  interphase -> prometaphase   SIM/stage_transition/transition_prometaphase.cpp:24-105
  prometaphase -> anaphase     SIM/stage_transition/transition_cycle.cpp:25-76
The cubic spline is the reference's not-a-knot spline (cxx-spline, whose own tests
check it against scipy.interpolate.CubicSpline, submodules/.../cxx-spline/tests).
"""
from __future__ import annotations

import numpy as np


def resample_chain(chain: np.ndarray, new_len: int) -> np.ndarray:
    """Fit a not-a-knot cubic spline through `chain` at t = (k + 1/2) / n and
    resample it at t = (i + 1/2) / new_len (transition_interphase.cpp:16-41)."""
    from scipy.interpolate import CubicSpline
    n = len(chain)
    t_new = (0.5 + np.arange(new_len)) / new_len
    if n == 1:
        return np.repeat(chain, new_len, axis=0)
    t = (0.5 + np.arange(n)) / n
    if n == 2:       # a line through two knots
        w = ((t_new - t[0]) / (t[1] - t[0]))[:, None]
        return chain[0] * (1 - w) + chain[1] * w
    return CubicSpline(t, chain, bc_type="not-a-knot", extrapolate=True)(t_new)


def transition_interphase(mitotic_design, interphase_design, telophase_positions):
    """Refines the coarse telophase structure into the interphase one; nucleolar
    beads start on top of their NOR bead (transition_interphase.cpp:62-78)."""
    x = np.zeros((interphase_design.n, 3))
    for (ms, me), (s, e) in zip(mitotic_design.chains, interphase_design.chains):
        x[s:e] = resample_chain(np.asarray(telophase_positions[ms:me], dtype=np.float64), e - s)
    for nor, nuc in interphase_design.nucleolar_bonds:
        x[nuc] = x[nor]
    return x


def transition_prometaphase(interphase_design, prometaphase_design, mitotic_cfg,
                            interphase_positions):
    """Coarse-grains every interphase chromosome into a chromatid (centroids of
    `coarse_graining` consecutive beads) and places its sister copy displaced by
    -sister_separation along the spindle axis (transition_prometaphase.cpp:62-98)."""
    axis = np.asarray(mitotic_cfg.spindle_axis, dtype=np.float64)
    disp = -mitotic_cfg.sister_separation * axis / np.linalg.norm(axis)
    cg = int(mitotic_cfg.coarse_graining)
    xi = np.asarray(interphase_positions, dtype=np.float64)
    x = np.zeros((prometaphase_design.n, 3))
    for c, (s, e) in enumerate(interphase_design.chains):
        t_idx, s_idx = prometaphase_design.sisters[c]
        ts, te = prometaphase_design.chains[t_idx]
        ss, _ = prometaphase_design.chains[s_idx]
        for off in range(te - ts):
            # the reference's window, as written (transition_prometaphase.cpp:87-88):
            # [start, min(start + cg, start + source_length)).  With prepare's coarse length
            # floor(len / cg) it never leaves the chain; where a design asks for more coarse
            # beads than that it runs on into the following chain, exactly as the reference
            # does.  Only the end of the position array stops it here (the reference would
            # read out of bounds there; synthetic test designs with >= 3 coarse beads on
            # very short chains reach that case).
            a = min(s + cg * off, len(xi) - 1)
            b = min(a + cg, a + (e - s), len(xi))
            centroid = xi[a:b].sum(axis=0) / float(b - a)
            x[ts + off] = centroid
            x[ss + off] = centroid + disp
    return x


def transition_cycle(prometaphase_design, anaphase_design, mitotic_cfg, metaphase_positions):
    """The target chromatids, displaced by -spindle_axis, start the daughter cell's
    anaphase (transition_cycle.cpp:55-70)."""
    disp = -np.asarray(mitotic_cfg.spindle_axis, dtype=np.float64)
    xm = np.asarray(metaphase_positions, dtype=np.float64)
    x = np.zeros((anaphase_design.n, 3))
    for c, (s, e) in enumerate(anaphase_design.chains):
        t_idx, _ = prometaphase_design.sisters[c]
        ms, me = prometaphase_design.chains[t_idx]
        x[s:e] = xm[ms:ms + (e - s)] + disp
    return x

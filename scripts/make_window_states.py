"""Generates tests/golden/interphase_window_60672.npz: states of ONE cell along the
benchmark's trajectory window (the first 10^4 interphase steps from the relaxed
t = 0 state = BASELINE configs[0]), at steps 0, 2500, 5000, 7500 and 10000.

bench.py times the CUDA arm over that window; the reference arm (`--impl
reference`) and the in-line `cpu_baseline` leg cannot afford 10^4 CPU steps, so
they sample the SAME window: a few steps from each of these states (positions +
driver context), time-weighted.  Both arms therefore see the same pair density
(printed per state), which the round-1 bench did not guarantee.

Needs a B200 (run under gpurun; the output lands in gpurun_out/ and is copied to
tests/golden/ by hand):

    gpurun -- python scripts/make_window_states.py
"""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from gcpkg import gc_b200  # noqa: E402

synth, stages, ens = gc_b200.synth, gc_b200.stages, gc_b200.ensemble

MARKS = [0, 2500, 5000, 7500, 10000]
cfg = synth.InterphaseConfig()
d = synth.make_diploid_design(cfg=cfg)
z = np.load(ROOT / "tests" / "golden" / "interphase_relaxed_60672.npz")
x0 = z["positions"].astype(np.float64)

g = gc_b200.System(d.n, 1, 0)
stages.setup_interphase(g, d, cfg)
g.set_seeds(ens.replica_seeds(20250001, [0]))
g.set_positions(x0)
stages.begin_interphase(g, cfg)

pos, ctxs, pairs, energy = [], [], [], []
done = 0
for k, mark in enumerate(MARKS):
    if mark > done:
        g.run(mark - done, cfg.timestep, restart=(done == 0))
        done = mark
    c = g.get_context(0)
    x = synth.quantize_f32(g.get_positions(0))
    dcut = max(cfg.a_core_diameter, cfg.b_core_diameter) * c.core_scale
    pos.append(x.astype(np.float32))
    ctxs.append([c.time, c.semiaxes[0], c.semiaxes[1], c.semiaxes[2], c.core_scale, c.bond_scale,
                 c.contact_distance])
    pairs.append(len(g.neighbor_pairs(dcut)) / d.n)
    energy.append(g.compute_energy() / d.n)
    print(f"step {mark}: t={c.time:.5f} semiaxes={list(c.semiaxes)} core={c.core_scale:.5f} "
          f"dcut={dcut:.4f} pairs/bead={pairs[-1]:.2f} E/N={energy[-1]:.3f}", flush=True)

out = ROOT / "gpurun_out" / "interphase_window_60672.npz"
out.parent.mkdir(exist_ok=True)
np.savez_compressed(out, steps=np.asarray(MARKS), positions=np.stack(pos),
                    context=np.asarray(ctxs), pairs_per_bead=np.asarray(pairs),
                    energy_per_bead=np.asarray(energy),
                    context_fields=np.asarray(["time", "semiaxis_x", "semiaxis_y", "semiaxis_z",
                                               "core_scale", "bond_scale", "contact_distance"]))
print("->", out, out.stat().st_size, "bytes")

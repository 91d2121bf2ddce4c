"""BASELINE.json configs[3]: force-only microbenchmark -- nonbonded + bonded + wall evaluation
on a 100k-bead synthetic polymer in an ellipsoid nucleus (semiaxes 5, 4, 3; A/B blocks of 20;
core_scale = 1; seed 4; SURVEY.md 8d "Config 4").  One evaluation = cell-list rebuild + pair
kernel + single-bead terms, timed per kernel with CUDA events on the library stream; the
oracle evaluates the same state on one host core (its Verlet list rebuilt every call, as the
reference does after any move) and the forces are compared (max_i|dF_i| / max_i|F_i|).
usage: python scripts/bench_force.py [--beads N] [--evals K] [--out file.json]"""
import argparse
import json
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "oracle"))
from gcpkg import gc_b200  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--beads", type=int, default=100000)
ap.add_argument("--evals", type=int, default=20)
ap.add_argument("--no-oracle", action="store_true")
ap.add_argument("--out", default="")
a = ap.parse_args()

synth, stages = gc_b200.synth, gc_b200.stages
semi = (5.0, 4.0, 3.0) if a.beads <= 200000 else tuple(s * (a.beads / 1e5) ** (1 / 3) for s in (5.0, 4.0, 3.0))
cfg = synth.InterphaseConfig(wall_semiaxes_init=semi, nucleolus_droplet_energy=0.0)
design, x = synth.make_polymer(a.beads, seed=4, semiaxes=semi)


def build(s):
    stages.setup_interphase(s, design, cfg, wall_ode=False, schedule=False, contacts=False)
    s.set_context(core_scale=1.0, bond_scale=1.0)
    s.set_positions(x)


g = gc_b200.System(design.n, 1, 0)
build(g)
F = g.compute_force()                     # warm-up (also the parity sample)
g.profile_enable(True)
for _ in range(a.evals):
    g.compute_force()
prof = g.profile_read()
g.profile_enable(False)
kern = {k: v[1] / a.evals for k, v in prof.items()}
gpu_ms = sum(kern.values())
pairs = len(g.neighbor_pairs(max(cfg.a_core_diameter, cfg.b_core_diameter))) / design.n
rep = {"workload": f"force-only, {design.n} beads, semiaxes {semi}", "gpu_ms_per_eval": gpu_ms,
       "gpu_kernels_ms": {k: round(v, 5) for k, v in kern.items()},
       "bead_evals_per_s": design.n / gpu_ms * 1e3, "pairs_within_cutoff_per_bead": pairs}
if not a.no_oracle:
    import oracle_py as O
    kind = "reference" if O.available("reference") else "port"
    o = O.OracleSystem(design.n, kind)
    build(o)
    Fo = o.compute_force()
    t0 = time.perf_counter()
    n_cpu = 3
    for k in range(n_cpu):
        o.set_positions(x)                # invalidates the Verlet list like a move does
        o.compute_force()
    cpu_ms = (time.perf_counter() - t0) / n_cpu * 1e3
    err = np.linalg.norm(F - Fo, axis=1).max() / np.linalg.norm(Fo, axis=1).max()
    rep.update({"oracle": kind, "cpu_ms_per_eval_1core": cpu_ms, "rel_force_err": float(err),
                "speedup_vs_1core": cpu_ms / gpu_ms})
line = json.dumps(rep)
print(line)
if a.out:
    Path(a.out).write_text(line + "\n")

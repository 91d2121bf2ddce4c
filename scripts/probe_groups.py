"""Timing probe: R cells on one GPU as G concurrent contexts (one stream each, one host
thread each) instead of one context.  The latency-bound kernels of one group (cell scan,
finalize, count/fill) then overlap the pair kernel of the others.

    python scripts/probe_groups.py [R=8] [groups=1,2,4] [steps=1000]
"""
import sys, time, threading
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import os
import numpy as np
from gcpkg import gc_b200
synth, stages = gc_b200.synth, gc_b200.stages

R = int(sys.argv[1]) if len(sys.argv) > 1 else 8
groups = [int(a) for a in sys.argv[2].split(",")] if len(sys.argv) > 2 else [1, 2, 4]
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 1000
relax = int(os.environ.get("QB_RELAX", 3000))
cfg = synth.InterphaseConfig()
d = synth.make_diploid_design(scale=1.0, cfg=cfg)
x = synth.globule_positions(d, seed=1, radius=1.5)
print(f"n={d.n}", flush=True)


def each(ctxs, fn):
    th = [threading.Thread(target=fn, args=(g,)) for g in ctxs]
    for t in th:
        t.start()
    for t in th:
        t.join()


for G in groups:
    per = R // G
    ctxs = []
    for k in range(G):
        g = gc_b200.System(d.n, per, 0)
        stages.setup_interphase(g, d, cfg)
        g.set_positions(x)
        g.set_seeds(np.arange(1 + k * per, 1 + (k + 1) * per, dtype=np.uint64))
        ctxs.append(g)
    each(ctxs, lambda g: g.run(relax, cfg.timestep, spacestep=cfg.relaxation_spacestep))
    for g in ctxs:
        stages.interphase_callback0(g, cfg)
    each(ctxs, lambda g: g.run(200, cfg.timestep))
    best = 0.0
    for rep in range(3):
        t0 = time.time()
        each(ctxs, lambda g: g.run(steps, cfg.timestep, restart=False))
        dt = time.time() - t0
        best = max(best, R * d.n * steps / dt)
        print(f"R={R} groups={G}: {dt/steps*1e6:.1f} us/step {R*d.n*steps/dt:.3e} bead-steps/s", flush=True)
    e = [g.compute_energy() / d.n for g in ctxs]
    print(f"R={R} groups={G}: best {best:.3e}  E/N={e[0]:.4f}", flush=True)
    for g in ctxs:
        g.close()

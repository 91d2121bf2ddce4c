"""Throwaway timing probe (not the contract bench): diploid interphase, R replicas."""
import sys, time
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import os
import numpy as np
from gcpkg import gc_b200
synth, stages = gc_b200.synth, gc_b200.stages

scale = float(sys.argv[1]) if len(sys.argv) > 1 else 1.0
reps = [int(a) for a in sys.argv[2].split(",")] if len(sys.argv) > 2 else [1, 8]
skins = [float(a) for a in sys.argv[3].split(",")] if len(sys.argv) > 3 else [0.0, 0.15, 0.3]
cfg = synth.InterphaseConfig()
t0 = time.time()
d = synth.make_diploid_design(scale=scale, cfg=cfg)
x = synth.globule_positions(d, seed=1, radius=1.5 * scale ** (1 / 3))
print(f"n={d.n} gen {time.time()-t0:.1f}s", flush=True)
for R in reps:
    for skin in skins:
        g = gc_b200.System(d.n, R, 0)
        stages.setup_interphase(g, d, cfg)
        g.set_positions(x)
        g.set_skin(skin)
        g.set_seeds(np.arange(1, R + 1, dtype=np.uint64))
        t0 = time.time()
        g.run(int(os.environ.get("QB_RELAX", 200)), cfg.timestep, spacestep=cfg.relaxation_spacestep)      # relaxation
        t_relax = time.time() - t0
        stages.interphase_callback0(g, cfg)
        g.run(int(os.environ.get("QB_WARM", 100)), cfg.timestep)                                          # warm-up
        steps = int(os.environ.get("QB_STEPS", 400))
        rb0 = g.get_context().rebuilds
        t0 = time.time()
        g.run(steps, cfg.timestep, restart=False)
        dt = time.time() - t0
        c = g.get_context()
        pairs = len(g.neighbor_pairs(0.3 * c.core_scale)) / d.n
        print(f"R={R} skin={skin}: {dt/steps*1e6:.1f} us/step  {R*d.n*steps/dt:.3e} bead-steps/s  "
              f"rebuilds/step={(c.rebuilds-rb0)/steps:.2f} pairs/bead={pairs:.1f} "
              f"relax200={t_relax:.2f}s semi={c.semiaxes[0]:.3f} E/N={g.compute_energy()/d.n:.3f}",
              flush=True)
        g.close()

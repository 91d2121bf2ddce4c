"""Timing probe: prometaphase (1,172 beads, one cell), us per step.
usage: python scripts/probe_mitotic.py [steps=20000] [replicas=1]"""
import sys, time
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import numpy as np
from gcpkg import gc_b200
synth, stages = gc_b200.synth, gc_b200.stages

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
R = int(sys.argv[2]) if len(sys.argv) > 2 else 1
d = synth.make_diploid_design(scale=1.0)
mcfg = synth.MitoticConfig()
md = synth.make_mitotic_design(d, mcfg, sisters=True)
x = np.random.default_rng(5).normal(size=(md.n, 3)) * 1.5
g = gc_b200.System(md.n, R, 0)
stages.setup_prometaphase(g, md, mcfg)
g.set_positions(x)
g.set_seeds(np.arange(1, R + 1, dtype=np.uint64))
g.run(min(steps, 2000), mcfg.timestep)
l0 = g.launch_count()
t0 = time.time()
g.run(steps, mcfg.timestep, restart=False)
dt = time.time() - t0
print(f"n={md.n} R={R}: {dt/steps*1e6:.2f} us/step, {(g.launch_count()-l0)/steps:.1f} launches/step, "
      f"E/N={g.compute_energy()/md.n:.4f}", flush=True)

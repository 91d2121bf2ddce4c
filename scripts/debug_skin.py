import sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "oracle")); sys.path.insert(0, str(ROOT / "tests"))
import numpy as np
import oracle_py as O
from gcpkg import gc_b200
from helpers import build_pair, small_interphase, stages, synth

d, cfg, x = small_interphase(scale=0.02, radius=0.6)
steps = 60
z = np.random.default_rng(10).normal(size=(steps, d.n, 3))
o, _ = build_pair(lambda n: O.OracleSystem(n, "reference"), d, cfg, x)
stages.interphase_callback0(o, cfg)
out = o.run(steps, cfg.timestep, noise=z, interphase_callback=True, want_traj=True)
for skin in (0.0, 0.1, 0.2, 0.5):
    row = []
    for m in (1, 2, 5, 10, 20, 30, 40, 50, 60):
        g, _ = build_pair(lambda n: gc_b200.System(n, 1, 0), d, cfg, x)
        g.set_skin(skin)
        stages.interphase_callback0(g, cfg)
        g.run(m, cfg.timestep, noise=z[:m])
        e = np.abs(g.get_positions() - out["traj"][m-1]).max()
        row.append(f"{m}:{e:.1e}/{g.get_context().rebuilds}")
    print(skin, " ".join(row))

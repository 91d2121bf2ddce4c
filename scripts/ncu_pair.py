"""ncu target: plain (ungraphed) steps of R cells set to one committed window state.
    python scripts/ncu_pair.py [R=8] [state=5000] [steps=24]"""
import sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import numpy as np
from gcpkg import gc_b200
synth, stages = gc_b200.synth, gc_b200.stages
R = int(sys.argv[1]) if len(sys.argv) > 1 else 8
mark = int(sys.argv[2]) if len(sys.argv) > 2 else 5000
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 24
cfg = synth.InterphaseConfig()
d = synth.make_diploid_design(cfg=cfg)
z = np.load(ROOT / "tests" / "golden" / "interphase_window_60672.npz")
k = list(z["steps"]).index(mark)
c = z["context"][k]
g = gc_b200.System(d.n, R, 0)
stages.setup_interphase(g, d, cfg)
g.set_seeds(np.arange(1, R + 1, dtype=np.uint64))
g.set_positions(z["positions"][k].astype(np.float64))
g.run(0, cfg.timestep, restart=True)
g.set_context(time=float(c[0]), semiaxes=[float(c[1]), float(c[2]), float(c[3])], core_scale=float(c[4]),
              bond_scale=float(c[5]), contact_distance=float(c[6]))
g.set_step(mark)
g.run(steps, cfg.timestep, restart=False)       # < 64 steps: plain launches, no graph
g.synchronize()
print("ok", g.compute_energy() / d.n)

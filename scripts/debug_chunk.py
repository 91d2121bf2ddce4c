import sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "oracle")); sys.path.insert(0, str(ROOT / "tests"))
import numpy as np
from gcpkg import gc_b200
from helpers import build_pair, small_interphase, stages, synth

d, cfg, x = small_interphase(scale=0.02, radius=0.6)
kw = dict(contacts=False, wall_ode=False, schedule=False)
def run(chunks):
    g, _ = build_pair(lambda n: gc_b200.System(n, 1, 0), d, cfg, x, **kw)
    g.set_seeds([1234])
    first = True
    grids = []
    for c in chunks:
        g.run(c, cfg.timestep, restart=first)
        first = False
        grids.append(g.debug_grid())
    return g.get_positions_f32(), grids
ref = {}
for m in range(1, 11):
    ref[m], gr = run((m,))
    print(m, gr[-1])
for chunks in ((5, 5), (5, 5), (5, 1), (5, 2), (4, 1), (6, 1), (2, 3, 5)):
    b, gr = run(chunks)
    a = ref[sum(chunks)]
    bad = np.argwhere(a != b)
    print(chunks, "nd", len(bad), bad[:6].tolist())

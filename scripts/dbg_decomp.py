import sys
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests")
import numpy as np, torch
from gcpkg import gc_b200
from helpers import small_interphase, stages
decomp = gc_b200.decomp
d, cfg, x = small_interphase()
def mk():
    g = gc_b200.System(d.n, 1, 0)
    stages.setup_interphase(g, d, cfg, contacts=False)
    g.set_positions(x); g.set_seeds([77]); stages.interphase_callback0(g, cfg)
    return g
world = 2
refs = []
for trial in range(6):
    ref = mk(); ref.run(40, cfg.timestep, restart=True); xr = ref.get_positions(0); ref.close()
    refs.append(xr)
    ranks = [mk() for _ in range(world)]
    ex = decomp.LocalExchange(ranks, halo_layers=3)
    ex.begin(cfg.timestep)
    ex.step(40)
    xd = ex.gather_positions()
    for g in ranks: g.close()
    ranks = [mk() for _ in range(world)]
    ex = decomp.LocalExchange(ranks, halo_layers=3)
    ex.begin(cfg.timestep)
    for _ in range(40): ex.step(1)
    xd2 = ex.gather_positions()
    for g in ranks: g.close()
    print("trial", trial, "ref==ref0", np.array_equal(xr, refs[0]), "dec==ref", np.array_equal(xd, xr),
          "dec2==ref", np.array_equal(xd2, xr), "dec==dec2", np.array_equal(xd, xd2), flush=True)

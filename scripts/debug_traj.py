import sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "oracle")); sys.path.insert(0, str(ROOT / "tests"))
import numpy as np
import oracle_py as O
from gcpkg import gc_b200
from helpers import build_pair, small_interphase, stages, synth, rel_force_err

d, cfg, x = small_interphase(scale=0.02, radius=0.6)
steps = 50
z = np.random.default_rng(7).normal(size=(steps, d.n, 3))
for contacts in (True, False):
    g, _ = build_pair(lambda n: gc_b200.System(n, 1, 0), d, cfg, x, contacts=contacts)
    o, _ = build_pair(lambda n: O.OracleSystem(n, "reference"), d, cfg, x, contacts=contacts)
    for s in (g, o):
        stages.interphase_callback0(s, cfg)
    out = o.run(steps, cfg.timestep, noise=z, interphase_callback=True, want_traj=True, want_ctx=True)
    for k in range(steps):
        g.run(1, cfg.timestep, noise=z[k:k+1]) if False else None
    # chunked GPU run is not possible with injected noise; run prefixes instead
    for m in (1, 5, 10, 19, 20, 21, 22, 25, 30, 40, 50):
        g2, _ = build_pair(lambda n: gc_b200.System(n, 1, 0), d, cfg, x, contacts=contacts)
        stages.interphase_callback0(g2, cfg)
        g2.run(m, cfg.timestep, noise=z[:m])
        c = g2.get_context()
        print(contacts, m, "maxdx", np.abs(g2.get_positions() - out["traj"][m-1]).max(),
              "semi", c.semiaxes[0], out["ctx"][m-1][1], "core", c.core_scale - out["ctx"][m-1][4])
# force test detail
d, cfg, x = small_interphase(scale=0.05, radius=1.3)
x = x.copy(); x[::53] *= 1.75
x = x.astype(np.float32).astype(np.float64)
g, terms = build_pair(lambda n: gc_b200.System(n, 1, 0), d, cfg, x)
o, _ = build_pair(lambda n: O.OracleSystem(n, "reference"), d, cfg, x)
for s in (g, o):
    s.set_context(core_scale=0.73, bond_scale=0.61, semiaxes=(2.0, 1.8, 2.1), contact_distance=0.1)
for name, t in terms.items():
    Fg, Fo = g.compute_force(1 << t), o.compute_force(1 << t)
    print(name, rel_force_err(Fg, Fo), np.abs(Fo).max(), g.compute_energy(1<<t), o.compute_energy(1<<t))
Fg, Fo = g.compute_force(), o.compute_force()
print("all", rel_force_err(Fg, Fo))
cg, co = g.get_context(), o.get_context()
print(list(cg.axial_reaction), list(co.axial_reaction), cg.reaction_force, co.reaction_force)

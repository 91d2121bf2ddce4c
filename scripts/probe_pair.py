"""Timing probe for the pair kernel and the step (not the contract bench): all cells of a
context are set to one committed window state (tests/golden/interphase_window_60672.npz),
then (a) a serialised per-kernel event pass, (b) graph-replay throughput.
    GC_B200_LIB=<variant.so> python scripts/probe_pair.py [R=8,1] [states=0,5000,10000] [steps=600]
"""
import os, sys, time
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import numpy as np
from gcpkg import gc_b200
synth, stages = gc_b200.synth, gc_b200.stages

reps = [int(a) for a in sys.argv[1].split(",")] if len(sys.argv) > 1 else [8, 1]
marks = [int(a) for a in sys.argv[2].split(",")] if len(sys.argv) > 2 else [0, 5000, 10000]
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 600
tag = os.environ.get("PROBE_TAG", os.path.basename(os.environ.get("GC_B200_LIB", "default")))
cfg = synth.InterphaseConfig()
d = synth.make_diploid_design(cfg=cfg)
z = np.load(ROOT / "tests" / "golden" / "interphase_window_60672.npz")
for R in reps:
    g = gc_b200.System(d.n, R, 0)
    stages.setup_interphase(g, d, cfg)
    g.set_seeds(np.arange(1, R + 1, dtype=np.uint64))
    for mark in marks:
        k = list(z["steps"]).index(mark)
        c = z["context"][k]
        g.set_positions(z["positions"][k].astype(np.float64))
        g.run(0, cfg.timestep, restart=True)
        g.set_context(time=float(c[0]), semiaxes=[float(c[1]), float(c[2]), float(c[3])], core_scale=float(c[4]),
                      bond_scale=float(c[5]), contact_distance=float(c[6]))
        g.set_step(mark)
        g.contacts_clear()
        g.run(100, cfg.timestep, restart=False)
        g.profile_enable(True)
        g.run(40, cfg.timestep, restart=False)
        prof = g.profile_read()
        g.profile_enable(False)
        ker = {kn: v[1] / 40 * 1e3 for kn, v in prof.items()}
        pair = sum(v for kn, v in ker.items() if kn.startswith("k_pair"))
        g.synchronize()
        t0 = time.time()
        g.run(steps, cfg.timestep, restart=False)
        g.synchronize()
        dt = time.time() - t0
        e = g.compute_energy() / d.n
        print(f"[{tag}] R={R} state={mark}: pair {pair:.1f} us  sum {sum(ker.values()):.1f} us | graph {dt/steps*1e6:.1f} us/step "
              f"{R*d.n*steps/dt:.3e} bead-steps/s  E/N={e:.4f} | " +
              " ".join(f"{kn.replace('k_','')}={v:.1f}" for kn, v in ker.items() if v > 0.5), flush=True)
    g.close()

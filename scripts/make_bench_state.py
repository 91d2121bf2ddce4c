"""Generates tests/golden/interphase_relaxed_60672.npz: the t = 0 interphase
state bench.py starts both arms from.

The reference pipeline hands the interphase stage a telophase globule relaxed
by run_relaxation (SIM/stage_interphase/simulation_driver_relaxation.cpp:8-56:
10^4 adaptive-dt steps, spacestep 1e-3, scales frozen at their initial values).
That relaxation is run here ON THE DEVICE (it needs a B200: run under gpurun,
output lands in gpurun_out/ and is then copied to tests/golden/), the result is
quantised exactly as the reference stores frames (16 mantissa bits, f32;
SIM/common/simulation_store.cpp:22-32) and saved.

    gpurun -- python scripts/make_bench_state.py [relax_steps]
"""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from gcpkg import gc_b200  # noqa: E402

synth, stages = gc_b200.synth, gc_b200.stages

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 10000
cfg = synth.InterphaseConfig()
d = synth.make_diploid_design(cfg=cfg)
x0 = synth.globule_positions(d, seed=1, radius=1.5)
g = gc_b200.System(d.n, 1, 0)
stages.setup_interphase(g, d, cfg)
g.set_positions(x0)
g.set_seeds([20250001])
stages.begin_relaxation(g, cfg)
g.run(steps, cfg.timestep, spacestep=cfg.relaxation_spacestep)
x = synth.quantize_f32(g.get_positions(0))
e = g.compute_energy() / d.n
c = g.get_context()
pairs = len(g.neighbor_pairs(0.3 * 0.5)) / d.n
out = ROOT / "gpurun_out" / "interphase_relaxed_60672.npz"
out.parent.mkdir(exist_ok=True)
how = (f"synthetic globule (seed 1, radius 1.5) + {steps} adaptive relaxation steps "
       f"(spacestep 1e-3) on device, 16-bit-mantissa f32; E/N={e:.3f}, "
       f"{pairs:.1f} pairs/bead within dcut=0.15")
np.savez_compressed(out, positions=x, how=np.asarray(how))
print(how, "last dt", c.last_timestep, "->", out, out.stat().st_size, "bytes")

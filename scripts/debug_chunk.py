import sys, time
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
import numpy as np
from gcpkg import gc_b200
synth, stages = gc_b200.synth, gc_b200.stages
cfg = synth.InterphaseConfig(); d = synth.make_diploid_design(cfg=cfg)
x = np.load('/root/repo/tests/golden/interphase_relaxed_60672.npz')
x = x[x.files[0]] if 'x' not in x.files else x['x']
g = gc_b200.System(d.n, 1, 0); stages.setup_interphase(g, d, cfg); g.set_positions(x.astype(np.float64)); stages.begin_interphase(g, cfg)
g.run(1000, cfg.timestep, restart=True); g.synchronize()
for steps in (1000, 100, 20):
    t0 = time.perf_counter()
    for k in range(10): g.run(steps, cfg.timestep, restart=False)
    g.synchronize(); t1 = time.perf_counter()
    for k in range(10): g.get_frame(0, 16)
    t2 = time.perf_counter()
    print(steps, 'run ms/call', (t1 - t0) * 100, 'us/step', (t1 - t0) / 10 / steps * 1e6, 'frame ms', (t2 - t1) * 100)

"""BASELINE.json configs[1]: one cell through the whole cycle on one B200, with the
reference's default step counts (--scale-steps 1) or a fraction of them.
usage: python scripts/run_cycle.py [--scale-steps F] [--scale S] [--out file.json]"""
import argparse
import importlib
import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))
from gcpkg import gc_b200  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--scale-steps", type=float, default=1.0)
ap.add_argument("--scale", type=float, default=1.0, help="chromosome length scale (1 = diploid human)")
ap.add_argument("--seed", type=int, default=1)
ap.add_argument("--out", default="")
a = ap.parse_args()

design = gc_b200.synth.make_diploid_design(scale=a.scale)
rep = gc_b200.cycle.run_cell_cycle(design=design, seed=a.seed, scale_steps=a.scale_steps,
                                   log=lambda row: print(row, file=sys.stderr, flush=True))
rep.pop("next_anaphase_start")
rep["config"] = {"workload": "full cell cycle, single cell, 1 GPU (BASELINE configs[1])",
                 "scale_steps": a.scale_steps, "scale": a.scale, "seed": a.seed}
line = json.dumps(rep)
print(line)
if a.out:
    Path(a.out).write_text(line + "\n")

"""BASELINE.json configs[4]: one synthetic polymer spread over the GPUs of a box by the
slab + halo decomposition of csrc/gc_decomp.cuh: per step one all-gather of 128-byte headers
and one NCCL send/recv of halo beads with each neighbouring rank (SURVEY.md 8d "Config 5": 1 chain x N beads at config-4 density, fixed dt,
interphase force set).
  torchrun --nproc-per-node G --master-addr 127.0.0.1 scripts/run_decomp.py --beads 2000000
Prints one JSON line on rank 0: decomposed ms/step (CUDA events, max over ranks), the same
polymer undecomposed on one GPU (rank 0), and with --check a bit-exact comparison of the
two trajectories."""
import argparse
import json
import os
import sys
from pathlib import Path

import numpy as np
import torch
import torch.distributed as dist

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from gcpkg import gc_b200  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--beads", type=int, default=2_000_000)
ap.add_argument("--steps", type=int, default=100)
ap.add_argument("--warmup", type=int, default=20)
ap.add_argument("--check", action="store_true")
ap.add_argument("--profile", type=int, default=0, help="per-kernel event timing over this many steps")
ap.add_argument("--halo", type=int, default=3, help="position halo in cell layers")
ap.add_argument("--cap-frac", type=float, default=0.0, help="halo buffer beads / (N / world); 0 = default")
ap.add_argument("--peer", action="store_true", help="halo through NVLink peer memory (CUDA IPC)")
ap.add_argument("--graph", type=int, default=0, help="steps per CUDA graph (0: eager)")
ap.add_argument("--out", default="")
a = ap.parse_args()

rank = int(os.environ.get("RANK", "0"))
world = int(os.environ.get("WORLD_SIZE", "1"))
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))

synth, stages, decomp = gc_b200.synth, gc_b200.stages, gc_b200.decomp
semi = tuple(s * (a.beads / 1e5) ** (1 / 3) for s in (5.0, 4.0, 3.0))
cfg = synth.InterphaseConfig(wall_semiaxes_init=semi, nucleolus_droplet_energy=0.0,
                             wall_semiaxes_spring=(3e4 * a.beads / 60672,) * 3)
design, x = synth.make_polymer(a.beads, seed=4, semiaxes=semi)


def make():
    g = gc_b200.System(design.n, 1, local)
    stages.setup_interphase(g, design, cfg, contacts=False)
    g.set_positions(x)
    g.set_seeds([4])
    stages.interphase_callback0(g, cfg)
    return g


def timed(fn, stream):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0.record(stream)
    fn()
    e1.record(stream)
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=f"cuda:{local}")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


rep = {"workload": f"1 chain x {design.n} beads, semiaxes {tuple(round(s, 2) for s in semi)}, interphase "
                   "force set, fixed dt, wall ODE + expansion schedule on", "world": world,
       "steps": a.steps, "warmup": a.warmup}
x_dec = None
if world > 1:
    g = make()
    run = decomp.DecomposedRun(g, rank, world, local, halo_layers=a.halo,
                               halo_capacity=int(a.cap_frac * design.n / world),
                               graph_steps=a.graph, peer=a.peer)
    run.begin(cfg.timestep)
    run.step(a.warmup // 2)
    rep["halo_capacity_beads"] = run.tune_capacity()
    run.step(a.warmup - a.warmup // 2)
    import time
    enq = [0.0]

    def block():
        t0 = time.perf_counter()
        run.step(a.steps)
        enq[0] = time.perf_counter() - t0

    ms = timed(block, run.stream)
    rep["graph_steps"] = a.graph
    rep["host_enqueue_ms_per_step"] = enq[0] * 1e3 / a.steps
    g.synchronize()
    rep["decomposed_ms_per_step"] = ms / a.steps
    rep["decomposed_bead_steps_per_s"] = design.n * a.steps / (ms * 1e-3)
    rep["scheme"] = run.scheme
    rep["halo_layers"] = a.halo
    rep["exchange_bytes_per_rank_step"] = int(run.exchange_bytes_per_step)
    counts = torch.tensor(list(run.halo_counts()) + [int(g.decomp_owned().sum())], device=f"cuda:{local}")
    allc = [torch.zeros_like(counts) for _ in range(world)]
    dist.all_gather(allc, counts)
    rep["halo_beads_sent_lo_hi_and_owned_per_rank"] = [c.tolist() for c in allc]
    if a.profile:
        ph = run.timed_phases(a.profile)
        allp = [None] * world
        dist.all_gather_object(allp, ph)
        rep["phase_ms_per_rank"] = [{k: round(v, 5) for k, v in p.items()} for p in allp]
    if a.profile and rank == 0:
        g.profile_enable(True)
    if a.profile:
        run.step(a.profile)
        g.synchronize()
    if a.profile and rank == 0:
        rep["decomposed_kernels_ms"] = {k: round(v[1] / a.profile, 5) for k, v in g.profile_read().items()}
        g.profile_enable(False)
    run.invalidate_graph()
    if a.check:
        x_dec = run.gather_positions()
    run.close()
    g.close()
if rank == 0:
    g = make()
    stream = torch.cuda.ExternalStream(g.stream(), device=local)
    g.run(a.warmup, cfg.timestep, restart=True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record(stream)
    g.run(a.steps, cfg.timestep, restart=False)
    e1.record(stream)
    torch.cuda.synchronize()
    ms1 = e0.elapsed_time(e1)
    rep["single_gpu_ms_per_step"] = ms1 / a.steps
    rep["single_gpu_bead_steps_per_s"] = design.n * a.steps / (ms1 * 1e-3)
    if a.profile:
        g.profile_enable(True)
        g.run(a.profile, cfg.timestep, restart=False)
        rep["single_gpu_kernels_ms"] = {k: round(v[1] / a.profile, 5) for k, v in g.profile_read().items()}
        g.profile_enable(False)
    if a.profile and world > 1:
        g.run(a.profile, cfg.timestep, restart=False)      # the decomposed run also took timed_phases()
    if world > 1:
        rep["speedup_vs_single_gpu"] = rep["single_gpu_ms_per_step"] / rep["decomposed_ms_per_step"]
    if a.check and x_dec is not None:
        x_one = g.get_positions(0)
        rep["bit_identical_to_single_gpu"] = bool(np.array_equal(x_one, x_dec))
        if not rep["bit_identical_to_single_gpu"]:
            bad = np.where(np.any(x_one != x_dec, axis=1))[0]
            rep["mismatch"] = {"beads": int(len(bad)), "max_abs": float(np.abs(x_one - x_dec).max()),
                               "z_min": float(x_one[bad, 2].min()), "z_max": float(x_one[bad, 2].max()),
                               "first": bad[:8].tolist()}
    g.close()
    line = json.dumps(rep)
    print(line)
    if a.out:
        Path(a.out).write_text(line + "\n")
if world > 1:
    dist.barrier()
    dist.destroy_process_group()

#!/usr/bin/env python
"""bench.py -- throughput of the force-evaluation + Brownian-integration hot path.

    python bench.py --gpus N --steps K --warmup W            # CUDA path (this repo)
    python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU path

Workload (config.workload): the DIPLOID INTERPHASE STAGE of BASELINE.json
(configs[0]/[2] topology: 46 chains + nucleolar beads, 60,672 beads) as an
ensemble of R independent cell trajectories per GPU (weak scaling: 8 per GPU,
i.e. configs[2]'s 64 cells at 8 GPUs).  One bench "step" = S consecutive
simulation steps of every replica: cell-list rebuild, soft-core pair forces,
bonds, nucleolar droplet, ellipsoid wall + wall ODE, parameter schedule,
BAOAB-limit update with Philox noise, contact-map update every 20 steps.
metric = bead-steps/s (beads x replicas x simulation steps / second).

value  : positions resident in HBM; CUDA events on the library's stream.
e2e    : the same through the C-ABI with HOST buffers: every bench step uploads
         the positions of all replicas, runs S steps, downloads positions,
         context and energy (what the reference driver does at sampling cadence).
roofline / cpu_baseline: see DESIGN.md sections 5-6.

oracle/ is used here ONLY for the cpu_baseline leg and for --impl reference.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

N_BEADS_FULL = 60672
STATE_FIXTURE = ROOT / "tests" / "golden" / "interphase_relaxed_60672.npz"
ALG_BYTES_PER_BEAD_STEP = 32.0         # float4 in + float4 out
FLOP_PER_PAIR = 40.0
FLOP_SINGLE_BEAD = 125.0               # interphase single-bead terms


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--replicas", type=int, default=8, help="cell trajectories per GPU")
    ap.add_argument("--sim-steps", type=int, default=500, help="simulation steps per bench step")
    ap.add_argument("--scale", type=float, default=1.0, help="genome scale (tests only; 1 = diploid)")
    ap.add_argument("--relax-steps", type=int, default=2000,
                    help="adaptive relaxation steps when no cached state exists")
    ap.add_argument("--cpu-sample-steps", type=int, default=20,
                    help="simulation steps per core of the CPU baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--profile-steps", type=int, default=60)
    return ap.parse_args()


# ---------------------------------------------------------------------------
# workload
# ---------------------------------------------------------------------------
def make_workload(scale):
    from gcpkg import gc_b200
    synth = gc_b200.synth
    cfg = synth.InterphaseConfig()
    design = synth.make_diploid_design(scale=scale, cfg=cfg)
    x0 = synth.globule_positions(design, seed=1, radius=1.5 * scale ** (1.0 / 3.0))
    return design, cfg, x0


def cached_state(design):
    """The relaxed t = 0 interphase state (scripts/make_bench_state.py), if committed."""
    if design.n == N_BEADS_FULL and STATE_FIXTURE.exists():
        z = np.load(STATE_FIXTURE)
        if z["positions"].shape == (design.n, 3):
            return z["positions"].astype(np.float64), str(z["how"])
    return None, None


# ---------------------------------------------------------------------------
# clocks
# ---------------------------------------------------------------------------
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device):
        self.device = device
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.device}", f"--query-gpu={self.Q}",
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
            out, _ = self.proc.communicate()
        sm, mx, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in out.strip().splitlines():
            f = [t.strip() for t in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
                power.append(float(f[3]))
            except ValueError:
                continue
            for name, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None,
                "sm_max_mhz": float(max(mx)) if mx else None,
                "power_w_max": float(max(power)) if power else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# ---------------------------------------------------------------------------
# CPU reference leg (oracle/_ref = micromd compiled unmodified, else the port)
# ---------------------------------------------------------------------------
def cpu_reference_run(design, cfg, x_start, ctx0, sim_steps, rounds, warm_rounds, cores=None):
    """`cores` independent single-threaded trajectories (how the reference uses
    more than one core: one process per cell), each `sim_steps` interphase steps
    per round, driver callback (schedule, wall ODE, contact map) included.
    Returns (bead-steps/s over the timed rounds, ms per round, kind, cores)."""
    sys.path.insert(0, str(ROOT / "oracle"))
    import oracle_py as O
    from gcpkg import gc_b200
    stages = gc_b200.stages
    kind = "reference" if O.available("reference") else "port"
    if kind == "port":
        O.build(("port",))
    cores = cores or os.cpu_count() or 1
    systems = []
    for c in range(cores):
        o = O.OracleSystem(design.n, kind)
        stages.setup_interphase(o, design, cfg)
        o.set_positions(x_start)
        if ctx0 is None:
            stages.interphase_callback0(o, cfg)
        else:
            o.set_context(**ctx0)
        o.set_seeds([1000 + c])
        systems.append(o)
    offset = [0]

    def one_round():
        def work(o):
            o.run(sim_steps, cfg.timestep, interphase_callback=True, step_offset=offset[0],
                  stepper=0)
        th = [threading.Thread(target=work, args=(o,)) for o in systems]
        t0 = time.perf_counter()
        for t in th:
            t.start()
        for t in th:
            t.join()
        offset[0] += sim_steps
        return time.perf_counter() - t0

    for _ in range(warm_rounds):
        one_round()
    total = sum(one_round() for _ in range(rounds))
    for o in systems:
        o.close()
    rate = cores * design.n * sim_steps * rounds / total
    return rate, total / rounds * 1e3, kind, cores


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    design, cfg, x0 = make_workload(args.scale)
    x, how = cached_state(design)
    if x is None:
        x, how = x0, "unrelaxed synthetic globule (no cached relaxed state)"
    cores = os.cpu_count() or 1
    rate, ms, kind, cores = cpu_reference_run(design, cfg, x, None, args.cpu_sample_steps,
                                              args.steps, args.warmup, cores)
    sample = (f"{cores} independent single-threaded trajectories (one per host core), "
              f"{args.cpu_sample_steps} interphase steps each per bench step, from t = 0")
    line = {
        "impl": "reference", "metric": "bead-steps/s", "value": rate, "unit": "bead-steps/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": workload_config(design, args, how, reference=True),
        "cpu_baseline": {"value": rate, "unit": "bead-steps/s", "cores": cores,
                         "kind": "reference" if kind == "reference" else "port", "sample": sample},
        "e2e": {"value": rate, "unit": "bead-steps/s", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


def workload_config(design, args, how, reference=False):
    if reference:
        return {
            "workload": "diploid interphase stage (BASELINE configs[0]/[2] topology), ensemble of "
                        "independent cell trajectories",
            "beads_per_cell": int(design.n), "chains": int(len(design.chains)),
            "cells_per_gpu": None, "sim_steps_per_bench_step": args.cpu_sample_steps,
            "timestep": 1e-5, "contact_map_interval": 20, "start_state": how,
            "l2": "n/a (host run)",
            "parallelism": "reference CPU code, one single-threaded trajectory per host core",
        }
    return {
        "workload": "diploid interphase stage (BASELINE configs[0]/[2] topology), ensemble of "
                    "independent cell trajectories",
        "beads_per_cell": int(design.n), "chains": int(len(design.chains)),
        "cells_per_gpu": None if reference else args.replicas,
        "sim_steps_per_bench_step": args.cpu_sample_steps if reference else args.sim_steps,
        "timestep": 1e-5, "contact_map_interval": 20,
        "start_state": how,
        "l2": "flushed between bench steps (256 MiB memset inside the timed region); the "
              "per-step working set itself is L2-resident by design",
        "parallelism": f"ensemble x{args.gpus} (one process per GPU, no data-path collective); on each "
                       "GPU the step graph runs the cells as parallel branches (replica groups, "
                       f"GC_GROUPS={os.environ.get('GC_GROUPS', 'default: min(cells, 8)')})",
    }


# ---------------------------------------------------------------------------
# CUDA arm
# ---------------------------------------------------------------------------
def run_b200_arm(args):
    import torch
    import torch.distributed as dist
    from gcpkg import gc_b200
    stages = gc_b200.stages

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (this path has no CPU fallback)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    design, cfg, x0 = make_workload(args.scale)
    R, S, n = args.replicas, args.sim_steps, design.n
    g = gc_b200.System(n, R, local)
    stages.setup_interphase(g, design, cfg)
    ens = gc_b200.ensemble
    cells = ens.shard_replicas(world * R, world, rank)          # global cell ids of this rank
    g.set_seeds(ens.replica_seeds(20250001, cells))

    x, how = cached_state(design)
    if x is None:
        # relaxation stage on the device (run_relaxation semantics: adaptive dt,
        # scales frozen at their initial values, wall fixed)
        g.set_positions(x0)
        stages.begin_relaxation(g, cfg)
        g.run(args.relax_steps, cfg.timestep, spacestep=cfg.relaxation_spacestep)
        x = g.get_positions(0)
        how = f"synthetic globule + {args.relax_steps} adaptive relaxation steps on device"
    g.set_positions(x)
    stages.begin_interphase(g, cfg)
    ctx0 = g.get_context(0)
    ctx0 = dict(time=ctx0.time, semiaxes=list(ctx0.semiaxes), core_scale=ctx0.core_scale,
                bond_scale=ctx0.bond_scale, contact_distance=ctx0.contact_distance)
    x_start = g.get_positions(0)

    stream = torch.cuda.ExternalStream(g.stream(), device=local)
    flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device=f"cuda:{local}")

    def bench_step(first):
        with torch.cuda.stream(stream):
            flush_buf.zero_()
        g.run(S, cfg.timestep, restart=first)
        # the driver dumps and clears the contact map once per output window
        # (simulation_driver_interphase.cpp:42-45); here once per bench step
        g.contacts_clear()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        g.synchronize()

    # ---- device-resident throughput ----------------------------------------
    for w in range(args.warmup):
        bench_step(w == 0)
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = g.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for k in range(args.steps):
        bench_step(args.warmup == 0 and k == 0)
    e1.record(stream)
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    launches = g.launch_count() - l0
    ms_total = e0.elapsed_time(e1)
    t = torch.tensor([ms_total], dtype=torch.float64, device=f"cuda:{local}")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t.item())
    value = world * R * n * S * args.steps / (ms_total * 1e-3)

    # ---- end to end through the C-ABI with host buffers -------------------------
    e2e = None
    if not args.no_e2e:
        host = [g.get_positions_f32(r) for r in range(R)]

        def e2e_pass(n_steps):
            """n_steps bench steps with host buffers, then one ensemble output window."""
            barrier()
            t0 = time.perf_counter()
            energy = 0.0
            for k in range(n_steps):
                for r in range(R):
                    g.set_positions_f32(host[r], r)
                g.run(S, cfg.timestep, restart=False)
                if k + 1 < n_steps:
                    g.contacts_clear()
                for r in range(R):
                    host[r] = g.get_positions_f32(r)
                c = g.get_context(0)
                energy = g.compute_energy(replica=0) / n
            # output window: ensemble contact map (device reduce per rank, NCCL gather, device
            # merge) and observables, inside the timed region
            t1 = time.perf_counter()
            lap = {}

            def mark(name, t_prev):
                g.synchronize()
                now = time.perf_counter()
                lap[name] = (now - t_prev) * 1e3
                return now
            keys, counts = g.contacts_reduce_device(-1, local)
            tl = mark("device_reduce", t1)
            keys, counts = ens.gather_contacts(keys, counts,
                                               merge=lambda a, b: g.contacts_merge_device(a, b, local))
            tl = mark("gather_merge", tl)
            rg2 = np.stack([g.radius_of_gyration2(design.chains, r) for r in range(R)])
            obs = torch.tensor(np.concatenate([rg2.sum(axis=0), [float(R)]]), dtype=torch.float64,
                               device=f"cuda:{local}")
            obs = ens.reduce_sum(obs).cpu().numpy()
            tl = mark("rg2", tl)
            ijv = ens.keys_to_ijv(keys, counts) if rank == 0 else None
            tl = mark("d2h_triples", tl)
            g.contacts_clear()
            barrier()
            tl = mark("clear_barrier", tl)
            t_reduce = time.perf_counter() - t1
            dt = time.perf_counter() - t0
            tt = torch.tensor([dt], dtype=torch.float64, device=f"cuda:{local}")
            if world > 1:
                dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            dt = float(tt.item())
            return dt, t_reduce, lap, energy, c, obs, ijv

        e2e_pass(1)                       # untimed: page-locked staging buffers, NCCL channels
        dt, t_reduce, lap, energy, c, obs, ijv = e2e_pass(args.steps)
        e2e = {"value": world * R * n * S * args.steps / dt, "unit": "bead-steps/s",
               "h2d_bytes_per_step": R * n * 16, "d2h_bytes_per_step": R * n * 16 + 8 * 16 + 8,
               "ms_per_step": dt / args.steps * 1e3, "energy_per_bead": energy,
               "semiaxis_x": c.semiaxes[0]}
        if rank == 0:
            updates = S // cfg.contactmap_update_interval
            ps = ens.contact_probability(ijv, design.chains, updates, world * R, max_s=1000)
            e2e["ensemble"] = {
                "cells": world * R, "contact_pairs": int(len(ijv)),
                "contact_counts": int(ijv[:, 2].sum()) if len(ijv) else 0,
                "reduce_ms": t_reduce * 1e3, "reduce_breakdown_ms": {k: round(v, 3) for k, v in lap.items()},
                "collective": "nccl all_gather of per-rank (key, count) runs + all_reduce of Rg^2"
                              if world > 1 else "none (1 rank)",
                "rg_mean": float(np.sqrt(obs[:-1] / obs[-1]).mean()),
                "P_s": {str(sep): float(ps[sep]) for sep in (1, 2, 5, 10, 30, 100, 300, 1000)},
            }

    # ---- roofline of the dominant kernel: separate short pass with per-kernel events
    roof = None
    useful = None
    if rank == 0:
        args.profile_steps = max(args.profile_steps, 1)
        g.profile_enable(True)
        g.run(args.profile_steps, cfg.timestep, restart=False)
        prof = g.profile_read()
        g.profile_enable(False)
        g.contacts_clear()
        total_ms = sum(v[1] for v in prof.values())
        top = max(prof.items(), key=lambda kv: kv[1][1])
        name, (cnt, ms) = top
        per_launch_s = ms / cnt * 1e-3
        cnow = g.get_context(0)
        dcut = max(cfg.a_core_diameter, cfg.b_core_diameter) * cnow.core_scale
        useful = len(g.neighbor_pairs(dcut, replica=0)) / n      # unordered pairs within dcut per bead
        # pair evaluations the kernel actually EXECUTES: every home bead against every bead of
        # its 27 cells of edge dcut (gather over cubic cells, full neighbourhood)
        xs = g.get_positions(0)
        ci = np.floor((xs - xs.min(axis=0)) / dcut).astype(np.int64) + 1
        dims = ci.max(axis=0) + 2
        occ = np.zeros(tuple(dims), dtype=np.int64)
        np.add.at(occ, (ci[:, 0], ci[:, 1], ci[:, 2]), 1)
        cand = np.zeros_like(occ)
        for dx in (-1, 0, 1):
            for dy in (-1, 0, 1):
                for dz in (-1, 0, 1):
                    cand += np.roll(occ, (dx, dy, dz), axis=(0, 1, 2))
        executed_pairs = float((occ * cand).sum()) * R
        peaks = {}
        pk = ROOT / "MEASURED_PEAKS.json"
        if pk.exists():
            peaks = json.loads(pk.read_text())
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        # FP32 pipe peak, measured live with scalar FFMA chains and with packed FFMA2
        # chains (the pair kernel issues FFMA2); the larger one is the denominator
        os.environ.pop("GC_PEAK_PACKED", None)
        fp32_scalar = gc_b200.api.measure_fp32_peak(local)
        os.environ["GC_PEAK_PACKED"] = "1"
        fp32_packed = gc_b200.api.measure_fp32_peak(local)
        os.environ.pop("GC_PEAK_PACKED", None)
        fp32_peak = max(fp32_scalar, fp32_packed)
        beads_per_launch = R * n
        hbm_ach = ALG_BYTES_PER_BEAD_STEP * beads_per_launch / per_launch_s / 1e9
        flop_alg = (FLOP_PER_PAIR * useful + FLOP_SINGLE_BEAD) * beads_per_launch
        fp32_ach = flop_alg / per_launch_s / 1e12
        roof = {
            "kernel": name, "bound": "fp32", "achieved": fp32_ach, "peak": fp32_peak,
            "unit": "TFLOP/s", "frac": fp32_ach / fp32_peak if fp32_peak else None,
            "peak_source": "measured live: FFMA / FFMA2 chain microbenchmarks (gc_measure_fp32_peak), max of "
                           f"scalar {fp32_scalar:.2f} and packed {fp32_packed:.2f} TFLOP/s",
            # dram__bytes_read.sum + dram__bytes_write.sum of one k_pair_cells launch from the
            # committed ncu --set full capture of this workload (8 cells, relaxed state)
            "traffic": 13.12e6 if (R == 8 and args.scale == 1.0 and name.startswith("k_pair_cells")) else None,
            "traffic_source": "profiles/r1_pair_cells_v8_sparse_ncu_summary.txt (13.12 MB read, 0.5 KB written "
                              "per launch of all 8 cells; algorithmic 32 B x R x N = %.2f MB)" % (32 * R * n / 1e6),
            "timing_note": "kernel_ms_per_launch / kernel_share_of_step come from a serialised pass (plain "
                           "launches, all cells in one launch, CUDA events on the library's stream); the timed "
                           "bench steps replay the step graph, whose replica-group branches overlap",
            "executed": {"pair_evaluations_per_launch": executed_pairs, "flop_per_pair_evaluation": FLOP_PER_PAIR,
                         "achieved": executed_pairs * FLOP_PER_PAIR / per_launch_s / 1e12, "unit": "TFLOP/s",
                         "frac": executed_pairs * FLOP_PER_PAIR / per_launch_s / 1e12 / fp32_peak if fp32_peak else None,
                         "note": "work the branch-free gather executes (27 cells of edge dcut, both directions "
                                 "of every pair) = %.1fx the algorithmic pair count" % (
                                     executed_pairs / max(useful * beads_per_launch, 1.0))},
            "kernel_ms_per_launch": per_launch_s * 1e3, "kernel_launches_timed": cnt,
            "kernel_share_of_step": ms / total_ms if total_ms else None,
            "algorithmic_flop_per_bead_step": FLOP_PER_PAIR * useful + FLOP_SINGLE_BEAD,
            "useful_pairs_per_bead": useful,
            "hbm": {"bound": "hbm", "achieved": hbm_ach, "peak": hbm_peak, "unit": "GB/s",
                    "frac": hbm_ach / hbm_peak,
                    "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback 6650 GB/s",
                    "algorithmic_bytes_per_bead_step": ALG_BYTES_PER_BEAD_STEP},
            "step_kernels_ms": {k: round(v[1] / args.profile_steps, 5) for k, v in prof.items()},
            "whole_step": {
                "fp32_frac": (FLOP_PER_PAIR * useful + FLOP_SINGLE_BEAD) * value / world / 1e12 / fp32_peak,
                "hbm_frac": ALG_BYTES_PER_BEAD_STEP * value / world / 1e9 / hbm_peak},
        }

    # ---- CPU baseline on the host cores (rank 0, N = 1 only) --------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        g.synchronize()
        rate, ms, kind, cores = cpu_reference_run(design, cfg, x_start, ctx0,
                                                  args.cpu_sample_steps, 1, 0)
        cpu = {"value": rate, "unit": "bead-steps/s", "cores": cores,
               "kind": "reference" if kind == "reference" else "port",
               "sample": f"{cores} independent single-threaded trajectories (one per host core) x "
                         f"{args.cpu_sample_steps} interphase steps from the same start state, "
                         f"driver callback incl. contact map; {ms / 1e3:.1f} s wall"}

    if rank == 0:
        line = {
            "metric": "bead-steps/s", "value": value, "unit": "bead-steps/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_total / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": workload_config(design, args, how),
            "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches),
            "roofline": roof, "cpu_baseline": cpu,
        }
        emit(line)
    g.close()
    if world > 1:
        dist.destroy_process_group()


_REAL_STDOUT = None


def emit(line):
    """The one JSON line goes to the real stdout; everything else printed while the
    bench runs (NCCL's version banner, library chatter) is diverted to stderr."""
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def main():
    global _REAL_STDOUT
    args = parse_args()
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_b200_arm(args)


if __name__ == "__main__":
    main()

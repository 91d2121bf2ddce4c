#!/usr/bin/env python
"""bench.py -- throughput of the force-evaluation + Brownian-integration hot path.

    python bench.py --gpus N --steps K --warmup W            # CUDA path (this repo)
    python bench.py --impl reference --gpus N --steps K ...  # the reference's CPU path

Workload (config.workload): the DIPLOID INTERPHASE STAGE of BASELINE.json
(configs[0]/[2] topology: 46 chains + nucleolar beads, 60,672 beads) as an
ensemble of R independent cell trajectories per GPU (weak scaling: 8 per GPU,
i.e. configs[2]'s 64 cells at 8 GPUs).  One bench "step" = S = 500 consecutive
simulation steps of every replica: cell-list rebuild, soft-core pair forces,
bonds, nucleolar droplet, ellipsoid wall + wall ODE, parameter schedule,
BAOAB-limit update with Philox noise, contact-map update every 20 steps.
metric = bead-steps/s (beads x replicas x simulation steps / second).

THE WINDOW.  Every arm works on the same piece of trajectory: the first 10^4
interphase steps from the relaxed t = 0 state (BASELINE configs[0]: "10k Brownian
steps").  Bench step k covers simulation steps [k S, (k + 1) S) of it (wrapping to
t = 0 after 10^4); warm-up steps run from t = 0 and the state is reset before the
timed region.  The pair density falls from 33.9 to 11.8 pairs/bead along the
window, so WHERE an arm is timed matters: the CPU arms, which cannot afford 10^4
steps, sample the window at the committed states of
tests/golden/interphase_window_60672.npz (steps 0, 2500, ..., 10^4; positions +
driver context; scripts/make_window_states.py) and their config block is the CUDA
arm's, verbatim.

value  : positions resident in HBM; CUDA events on the library's stream.
e2e    : the same through the C-ABI with HOST buffers: every bench step uploads
         the positions of all replicas (one async copy from page-locked memory),
         runs S steps, downloads positions, the context and the energy of every
         replica (what the reference driver does at sampling cadence).
r1     : the same window with ONE cell per GPU (BASELINE configs[0] as written).
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
import tempfile
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

N_BEADS_FULL = 60672
STATE_FIXTURE = ROOT / "tests" / "golden" / "interphase_relaxed_60672.npz"
WINDOW_FIXTURE = ROOT / "tests" / "golden" / "interphase_window_60672.npz"
TRAFFIC_FILE = ROOT / "profiles" / "pair_kernel_traffic.json"
WINDOW_STEPS = 10000                   # BASELINE configs[0]: 10k Brownian steps
ALG_BYTES_PER_BEAD_STEP = 32.0         # float4 in + float4 out
FLOP_PER_PAIR = 40.0                   # SURVEY.md 8(d), per unordered pair within the cutoff
FLOP_SINGLE_BEAD = 125.0               # interphase single-bead terms (k_finish, not the pair kernel)
CTX_FIELDS = ("time", "semiaxes", "core_scale", "bond_scale", "contact_distance")


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)      # 20 x 500 = the whole 10^4-step window
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--replicas", type=int, default=8, help="cell trajectories per GPU")
    ap.add_argument("--sim-steps", type=int, default=500, help="simulation steps per bench step")
    ap.add_argument("--scale", type=float, default=1.0, help="genome scale (tests only; 1 = diploid)")
    ap.add_argument("--relax-steps", type=int, default=2000,
                    help="adaptive relaxation steps when no cached state exists")
    ap.add_argument("--cpu-sample-steps", type=int, default=20,
                    help="simulation steps per core and per sampled state of the CPU legs")
    ap.add_argument("--cpu-cores", type=int, default=0, help="0 = every core this process may use")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-r1", action="store_true")
    ap.add_argument("--no-extra", action="store_true",
                    help="skip the force_100k (N = 1) / decomp_2m (N > 1) blocks")
    ap.add_argument("--extra-timeout", type=float, default=240.0,
                    help="seconds the decomp_2m block may take before the line is emitted without it")
    ap.add_argument("--profile-steps", type=int, default=60)
    ap.add_argument("--decomp-beads", type=int, default=2_000_000)
    # internal: one pinned CPU worker of the reference legs
    ap.add_argument("--cpu-worker", type=int, default=-1, help=argparse.SUPPRESS)
    ap.add_argument("--states", default="", help=argparse.SUPPRESS)
    ap.add_argument("--kind", default="", help=argparse.SUPPRESS)
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


def callback0_context(cfg):
    """Driver context after callback(0) (stages.interphase_callback0), as a dict."""
    semi = [a + cfg.timestep * cfg.wall_mobility * (0.0 - k * a)
            for a, k in zip(cfg.wall_semiaxes_init, cfg.wall_semiaxes_spring)]
    return dict(time=0.0, semiaxes=semi, core_scale=cfg.core_scale_init,
                bond_scale=cfg.bond_scale_init,
                contact_distance=cfg.contactmap_distance * cfg.core_scale_init)


class Window:
    """States along the benchmark window: .steps [m], .positions [m, n, 3], .contexts [m]
    (dicts of CTX_FIELDS), .pairs [m] (unordered pairs within the cutoff per bead)."""

    def __init__(self, steps, positions, contexts, pairs, source):
        self.steps = [int(s) for s in steps]
        self.positions = positions
        self.contexts = contexts
        self.pairs = [None if p is None else float(p) for p in pairs]
        self.source = source

    @staticmethod
    def committed(design):
        if design.n != N_BEADS_FULL or not WINDOW_FIXTURE.exists():
            return None
        z = np.load(WINDOW_FIXTURE)
        if z["positions"].shape[1:] != (design.n, 3):
            return None
        ctxs = [dict(time=float(c[0]), semiaxes=[float(c[1]), float(c[2]), float(c[3])],
                     core_scale=float(c[4]), bond_scale=float(c[5]), contact_distance=float(c[6]))
                for c in z["context"]]
        return Window(z["steps"], z["positions"], ctxs, z["pairs_per_bead"],
                      "tests/golden/interphase_window_60672.npz")

    def nearest(self, step):
        return int(np.argmin([abs(s - step) for s in self.steps]))

    def save(self, path):
        np.savez(path, steps=np.asarray(self.steps), positions=np.asarray(self.positions, np.float32),
                 context=np.asarray([[c["time"], *c["semiaxes"], c["core_scale"], c["bond_scale"],
                                      c["contact_distance"]] for c in self.contexts]))

    def describe(self):
        return {str(s): (None if p is None else round(p, 2)) for s, p in zip(self.steps, self.pairs)}


def window_plan(window, n_steps, sim_steps):
    """State index sampled for each of n_steps bench steps: the committed state nearest to
    the midpoint of the simulation steps that bench step covers."""
    return [window.nearest(((k * sim_steps) % WINDOW_STEPS) + sim_steps / 2) for k in range(n_steps)]


def workload_config(design, args, how, window):
    """ONE config block for both arms (the reference arm runs on the CUDA arm's config)."""
    return {
        "workload": "diploid interphase stage (BASELINE configs[0]/[2] topology), ensemble of "
                    "independent cell trajectories",
        "beads_per_cell": int(design.n), "chains": int(len(design.chains)),
        "cells_per_gpu": args.replicas, "sim_steps_per_bench_step": args.sim_steps,
        "timestep": 1e-5, "contact_map_interval": 20,
        "start_state": how,
        "window": f"interphase steps [0, {WINDOW_STEPS}) from the start state; bench step k covers "
                  f"[k S, (k + 1) S) mod {WINDOW_STEPS}; warm-up steps run from t = 0 and the state is "
                  "reset before the timed region",
        "pairs_per_bead_along_window": window.describe() if window else None,
        "window_states": window.source if window else "none (t = 0 only)",
        "l2": "flushed between bench steps (256 MiB memset inside the timed region); the "
              "per-step working set itself is L2-resident by design",
        "parallelism": f"ensemble x{args.gpus} (one process per GPU, no data-path collective); on each "
                       "GPU the step graph runs the cells as parallel branches (replica groups, "
                       f"GC_GROUPS={os.environ.get('GC_GROUPS', 'default: min(cells, 8)')})",
    }


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
# CPU reference legs (oracle/_ref = micromd compiled unmodified, else the port):
# one PINNED single-threaded process per core, each an independent trajectory --
# exactly how the reference uses more than one core (one `scripts/simulate -s seed`
# per cell).
# ---------------------------------------------------------------------------
def cpu_worker_main(args):
    """Child process: pins itself to one core, builds the interphase system on the CPU
    oracle, then serves `<state index> <sim steps>` requests from stdin; every request
    resets positions + driver context to that window state, runs the steps with the
    full driver callback (schedule, wall ODE, contact map) and prints the seconds."""
    core = args.cpu_worker
    try:
        os.sched_setaffinity(0, {core})
    except (AttributeError, OSError):
        pass
    sys.path.insert(0, str(ROOT / "oracle"))
    import oracle_py as O
    from gcpkg import gc_b200
    stages = gc_b200.stages
    design, cfg, _ = make_workload(args.scale)
    z = np.load(args.states)
    o = O.OracleSystem(design.n, args.kind)
    stages.setup_interphase(o, design, cfg)
    o.set_seeds([1000 + core])
    print("ready", flush=True)
    for line in sys.stdin:
        tok = line.split()
        if not tok or tok[0] == "quit":
            break
        k, sim_steps = int(tok[0]), int(tok[1])
        c = z["context"][k]
        o.set_positions(z["positions"][k].astype(np.float64))
        o.set_context(time=float(c[0]), semiaxes=[float(c[1]), float(c[2]), float(c[3])],
                      core_scale=float(c[4]), bond_scale=float(c[5]), contact_distance=float(c[6]))
        o.contacts_clear()
        t0 = time.perf_counter()
        o.run(sim_steps, cfg.timestep, interphase_callback=True, step_offset=int(z["steps"][k]),
              stepper=0)
        print(repr(time.perf_counter() - t0), flush=True)
    o.close()


class CpuPool:
    def __init__(self, args, window, cores):
        sys.path.insert(0, str(ROOT / "oracle"))
        import oracle_py as O
        if O.native_usable():
            self.kind, self.lib_note = "native", f"reference sources, -march=native ({O.native_arch()})"
        elif O.available("reference"):
            self.kind, self.lib_note = "reference", ("reference sources, -march=x86-64-v3 (the -march=native "
                                                      "build is absent or this host lacks its ISA flags)")
        else:
            O.build(("port",))
            self.kind, self.lib_note = "port", "oracle port (no compiled reference in this tree)"
        allowed = sorted(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else list(range(os.cpu_count() or 1))
        self.cores = allowed[:cores] if cores else allowed
        self.tmp = tempfile.NamedTemporaryFile(suffix=".npz", delete=False)
        self.tmp.close()
        window.save(self.tmp.name)
        self.procs = []
        for c in self.cores:
            cmd = [sys.executable, str(ROOT / "bench.py"), "--cpu-worker", str(c), "--states", self.tmp.name,
                   "--kind", self.kind, "--scale", str(args.scale)]
            self.procs.append(subprocess.Popen(cmd, stdin=subprocess.PIPE, stdout=subprocess.PIPE, text=True))
        for p in self.procs:
            line = p.stdout.readline()
            if line.strip() != "ready":
                raise RuntimeError(f"cpu worker failed to start: {line!r}")

    def run(self, state_index, sim_steps):
        """All workers run the same sample concurrently; returns the wall time (max)."""
        for p in self.procs:
            p.stdin.write(f"{state_index} {sim_steps}\n")
            p.stdin.flush()
        return max(float(p.stdout.readline()) for p in self.procs)

    def close(self):
        for p in self.procs:
            try:
                p.stdin.write("quit\n")
                p.stdin.flush()
            except OSError:
                pass
        for p in self.procs:
            p.wait(timeout=30)
        os.unlink(self.tmp.name)

    def cpu_model(self):
        try:
            for line in open("/proc/cpuinfo"):
                if line.startswith("model name"):
                    return line.split(":", 1)[1].strip()
        except OSError:
            pass
        return "unknown"

    @property
    def baseline_kind(self):
        return "port" if self.kind == "port" else "reference"


def cpu_sample(pool, design, plan, sim_steps):
    """Runs one sample per entry of `plan` (state indices); returns (rate, per-state
    rates, total seconds)."""
    per_state, total = {}, 0.0
    for k in plan:
        dt = pool.run(k, sim_steps)
        total += dt
        per_state.setdefault(k, []).append(len(pool.cores) * design.n * sim_steps / dt)
    rate = len(pool.cores) * design.n * sim_steps * len(plan) / total
    return rate, {k: float(np.mean(v)) for k, v in per_state.items()}, total


def fallback_window(design, cfg, x, how):
    return Window([0], np.asarray([x], dtype=np.float32), [callback0_context(cfg)], [None],
                  f"t = 0 only ({how})")


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    design, cfg, x0 = make_workload(args.scale)
    x, how = cached_state(design)
    if x is None:
        x, how = x0, "unrelaxed synthetic globule (no cached relaxed state)"
    window = Window.committed(design) or fallback_window(design, cfg, x, how)
    pool = CpuPool(args, window, args.cpu_cores)
    S = args.cpu_sample_steps
    # warm-up: short samples at t = 0 (pages, caches, branch predictors)
    for _ in range(args.warmup):
        pool.run(0, max(S // 5, 2))
    plan = window_plan(window, args.steps, args.sim_steps)
    rate, per_state, total = cpu_sample(pool, design, plan, S)
    cores = len(pool.cores)
    sample = (f"{cores} pinned single-threaded processes (one independent trajectory per core, "
              f"{pool.cpu_model()}), {S} interphase steps per bench step from the committed window state "
              f"nearest to that bench step (states used: {sorted(set(window.steps[k] for k in plan))}), "
              f"driver callback incl. contact map; {pool.lib_note}; {total:.1f} s wall")
    line = {
        "impl": "reference", "metric": "bead-steps/s", "value": rate, "unit": "bead-steps/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": total / max(args.steps, 1) * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": workload_config(design, args, how, window),
        "cpu_baseline": {"value": rate, "unit": "bead-steps/s", "cores": cores,
                         "kind": pool.baseline_kind, "sample": sample, "cpu_model": pool.cpu_model(),
                         "per_state": {str(window.steps[k]): v for k, v in per_state.items()}},
        "e2e": {"value": rate, "unit": "bead-steps/s", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    pool.close()
    emit(line)


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
    # cell c = a reference run `scripts/simulate -s (1000 + c)`
    g.set_seeds(ens.replica_seeds_reference(1000, cells))

    x, how = cached_state(design)
    if x is None:
        # relaxation stage on the device (run_relaxation semantics: adaptive dt,
        # scales frozen at their initial values, wall fixed)
        g.set_positions(x0)
        stages.begin_relaxation(g, cfg)
        g.run(args.relax_steps, cfg.timestep, spacestep=cfg.relaxation_spacestep)
        x = g.get_positions(0)
        how = f"synthetic globule + {args.relax_steps} adaptive relaxation steps on device"
    window = Window.committed(design)

    stream = torch.cuda.ExternalStream(g.stream(), device=local)
    flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device=f"cuda:{local}")
    done = [0]                                                   # simulation steps into the window

    def reset_to_t0(sys_=None):
        s = sys_ or g
        s.set_positions(x)                                       # every cell starts from the same state
        stages.begin_interphase(s, cfg)
        s.contacts_clear()
        done[0] = 0

    def bench_step():
        if done[0] + S > WINDOW_STEPS:
            reset_to_t0()
        with torch.cuda.stream(stream):
            flush_buf.zero_()
        g.run(S, cfg.timestep, restart=(done[0] == 0))
        done[0] += S
        # the driver dumps and clears the contact map once per output window
        # (simulation_driver_interphase.cpp:42-45); here once per bench step
        g.contacts_clear()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        g.synchronize()

    # ---- device-resident throughput ----------------------------------------
    reset_to_t0()
    for w in range(args.warmup):
        bench_step()
    reset_to_t0()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    l0 = g.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for k in range(args.steps):
        bench_step()
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
    ctx_end = g.get_context(0)

    # ---- end to end through the C-ABI with host buffers -------------------------
    e2e = None
    if not args.no_e2e:
        host = g.host_buffer()                                   # page-locked [R, n, 3] f32
        reset_to_t0()
        g.get_positions_all_f32(host)

        def e2e_pass(n_steps):
            """n_steps bench steps with host buffers, then one ensemble output window."""
            reset_to_t0()
            g.get_positions_all_f32(host)
            barrier()
            t0 = time.perf_counter()
            energy, ctxs = None, None
            for k in range(n_steps):
                if done[0] + S > WINDOW_STEPS:
                    reset_to_t0()
                    g.get_positions_all_f32(host)
                g.set_positions_all_f32(host)                    # H2D, all replicas, one copy
                g.run(S, cfg.timestep, restart=(done[0] == 0))
                done[0] += S
                if k + 1 < n_steps:
                    g.contacts_clear()
                g.get_positions_all_f32(host)                    # D2H, all replicas, one copy
                ctxs = g.get_contexts()                          # context of every cell
                energy = g.compute_energy_all() / n              # energy of every cell
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
            # P(s) on the device from this rank's sorted (key, count) runs: only the histogram
            # (8 KB) and three scalars cross NVLink
            hist = obs_dev.ps_runs(keys, counts, PS_MAX).astype(np.float64)
            tot = torch.tensor(np.concatenate([hist, [float(counts.sum().item()), float(keys.numel())]]),
                               dtype=torch.float64, device=f"cuda:{local}")
            tot = ens.reduce_sum(tot).cpu().numpy()
            tl = mark("ps_histogram_allreduce", tl)
            rg2 = np.stack([g.radius_of_gyration2(design.chains, r) for r in range(R)])
            obs = torch.tensor(np.concatenate([rg2.sum(axis=0), [float(R)]]), dtype=torch.float64,
                               device=f"cuda:{local}")
            obs = ens.reduce_sum(obs).cpu().numpy()
            tl = mark("rg2", tl)
            t_reduce = time.perf_counter() - t1
            dt = time.perf_counter() - t0
            # the merged ensemble MAP (what the reference's offline merge produces, src/cool.py:80-114)
            # is an output product, not a per-step metric: gathered and merged after the timed
            # region, timed on its own
            t2 = time.perf_counter()
            keys, counts = ens.gather_contacts(keys, counts,
                                               merge=lambda a, b: g.contacts_merge_device(a, b, local))
            ijv = ens.keys_to_ijv(keys, counts) if rank == 0 else None
            g.synchronize()
            t_map = time.perf_counter() - t2
            # compartment observables of the merged map on the device (rank 0): binned matrix,
            # PC1 by power iteration, A/B enrichment (src/pc1.py:21-125)
            comp[0] = None
            if rank == 0:
                try:
                    t3 = time.perf_counter()
                    binned.clear()
                    binned.add_runs(keys, counts)
                    pc = binned.pc1()
                    enr = binned.type_enrichment(bin_type, 2, min_separation=2)
                    comp[0] = {"bins": int(n_bins), "beads_per_bin": BIN, "pc1_explained_variance": float(pc["evr"]),
                               "pc1_vs_A_fraction_corr": float(abs(np.corrcoef(
                                   np.nan_to_num(pc["pc1"]), frac_a)[0, 1])),
                               "contacts_per_bin_pair": {"AA": float(enr[0, 0]), "AB": float(enr[0, 1]),
                                                         "BB": float(enr[1, 1])},
                               "ms_outside_timed_region": (time.perf_counter() - t3) * 1e3}
                except Exception as ex:            # an output product: never take the bench down
                    comp[0] = {"error": repr(ex)[:200]}
            g.contacts_clear()
            barrier()
            tt = torch.tensor([dt], dtype=torch.float64, device=f"cuda:{local}")
            if world > 1:
                dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            dt = float(tt.item())
            return dt, t_reduce, lap, energy, ctxs, obs, ijv, tot, t_map

        PS_MAX = 1000
        bead_chain = np.full(n, -1, dtype=np.int32)
        for c_, (s_, e_) in enumerate(design.chains):
            bead_chain[s_:e_] = c_
        obs_dev = gc_b200.Observables(len(design.chains), [(k, k + 1) for k in range(len(design.chains))], local)
        obs_dev.set_bead_map(np.maximum(bead_chain, 0), bead_chain)
        # bins of BIN beads inside every chain (nucleolar beads dropped), typed A / B by majority
        BIN, n_bins, ranges_b = 50, 0, []
        bead_bin = np.full(n, -1, dtype=np.int32)
        for s_, e_ in design.chains:
            k_ = -(-(e_ - s_) // BIN)
            bead_bin[s_:e_] = n_bins + (np.arange(e_ - s_) // BIN)
            ranges_b.append((n_bins, n_bins + k_))
            n_bins += k_
        ok_ = bead_bin >= 0
        frac_a = np.bincount(bead_bin[ok_], weights=(design.a > design.b)[ok_].astype(np.float64), minlength=n_bins) / \
            np.maximum(np.bincount(bead_bin[ok_], minlength=n_bins), 1)
        bin_type = np.where(frac_a > 0.5, 0, 1).astype(np.int32)
        binned = gc_b200.Observables(n_bins, ranges_b, local) if rank == 0 else None
        if binned is not None:
            binned.set_bead_map(bead_bin, bead_chain)
        comp = [None]
        e2e_pass(1)                       # untimed: page-locked staging buffers, NCCL channels
        dt, t_reduce, lap, energy, ctxs, obs, ijv, tot, t_map = e2e_pass(args.steps)
        obs_dev.close()
        if binned is not None:
            binned.close()
        e2e = {"value": world * R * n * S * args.steps / dt, "unit": "bead-steps/s",
               "h2d_bytes_per_step": R * n * 12,
               "d2h_bytes_per_step": R * n * 12 + R * 14 * 8 + R * 8,
               "copies": "one async H2D + one D2H of all replicas (page-locked host buffer, device-side "
                         "float3<->float4 repack), contexts of all replicas in one copy, energies of all "
                         "replicas in one copy",
               "ms_per_step": dt / args.steps * 1e3,
               "energy_per_bead": [float(e) for e in energy],
               "semiaxis_x": [float(c.semiaxes[0]) for c in ctxs]}
        if rank == 0:
            updates = S // cfg.contactmap_update_interval
            lengths = design.chains[:, 1] - design.chains[:, 0]
            avail = np.zeros(PS_MAX + 1)
            for L_ in lengths:
                sep_ = np.arange(1, min(L_ - 1, PS_MAX) + 1)
                avail[sep_] += L_ - sep_
            with np.errstate(divide="ignore", invalid="ignore"):
                ps = np.where(avail > 0, tot[:PS_MAX + 1] / (avail * max(updates, 1) * max(world * R, 1)), 0.0)
            # the same from the merged map on the host: the two must agree exactly
            ps_host = ens.contact_probability(ijv, design.chains, updates, world * R, max_s=PS_MAX)
            e2e["ensemble"] = {
                "cells": world * R, "contact_pairs": int(len(ijv)),
                "contact_counts": int(tot[PS_MAX + 1]),
                "reduce_ms": t_reduce * 1e3, "reduce_breakdown_ms": {k: round(v, 3) for k, v in lap.items()},
                "collective": "inside the timed region: all_reduce of the device-side P(s) histogram (8 KB) and of "
                              "Rg^2; after it: nccl all_gather + device merge of the (key, count) runs into the "
                              "ensemble contact map" if world > 1 else "none (1 rank)",
                "merged_map_ms_outside_timed_region": t_map * 1e3,
                "rg_mean": float(np.sqrt(obs[:-1] / obs[-1]).mean()),
                "P_s": {str(sep): float(ps[sep]) for sep in (1, 2, 5, 10, 30, 100, 300, 1000)},
                "P_s_matches_merged_map": bool(np.array_equal(ps[1:], ps_host[1:])),
                "compartments": comp[0],
            }

    # ---- one cell per GPU over the whole window (configs[0] as written) -----------
    r1 = None
    if rank == 0 and not args.no_r1:
        g1 = gc_b200.System(n, 1, local)
        stages.setup_interphase(g1, design, cfg)
        g1.set_seeds(ens.replica_seeds_reference(1000, [cells[0]]))
        s1 = torch.cuda.ExternalStream(g1.stream(), device=local)
        reset_to_t0(g1)
        g1.run(2 * S, cfg.timestep, restart=True)                # warm-up: graph capture, caches
        reset_to_t0(g1)
        g1.synchronize()
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a0.record(s1)
        nblk = WINDOW_STEPS // S
        for k in range(nblk):
            g1.run(S, cfg.timestep, restart=(k == 0))
            g1.contacts_clear()
        a1.record(s1)
        g1.synchronize()
        torch.cuda.synchronize()
        ms1 = a0.elapsed_time(a1)
        r1 = {"cells": 1, "sim_steps": nblk * S, "value": n * nblk * S / (ms1 * 1e-3),
              "unit": "bead-steps/s", "us_per_sim_step": ms1 * 1e3 / (nblk * S),
              "window": f"the whole window, {nblk * S} steps from t = 0, device-resident"}
        g1.close()
        done[0] = 0

    # ---- roofline of the dominant kernel: separate short pass with per-kernel events,
    # at the window's MIDDLE state (all cells set to it) ---------------------------------
    roof = None
    if rank == 0:
        args.profile_steps = max(args.profile_steps, 1)
        if window is not None:
            km = window.nearest(WINDOW_STEPS / 2)
            g.set_positions(window.positions[km].astype(np.float64))
            g.run(0, cfg.timestep, restart=True)                 # new noise stream, Wiener history reset
            g.set_context(**window.contexts[km])
            g.set_step(window.steps[km])                         # resume the schedule at that step
            prof_state = f"committed window state at step {window.steps[km]} + 64 steps"
        else:
            reset_to_t0()
            g.run(S, cfg.timestep, restart=True)
            prof_state = f"t = 0 + {S + 64} steps"
        g.contacts_clear()
        g.run(64, cfg.timestep, restart=False)                   # decorrelate the identical cells a little
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
        # unordered pairs within dcut per bead, mean over the cells of this launch
        useful = float(np.mean([len(g.neighbor_pairs(dcut, replica=r)) for r in range(R)])) / n
        # pair evaluations a 27-cell gather EXECUTES: every home bead against every bead of
        # its 27 cells of edge dcut (before any culling; both directions of every pair)
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
        gathered_pairs = float((occ * cand).sum()) * R
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
        is_pair = name.startswith("k_pair")
        # the pair kernel's OWN algorithmic work: 40 flop per unordered pair within the cutoff
        # (the 125 single-bead flops belong to k_finish and count only in whole_step)
        flop_alg = FLOP_PER_PAIR * useful * beads_per_launch
        fp32_ach = flop_alg / per_launch_s / 1e12
        hbm_ach = ALG_BYTES_PER_BEAD_STEP * beads_per_launch / per_launch_s / 1e9
        traffic, traffic_src = None, "not captured for this build"
        if TRAFFIC_FILE.exists():
            tf = json.loads(TRAFFIC_FILE.read_text())
            if tf.get("kernel") == name.split("<")[0] and tf.get("cells") == R and args.scale == 1.0:
                traffic = tf.get("dram_bytes_per_launch")
                traffic_src = (f"NOT measured in this run: {tf.get('source')} (commit {tf.get('commit')}), "
                               f"dram__bytes_read.sum + dram__bytes_write.sum of one launch")
        roof = {
            "kernel": name, "bound": "fp32", "achieved": fp32_ach, "peak": fp32_peak,
            "unit": "TFLOP/s", "frac": fp32_ach / fp32_peak if fp32_peak else None,
            "formula": "achieved = 40 flop x useful_pairs_per_bead x beads_per_launch / kernel_s_per_launch"
                       if is_pair else "dominant kernel is not the pair kernel: see step_kernels_ms",
            "peak_source": "measured live: FFMA / FFMA2 chain microbenchmarks (gc_measure_fp32_peak), max of "
                           f"scalar {fp32_scalar:.2f} and packed {fp32_packed:.2f} TFLOP/s",
            "traffic": traffic, "traffic_source": traffic_src,
            "algorithmic_bytes_per_launch": ALG_BYTES_PER_BEAD_STEP * beads_per_launch,
            "state": prof_state,
            "timing_note": "kernel_ms_per_launch / kernel_share_of_step come from a serialised pass (plain "
                           "launches, all cells in one launch, CUDA events on the library's stream); the timed "
                           "bench steps replay the step graph, whose replica-group branches overlap",
            "gathered": {"pair_evaluations_per_launch": gathered_pairs,
                         "flop_per_pair_evaluation": FLOP_PER_PAIR,
                         "achieved": gathered_pairs * FLOP_PER_PAIR / per_launch_s / 1e12, "unit": "TFLOP/s",
                         "frac": gathered_pairs * FLOP_PER_PAIR / per_launch_s / 1e12 / fp32_peak if fp32_peak else None,
                         "note": "a full 27-cell gather (edge dcut, both directions of every pair, before "
                                 "culling) = %.1fx the algorithmic pair count" % (
                                     gathered_pairs / max(useful * beads_per_launch, 1.0))},
            "kernel_ms_per_launch": per_launch_s * 1e3, "kernel_launches_timed": cnt,
            "kernel_share_of_step": ms / total_ms if total_ms else None,
            "beads_per_launch": beads_per_launch,
            "useful_pairs_per_bead": useful,
            "algorithmic_flop_per_bead_step": {"pair_kernel": FLOP_PER_PAIR * useful,
                                               "whole_step": FLOP_PER_PAIR * useful + FLOP_SINGLE_BEAD},
            "hbm": {"bound": "hbm", "achieved": hbm_ach, "peak": hbm_peak, "unit": "GB/s",
                    "frac": hbm_ach / hbm_peak,
                    "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback 6650 GB/s",
                    "algorithmic_bytes_per_bead_step": ALG_BYTES_PER_BEAD_STEP},
            "step_kernels_ms": {k: round(v[1] / args.profile_steps, 5) for k, v in prof.items()},
            "step_kernels_ms_sum": round(total_ms / args.profile_steps, 5),
        }
        # whole step over the timed window: mean useful pairs along it (committed states)
        if window is not None:
            pw = [p for s_, p in zip(window.steps, window.pairs) if s_ <= max(S * args.steps, 1) or s_ == 0]
            p_mean = float(np.mean(pw))
        else:
            p_mean = useful
        roof["whole_step"] = {
            "mean_useful_pairs_per_bead_over_timed_window": p_mean,
            "fp32_frac": (FLOP_PER_PAIR * p_mean + FLOP_SINGLE_BEAD) * value / world / 1e12 / fp32_peak,
            "hbm_frac": ALG_BYTES_PER_BEAD_STEP * value / world / 1e9 / hbm_peak}

    def headline():
        return {
            "metric": "bead-steps/s", "value": value, "unit": "bead-steps/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_total / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": workload_config(design, args, how, window),
            "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches),
            "roofline": roof, "cpu_baseline": None, "r1": r1,
            "window_end": {"step": int(ctx_end.step), "semiaxis_x": float(ctx_end.semiaxes[0]),
                           "core_scale": float(ctx_end.core_scale)},
        }

    # ---- configs[3] on one GPU / configs[4] over the GPUs of this run -------------------
    extra = {}
    if not args.no_extra and args.scale == 1.0:
        g.close()
        g = None
        del flush_buf
        torch.cuda.empty_cache()
        if world == 1:
            try:
                extra["force_100k"] = force_100k_block(gc_b200, local)
            except Exception as ex:                       # the headline numbers must still get out
                extra["force_100k"] = {"error": repr(ex)[:300]}
        else:
            # The decomposed polymer is a second, collective workload behind the ensemble timing:
            # a rank that fails or stalls in it must not take the line above with it.  A watchdog
            # thread emits the line without the block and ends the process when it takes too long.
            def bail(why):
                if rank == 0:
                    line = headline()
                    line["decomp_2m"] = {"error": why}
                    emit(line)
                sys.stderr.flush()
                os._exit(0)

            wd = threading.Timer(args.extra_timeout, bail, args=(f"no result after {args.extra_timeout} s",))
            wd.daemon = True
            wd.start()
            try:
                extra["decomp_2m"] = decomp_block(gc_b200, args, rank, world, local)
            except Exception as ex:
                wd.cancel()
                print(f"decomp_2m failed on rank {rank}: {ex!r}", file=sys.stderr)
                bail(repr(ex)[:300])                      # the other ranks leave through their watchdogs
            wd.cancel()

    # ---- CPU baseline on the host cores (rank 0, N = 1 only) --------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        if g is not None:
            g.synchronize()
        w = window or fallback_window(design, cfg, x, how)
        pool = CpuPool(args, w, args.cpu_cores)
        span = min(S * args.steps, WINDOW_STEPS)
        plan = sorted(set(w.nearest(t_) for t_ in (0, span / 2, span)))
        rate, per_state, total = cpu_sample(pool, design, plan, args.cpu_sample_steps)
        cpu = {"value": rate, "unit": "bead-steps/s", "cores": len(pool.cores),
               "kind": pool.baseline_kind, "cpu_model": pool.cpu_model(),
               "per_state": {str(w.steps[k]): v for k, v in per_state.items()},
               "sample": f"{len(pool.cores)} pinned single-threaded processes (one independent trajectory per "
                         f"core) x {args.cpu_sample_steps} interphase steps from each of the window states "
                         f"{[w.steps[k] for k in plan]} (start / middle / end of the timed window), driver "
                         f"callback incl. contact map; {pool.lib_note}; {total:.1f} s wall"}
        pool.close()

    if rank == 0:
        line = headline()
        line["cpu_baseline"] = cpu
        line.update(extra)
        emit(line)
    if g is not None:
        g.close()
    if world > 1:
        dist.destroy_process_group()


def force_100k_block(gc_b200, local):
    """BASELINE configs[3]: force-only evaluation (pair + bonds + wall) of a 100k-bead
    polymer in an ellipsoid, per-kernel CUDA events; cell build reported separately."""
    synth, stages = gc_b200.synth, gc_b200.stages
    semi = (5.0, 4.0, 3.0)
    cfg = synth.InterphaseConfig(wall_semiaxes_init=semi, nucleolus_droplet_energy=0.0)
    design, x = synth.make_polymer(100000, seed=4, semiaxes=semi)
    g = gc_b200.System(design.n, 1, local)
    stages.setup_interphase(g, design, cfg, wall_ode=False, schedule=False, contacts=False)
    g.set_context(core_scale=1.0, bond_scale=1.0)
    g.set_positions(x)
    evals = 100
    for _ in range(5):
        g.compute_force()
    g.profile_enable(True)
    for _ in range(evals):
        g.compute_force()
    prof = g.profile_read()
    g.profile_enable(False)
    kern = {k: v[1] / evals for k, v in prof.items()}
    build = sum(v for k, v in kern.items() if k in ("k_reset_reductions", "k_bbox", "k_plan", "k_count",
                                                    "k_scan", "k_fill", "k_gather", "k_rebuild_small"))
    ms = sum(kern.values())
    pairs = len(g.neighbor_pairs(max(cfg.a_core_diameter, cfg.b_core_diameter))) / design.n
    g.close()
    return {"workload": f"force-only (pair + chain bonds + both wall passes), {design.n} beads, semiaxes {semi}, "
                        "core_scale 1, seed 4", "evaluations": evals,
            "ms_per_evaluation": ms, "ms_cell_build": build, "ms_force_kernels": ms - build,
            "bead_evaluations_per_s": design.n / (ms * 1e-3),
            "kernels_ms": {k: round(v, 5) for k, v in kern.items()},
            "useful_pairs_per_bead": pairs,
            "timing": "CUDA events around every kernel on the library's stream (serialised)",
            "parity": "tests/test_gpu_parity.py::test_force_only_100k_polymer (total force <= 1e-5, pair set bit-exact)"}


def decomp_block(gc_b200, args, rank, world, local):
    """BASELINE configs[4]: one 2M-bead polymer over the GPUs of this run."""
    import torch
    import torch.distributed as dist
    synth, stages, decomp = gc_b200.synth, gc_b200.stages, gc_b200.decomp
    beads = args.decomp_beads
    semi = tuple(s * (beads / 1e5) ** (1 / 3) for s in (5.0, 4.0, 3.0))
    cfg = synth.InterphaseConfig(wall_semiaxes_init=semi, nucleolus_droplet_energy=0.0,
                                 wall_semiaxes_spring=(3e4 * beads / 60672,) * 3)
    design, x = synth.make_polymer(beads, seed=4, semiaxes=semi)
    steps, warm = 100, 20

    def make():
        s = gc_b200.System(design.n, 1, local)
        stages.setup_interphase(s, design, cfg, contacts=False)
        s.set_positions(x)
        s.set_seeds([4])
        stages.interphase_callback0(s, cfg)
        return s

    g = make()
    # the walk's bonds (0.3) span two cell edges (0.15): a three-layer position halo; halo
    # buffers sized for the beads of ~half a slab (the library reports an overflow)
    # halo through NVLink peer memory (CUDA IPC) when the box allows it, NCCL send/recv otherwise
    kw = dict(halo_layers=3, halo_capacity=max(1024, int(0.6 * beads / world)), graph_steps=10)
    try:
        if os.environ.get("GC_DECOMP_PEER", "1") == "0":
            raise RuntimeError("disabled by GC_DECOMP_PEER=0")
        run = decomp.DecomposedRun(g, rank, world, local, peer=True, **kw)
    except RuntimeError as ex:
        print(f"decomp_2m: peer mode off ({ex})", file=sys.stderr)
        run = decomp.DecomposedRun(g, rank, world, local, peer=False, **kw)
    run.begin(cfg.timestep)
    run.step(warm // 2)
    cap = run.tune_capacity()
    run.step(warm - warm // 2)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    dist.barrier()
    torch.cuda.synchronize()
    e0.record(run.stream)
    run.step(steps)
    e1.record(run.stream)
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=f"cuda:{local}")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    g.synchronize()
    run.invalidate_graph()          # a live graph holds NCCL kernels
    x_dec = run.gather_positions()
    counts = torch.tensor(list(run.halo_counts()) + [int(g.decomp_owned().sum())], device=f"cuda:{local}")
    allc = [torch.zeros_like(counts) for _ in range(world)]
    dist.all_gather(allc, counts)
    exch = int(run.exchange_bytes_per_step)
    scheme = run.scheme
    run.close()
    g.close()
    out = None
    if rank == 0:
        g1 = make()
        s1 = torch.cuda.ExternalStream(g1.stream(), device=local)
        g1.run(warm, cfg.timestep, restart=True)
        torch.cuda.synchronize()
        e0.record(s1)
        g1.run(steps, cfg.timestep, restart=False)
        e1.record(s1)
        torch.cuda.synchronize()
        ms1 = e0.elapsed_time(e1)
        same = bool(np.array_equal(g1.get_positions(0), x_dec))
        g1.close()
        out = {"workload": f"1 chain x {design.n} beads, semiaxes {tuple(round(s, 2) for s in semi)}, interphase "
                           "force set, fixed dt, wall ODE + expansion schedule on", "gpus": world,
               "scheme": scheme, "steps": steps, "warmup": warm,
               "ms_per_step": ms / steps, "bead_steps_per_s": design.n * steps / (ms * 1e-3),
               "single_gpu_ms_per_step": ms1 / steps, "speedup_vs_single_gpu": ms1 / ms,
               "exchange_bytes_per_rank_step": exch, "halo_layers": 3, "halo_capacity_beads": cap,
               "halo_beads_lo_hi_and_owned_per_rank": [c.tolist() for c in allc],
               "bit_identical_to_single_gpu": same,
               "graph_steps": 10,
               "timing": "CUDA events on the library's stream, max over ranks"}
    dist.barrier()
    return out


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
    if args.cpu_worker >= 0:
        cpu_worker_main(args)
        return
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_b200_arm(args)


if __name__ == "__main__":
    main()

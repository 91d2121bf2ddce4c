"""One cell through the whole cycle on one GPU (BASELINE.json configs[1]), in the
reference's stage order (scripts/simulate:44-47 plus the prometaphase / cycle programs):

  anaphase -> telophase -> [spline refinement] -> relaxation -> interphase
           -> [coarse-graining + sister copy] -> prometaphase -> [daughter cell]

"Prophase condensation" and "G1 decondensation" have no dynamics of their own in the
reference: condensation is the instantaneous coarse-graining, decondensation is the
spline refinement followed by the core/bond-scale expansion schedule.
Step counts default to the reference's (SIM/common/simulation_config.hpp:19-24,65-72);
`scale_steps` shrinks them proportionally for tests.
"""
from __future__ import annotations

import os
import time

import numpy as np

from . import prepare, stages, synth, transitions
from .api import System


def run_cell_cycle(design=None, icfg=None, mcfg=None, seed=1, device=0, scale_steps=1.0,
                   log=None, frame_sink=None, host_transitions=False):
    """frame_sink(system, frame): optional consumer of every sampled frame (e.g. a
    store.TrajectoryWriter); host_transitions=True uses transitions.py (numpy / scipy on the
    host) instead of the gc_transition_* kernels (A/B and tests)."""
    icfg = icfg or synth.InterphaseConfig()
    mcfg = mcfg or synth.MitoticConfig()
    design = design or synth.make_diploid_design(cfg=icfg)
    md_ana = synth.make_mitotic_design(design, mcfg)
    md_pro = synth.make_mitotic_design(design, mcfg, sisters=True)
    nsteps = lambda k: max(int(round(k * scale_steps)), 1)   # noqa: E731
    report = {"beads_interphase": int(design.n), "beads_anaphase": int(md_ana.n),
              "beads_prometaphase": int(md_pro.n),
              "stages": []}

    frames = {"count": 0, "bytes": 0}
    # the 64-bit seed of every simulate call, derived from the master seed exactly as the
    # reference derives them (prepare.py: std::seed_seq -> per-stage mt19937_64)
    seeds = prepare.run_seeds(seed)
    report["seeds"] = {k: int(v) for k, v in seeds.items()}

    def run_sampled(sys_, n, dt, interval, on_chunk=None):
        """n steps in chunks of the sampling interval; after each chunk the driver's
        save_positions: one quantised f32 frame leaves the device.  The frame is started
        asynchronously (gc_frame_begin: quantise + copy on a second stream) and collected
        after the NEXT chunk has been queued, so its transfer overlaps that chunk's steps."""
        done = 0
        pending = False

        def collect():
            f = sys_.frame_end()
            frames["count"] += 1
            frames["bytes"] += f.nbytes
            if frame_sink is not None:
                frame_sink(sys_, f)

        if os.environ.get("GC_CYCLE_NO_FRAMES"):        # A/B: no frame output, window-sized chunks
            interval = max(interval, 10 * interval)
        while done < n:
            chunk = min(interval, n - done)
            sys_.run(chunk, dt, restart=(done == 0))
            done += chunk
            if pending:
                collect()
                pending = False
            if done % interval == 0:
                sys_.frame_begin(0, 16)
                pending = True
            if on_chunk:
                on_chunk(done)
        if pending:
            collect()

    def timed(name, sys_, n_beads, steps, fn):
        if frame_sink is not None and hasattr(frame_sink, "begin_stage"):
            frame_sink.begin_stage(name)
        sys_.synchronize()
        t0 = time.perf_counter()
        fn()
        sys_.synchronize()
        dt = time.perf_counter() - t0
        e = sys_.compute_energy() / n_beads
        n_beads = int(n_beads)
        row = {"stage": name, "beads": n_beads, "steps": int(steps), "seconds": dt,
               "bead_steps_per_s": n_beads * int(steps) / dt, "energy_per_bead": float(e)}
        report["stages"].append(row)
        if log:
            log(row)

    # ---- anaphase / telophase (SIM/stage_anatelophase/simulation_driver.cpp:192-310) ----
    ga = System(md_ana.n, 1, device)
    ta = stages.setup_anatelophase(ga, md_ana, mcfg)
    ga.set_seeds([seeds["anaphase"]])
    ga.set_positions(synth.rod_positions(md_ana, mcfg, seed=seed))
    stages.enter_anaphase(ga, ta)
    n = nsteps(mcfg.anaphase_steps)
    timed("anaphase", ga, md_ana.n, n, lambda: run_sampled(ga, n, mcfg.timestep, mcfg.sampling_interval))
    stages.enter_telophase(ga, ta)
    ga.set_seeds([seeds["telophase"]])
    n = nsteps(mcfg.telophase_steps)
    timed("telophase", ga, md_ana.n, n, lambda: run_sampled(ga, n, mcfg.timestep, mcfg.sampling_interval))
    # ---- relaxation / interphase (SIM/stage_interphase/simulation_driver*.cpp) -----------
    # transition interphase: spline refinement of the telophase chains, on the device
    t0 = time.perf_counter()
    gi = System(design.n, 1, device)
    stages.setup_interphase(gi, design, icfg)
    gi.set_seeds([seeds["relaxation"]])
    if host_transitions:
        x_telo = synth.quantize_f32(ga.get_positions()).astype(np.float64)
        x0 = synth.quantize_f32(transitions.transition_interphase(md_ana, design, x_telo)).astype(np.float64)
        gi.set_positions(x0)
    else:
        gi.transition_interphase_from(ga, md_ana.chains, design.chains, design.nucleolar_bonds, bits=16)
    ga.close()
    report["transition_interphase_s"] = time.perf_counter() - t0
    stages.begin_relaxation(gi, icfg)
    n = nsteps(icfg.relaxation_steps)
    timed("relaxation", gi, design.n, n,
          lambda: gi.run(n, icfg.timestep, spacestep=icfg.relaxation_spacestep))
    stages.begin_interphase(gi, icfg)
    gi.set_seeds([seeds["interphase"]])
    n = nsteps(icfg.steps)
    window = icfg.sampling_interval * icfg.contactmap_output_window

    def window_end(done):
        if done % window == 0:              # save_contacts + clear (simulation_driver_interphase.cpp:42-45)
            report["contact_pairs_last_window"] = int(len(gi.contacts_fetch()))
            gi.contacts_clear()

    def interphase():
        run_sampled(gi, n, icfg.timestep, icfg.sampling_interval, window_end)
    timed("interphase", gi, design.n, n, interphase)
    ctx = gi.get_context()
    report["interphase_end"] = {"time": ctx.time, "semiaxes": list(ctx.semiaxes),
                                "core_scale": ctx.core_scale}

    # ---- prometaphase (SIM/stage_prometaphase/simulation_driver.cpp:185-237) ---------------
    # transition prometaphase: x100 centroid coarse-graining + displaced sister copy
    t0 = time.perf_counter()
    gp = System(md_pro.n, 1, device)
    stages.setup_prometaphase(gp, md_pro, mcfg)
    gp.set_seeds([seeds["prometaphase"]])
    axis = np.asarray(mcfg.spindle_axis, dtype=np.float64)
    if host_transitions:
        x_inter = synth.quantize_f32(gi.get_positions()).astype(np.float64)
        xp = synth.quantize_f32(transitions.transition_prometaphase(design, md_pro, mcfg, x_inter))
        gp.set_positions(xp.astype(np.float64))
    else:
        gp.transition_prometaphase_from(gi, design.chains, md_pro.chains[md_pro.sisters[:, 0]],
                                        md_pro.chains[md_pro.sisters[:, 1]], mcfg.coarse_graining,
                                        -mcfg.sister_separation * axis / np.linalg.norm(axis), bits=16)
    gi.close()
    report["transition_prometaphase_s"] = time.perf_counter() - t0
    n = nsteps(mcfg.prometaphase_steps)
    timed("prometaphase", gp, md_pro.n, n, lambda: run_sampled(gp, n, mcfg.timestep, mcfg.sampling_interval))
    # transition cycle: the target chromatids start the daughter cell's anaphase
    gn = System(md_ana.n, 1, device)
    if host_transitions:
        x_meta = synth.quantize_f32(gp.get_positions()).astype(np.float64)
        gn.set_positions(synth.quantize_f32(transitions.transition_cycle(md_pro, md_ana, mcfg, x_meta)).astype(np.float64))
    else:
        gn.transition_cycle_from(gp, md_pro.chains[md_pro.sisters[:, 0]], md_ana.chains, -axis, bits=16)
    gp.close()
    report["frames_saved"] = frames["count"]
    report["frame_bytes"] = frames["bytes"]
    report["next_anaphase_start"] = gn.get_positions()
    gn.close()
    total_bs = sum(int(r["beads"]) * int(r["steps"]) for r in report["stages"])
    total_s = sum(r["seconds"] for r in report["stages"])
    report["total_seconds"] = total_s
    report["bead_steps_per_s"] = total_bs / total_s
    return report

"""The seeded cases behind tests/golden/reference_vectors.npz.  One function builds a
case on ANY system object (reference oracle, port oracle, CUDA System) and returns its
outputs, so the generator (scripts/make_golden.py) and the checks (tests/test_golden.py)
cannot drift apart."""
import numpy as np

from gcpkg import gc_b200

synth, stages = gc_b200.synth, gc_b200.stages

CASES = ["interphase_forces", "interphase_trajectory", "relaxation_step", "contacts", "pairs",
         "anatelophase", "prometaphase"]


def _f32(x):
    return np.asarray(x, dtype=np.float32).astype(np.float64)


def _interphase(scale=0.012, radius=0.5, seed=21):
    cfg = synth.InterphaseConfig()
    d = synth.make_diploid_design(scale=scale, cfg=cfg)
    x = synth.globule_positions(d, seed=seed, radius=radius)
    return d, cfg, x


def _run(s, is_oracle, steps, dt, noise, **kw):
    if is_oracle:
        return s.run(steps, dt, noise=noise, **kw)
    kw.pop("interphase_callback", None)
    kw.pop("want_ctx", None)
    return s.run(steps, dt, noise=noise, **kw)


def build_case(name, make_sys, is_oracle):
    rng = np.random.default_rng(CASES.index(name) + 100)
    out = {}
    if name == "interphase_forces":
        d, cfg, x = _interphase()
        x = x.copy()
        x[::41] *= 1.9                                   # some beads beyond the wall
        x = _f32(x)
        s = make_sys(d.n)
        terms = stages.setup_interphase(s, d, cfg)
        s.set_positions(x)
        s.set_context(core_scale=0.73, bond_scale=0.61, semiaxes=(0.6, 0.55, 0.65),
                      contact_distance=0.1)
        out["x"] = x
        for tname, t in terms.items():
            out[f"force_{tname}"] = s.compute_force(1 << t)
            out[f"energy_{tname}"] = s.compute_energy(1 << t)
        out["energy_total"] = s.compute_energy()
        out["force_total"] = s.compute_force()
        c = s.get_context()                               # wall statistics of that force call
        out["axial_reaction"] = list(c.axial_reaction)
        out["reaction_force"] = c.reaction_force
    elif name == "interphase_trajectory":
        d, cfg, x = _interphase()
        steps = 25
        z = rng.normal(size=(steps, d.n, 3))
        s = make_sys(d.n)
        stages.setup_interphase(s, d, cfg)
        s.set_positions(x)
        stages.interphase_callback0(s, cfg)
        _run(s, is_oracle, steps, cfg.timestep, z, interphase_callback=True)
        c = s.get_context()
        out["x_final"] = s.get_positions()
        out["ctx"] = [c.time, *c.semiaxes, c.core_scale, c.bond_scale]
        out["contacts"] = s.contacts_fetch()
    elif name == "relaxation_step":
        d, cfg, x = _interphase()
        z = rng.normal(size=(1, d.n, 3))
        s = make_sys(d.n)
        stages.setup_interphase(s, d, cfg)
        s.set_positions(x)
        stages.begin_relaxation(s, cfg)
        if is_oracle:
            o = s.run(1, cfg.timestep, spacestep=cfg.relaxation_spacestep, noise=z, want_ctx=True)
            out["dt"] = o["ctx"][0, 6]
        else:
            s.run(1, cfg.timestep, spacestep=cfg.relaxation_spacestep, noise=z)
            out["dt"] = s.get_context().last_timestep
        out["x_final"] = s.get_positions()
    elif name == "contacts":
        d, cfg, x = _interphase()
        s = make_sys(d.n)
        stages.setup_interphase(s, d, cfg)
        s.set_positions(x)
        s.set_context(contact_distance=0.12)
        s.contacts_update()
        s.contacts_update()
        out["triples"] = s.contacts_fetch()
    elif name == "pairs":
        x = _f32(rng.normal(size=(400, 3)) * 0.6)
        s = make_sys(400)
        s.set_bead_attrs(np.ones(400))
        s.set_positions(x)
        out["x"] = x
        for r in (0.1, 0.25, 0.4):
            out[f"pairs_{r}"] = s.neighbor_pairs(r)
    elif name in ("anatelophase", "prometaphase"):
        d = synth.make_diploid_design(scale=0.3)
        mcfg = synth.MitoticConfig(polar_ejection_force=0.7, polar_ejection_cross_section=0.04,
                                   kfiber_length_anaphase=0.2, kfiber_length_prometaphase=0.3,
                                   telophase_bond_spring_multiplier=0.5,
                                   telophase_bending_energy_multiplier=2.0)
        if name == "anatelophase":
            md = synth.make_mitotic_design(d, mcfg)
            x = _f32(synth.rod_positions(md, mcfg) + rng.normal(size=(md.n, 3)) * 0.05)
            s = make_sys(md.n)
            t = stages.setup_anatelophase(s, md, mcfg)
            s.set_positions(x)
            stages.enter_anaphase(s, t)
            out["force_anaphase"] = s.compute_force()
            out["energy_anaphase"] = s.compute_energy()
            steps = 20
            z = rng.normal(size=(steps, md.n, 3))
            _run(s, is_oracle, steps, mcfg.timestep, z)
            out["x_anaphase"] = s.get_positions()
            stages.enter_telophase(s, t)
            out["force_telophase"] = s.compute_force()
            out["energy_telophase"] = s.compute_energy()
        else:
            md = synth.make_mitotic_design(d, mcfg, sisters=True)
            x = _f32(rng.normal(size=(md.n, 3)) * 1.5)
            s = make_sys(md.n)
            t = stages.setup_prometaphase(s, md, mcfg)
            s.set_positions(x)
            for tname, tid in t.items():
                out[f"force_{tname}"] = s.compute_force(1 << tid)
                out[f"energy_{tname}"] = s.compute_energy(1 << tid)
            out["force_total"] = s.compute_force()
        out["x"] = x
    else:
        raise KeyError(name)
    return out

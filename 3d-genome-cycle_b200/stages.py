"""Host-side mirror of the reference stage drivers' setup code.

Each function registers, on a system object, exactly the force-field terms the
corresponding reference driver registers, in the same order.  `sys` is either
the product's `api.System` (CUDA) or the test oracle's `OracleSystem`; both
expose the same methods, so parity tests build the two sides with one call.

Reference:
  interphase     SIM/stage_interphase/simulation_driver_{particles,forcefield}.cpp
  anatelophase   SIM/stage_anatelophase/simulation_driver.cpp:36-189
  prometaphase   SIM/stage_prometaphase/simulation_driver.cpp:24-182
"""
from __future__ import annotations

import math

import numpy as np

GC_BOND_SEMISPRING, GC_BOND_HARMONIC, GC_BOND_SPRING = 0, 1, 2
GC_SOURCE_SEMISPRING, GC_SOURCE_FORCE_FLUX = 0, 1


def chain_bonds(chains):
    """add_bonded_range: (i, i+1) for i in [start, end-1)
    (MD/forcefield/bonded_pairwise_forcefield.hpp:45-52)."""
    out = [np.stack([np.arange(s, e - 1), np.arange(s + 1, e)], axis=1)
           for s, e in chains if e - s >= 2]
    return np.concatenate(out).astype(np.int32) if out else np.zeros((0, 2), np.int32)


def chain_triples(ranges):
    """add_bonded_range of the triple forcefield: (i, i+1, i+2), i + 2 < end
    (MD/forcefield/bonded_triplewise_forcefield.hpp:45-52)."""
    out = [np.stack([np.arange(s, e - 2), np.arange(s + 1, e - 1), np.arange(s + 2, e)], axis=1)
           for s, e in ranges if e - s >= 3]
    return np.concatenate(out).astype(np.int32) if out else np.zeros((0, 3), np.int32)


def setup_interphase(sys, design, cfg, wall_ode=True, schedule=True, contacts=True):
    """simulation_driver::setup() of the interphase stage.  Returns term ids."""
    sys.set_bead_attrs(design.mobility, design.a, design.b)
    terms = {}
    # setup_repulsive_forcefield (simulation_driver_forcefield.cpp:20-53)
    terms["repulsion"] = sys.add_pair_softcore_ab(cfg.a_core_repulsion, cfg.a_core_diameter,
                                                  cfg.b_core_repulsion, cfg.b_core_diameter, True)
    # setup_connectivity_forcefield (56-96)
    terms["chain"] = sys.add_bonds(chain_bonds(design.chains), GC_BOND_SEMISPRING,
                                   K_a=cfg.a_core_bond_spring, K_b=cfg.b_core_bond_spring,
                                   l_a=cfg.a_core_bond_length, l_b=cfg.b_core_bond_length,
                                   mix_ab=True, scale=True)
    # setup_loop_forcefield (99-136): (i, i+2) harmonic bonds; K is 0 with the
    # reference's defaults but the term is still registered and iterated
    loops = [np.stack([np.arange(s, e - 2), np.arange(s + 2, e)], axis=1)
             for s, e in design.chains if e - s >= 3]
    loops = np.concatenate(loops).astype(np.int32) if loops else np.zeros((0, 2), np.int32)
    terms["loops"] = sys.add_bonds(loops, GC_BOND_HARMONIC, K_a=cfg.a_core_2nd_bond_spring,
                                   K_b=cfg.b_core_2nd_bond_spring, mix_ab=True, scale=True)
    # setup_nucleolus_forcefield (139-185)
    terms["nor_bonds"] = sys.add_bonds(design.nucleolar_bonds, GC_BOND_SEMISPRING,
                                       K_a=cfg.nucleolus_bond_spring, l_a=cfg.nucleolus_bond_length,
                                       mix_ab=False, scale=True)
    if cfg.nucleolus_droplet_energy != 0:
        terms["droplet"] = sys.add_pair_softwell(cfg.nucleolus_droplet_energy,
                                                 cfg.nucleolus_droplet_decay,
                                                 cfg.nucleolus_droplet_cutoff,
                                                 design.nucleolar_bonds[:, 1])
    # setup_membrane_forcefield (189-244)
    terms["wall"] = sys.set_wall(cfg.wall_semiaxes_init, cfg.a_core_repulsion, cfg.a_core_diameter,
                                 cfg.b_core_repulsion, cfg.b_core_diameter, cfg.wall_ab_factor[0],
                                 cfg.wall_ab_factor[1], cfg.wall_packing_spring, scale_by_core=True,
                                 ode_enabled=wall_ode, ode_mobility=cfg.wall_mobility,
                                 ode_spring=cfg.wall_semiaxes_spring)
    # setup_context (simulation_driver.cpp:34-43)
    sys.set_schedule(cfg.core_scale_init, cfg.core_scale_tau, cfg.bond_scale_init,
                     cfg.bond_scale_tau, enabled=schedule)
    sys.set_context(time=0.0, semiaxes=cfg.wall_semiaxes_init, core_scale=cfg.core_scale_init,
                    bond_scale=cfg.bond_scale_init, contact_distance=0.0)
    if contacts:
        sys.set_contact_map(cfg.contactmap_distance, cfg.contactmap_update_interval,
                            scale_by_core=True)
    return terms


def begin_relaxation(sys, cfg):
    """run_relaxation (simulation_driver_relaxation.cpp:8-56): scales and wall stay
    at their initial values and the callback never touches the contact map."""
    sys.set_schedule(cfg.core_scale_init, cfg.core_scale_tau, cfg.bond_scale_init,
                     cfg.bond_scale_tau, enabled=False)
    sys.set_contact_map(cfg.contactmap_distance, 0, scale_by_core=True)


def begin_interphase(sys, cfg):
    """run_simulation up to and including callback(0)
    (simulation_driver_interphase.cpp:8-55): schedule and contact cadence on."""
    sys.set_schedule(cfg.core_scale_init, cfg.core_scale_tau, cfg.bond_scale_init,
                     cfg.bond_scale_tau, enabled=True)
    sys.set_contact_map(cfg.contactmap_distance, cfg.contactmap_update_interval,
                        scale_by_core=True)
    interphase_callback0(sys, cfg)


def interphase_callback0(sys, cfg):
    """callback(0) of run_simulation (simulation_driver_interphase.cpp:55):
    update_core_scale at t = 0 and one wall-ODE step with zero reaction
    statistics.  Done on the host through get/set_context."""
    c = sys.get_context(0)
    core = 1 - (1 - cfg.core_scale_init) * math.exp(-0.0 / cfg.core_scale_tau)
    bond = 1 - (1 - cfg.bond_scale_init) * math.exp(-0.0 / cfg.bond_scale_tau)
    semi = [c.semiaxes[k] + cfg.timestep * cfg.wall_mobility *
            (0.0 - cfg.wall_semiaxes_spring[k] * c.semiaxes[k]) for k in range(3)]
    sys.set_context(time=0.0, semiaxes=semi, core_scale=core, bond_scale=bond,
                    contact_distance=cfg.contactmap_distance * core)


def bending_ranges(md, cfg):
    """Kinetochore bead excluded from bending unless penalize_centromere_bending
    (SIM/stage_anatelophase/simulation_driver.cpp:125-135)."""
    out = []
    for (s, e), k in zip(md.chains, md.kinetochores):
        if cfg.penalize_centromere_bending:
            out.append((s, e))
        else:
            out.append((s, k))
            out.append((k + 1, e))
    return out


def kfiber_springs(md, cfg, decay_rate, chain_ids=None):
    """K = decay_rate / (core_mobility / chain length)
    (SIM/stage_anatelophase/simulation_driver.cpp:156-170,
     SIM/common/forcefield/kinetochore_fiber_forcefield.cpp:46-53)."""
    ids = range(len(md.chains)) if chain_ids is None else chain_ids
    beads, K = [], []
    for c in ids:
        s, e = md.chains[c]
        beads.append(int(md.kinetochores[c]))
        K.append(decay_rate / (cfg.core_mobility / float(e - s)))
    return np.asarray(beads, np.int32), np.asarray(K)


def setup_anatelophase(sys, md, cfg):
    """simulation_driver::setup() of anatelophase.  Anaphase terms start
    enabled; dragging/packing/telophase terms are registered disabled and
    switched by enter_anaphase()/enter_telophase()."""
    sys.set_bead_attrs(np.full(md.n, cfg.core_mobility))
    t = {}
    t["repulsion"] = sys.add_pair_softcore(cfg.core_repulsion, cfg.core_diameter)
    bonds = chain_bonds(md.chains)
    tri = chain_triples(bending_ranges(md, cfg))
    t["ana_bonds"] = sys.add_bonds(bonds, GC_BOND_SEMISPRING, K_a=cfg.bond_spring,
                                   l_a=cfg.bond_length)
    t["ana_bend"] = sys.add_triples(tri, cfg.bending_energy)
    t["telo_bonds"] = sys.add_bonds(bonds, GC_BOND_SEMISPRING,
                                    K_a=cfg.bond_spring * cfg.telophase_bond_spring_multiplier,
                                    l_a=cfg.bond_length)
    t["telo_bend"] = sys.add_triples(tri, cfg.bending_energy *
                                     cfg.telophase_bending_energy_multiplier)
    beads, K = kfiber_springs(md, cfg, cfg.kfiber_decay_rate_anaphase)
    t["dragging"] = sys.add_kfiber(cfg.anaphase_spindle_shift, beads, K,
                                   np.full(len(beads), cfg.kfiber_length_anaphase))
    t["packing"] = sys.add_point_source(GC_SOURCE_SEMISPRING, (0.0, 0.0, 0.0),
                                        cfg.telophase_packing_spring, cfg.telophase_packing_radius)
    for k in ("telo_bonds", "telo_bend", "dragging", "packing"):
        sys.set_term_enabled(t[k], False)
    return t


def enter_anaphase(sys, t):
    """run_dragging_stage: add the dragging force field (simulation_driver.cpp:244)."""
    sys.set_term_enabled(t["dragging"], True)


def enter_telophase(sys, t):
    """run_packing_stage force-field swap (simulation_driver.cpp:277-285)."""
    sys.set_term_enabled(t["dragging"], False)
    sys.set_term_enabled(t["packing"], True)
    sys.set_term_enabled(t["ana_bonds"], False)
    sys.set_term_enabled(t["ana_bend"], False)
    sys.set_term_enabled(t["telo_bonds"], True)
    sys.set_term_enabled(t["telo_bend"], True)


def setup_prometaphase(sys, md, cfg):
    """simulation_driver::setup() of prometaphase (sister chromatids)."""
    sys.set_bead_attrs(np.full(md.n, cfg.core_mobility))
    t = {}
    t["repulsion"] = sys.add_pair_softcore(cfg.core_repulsion, cfg.core_diameter)
    t["bonds"] = sys.add_bonds(chain_bonds(md.chains), GC_BOND_SEMISPRING, K_a=cfg.bond_spring,
                               l_a=cfg.bond_length)
    t["bend"] = sys.add_triples(chain_triples(bending_ranges(md, cfg)), cfg.bending_energy)
    # sister bonds use bond_spring, not sister_spring (simulation_driver.cpp:106)
    sis = np.stack([md.kinetochores[md.sisters[:, 0]], md.kinetochores[md.sisters[:, 1]]], axis=1)
    t["sisters"] = sys.add_bonds(sis, GC_BOND_SEMISPRING, K_a=cfg.bond_spring,
                                 l_a=cfg.sister_separation)
    for name, col, pole in (("kfiber_target", 0, md.poles[0]), ("kfiber_sister", 1, md.poles[1])):
        beads, K = kfiber_springs(md, cfg, cfg.kfiber_decay_rate_prometaphase, md.sisters[:, col])
        t[name] = sys.add_kfiber(pole, beads, K, np.full(len(beads), cfg.kfiber_length_prometaphase))
    b = math.sqrt(cfg.polar_ejection_cross_section)
    t["ejection_target"] = sys.add_point_source(GC_SOURCE_FORCE_FLUX, md.poles[0],
                                                cfg.polar_ejection_force, b)
    t["ejection_sister"] = sys.add_point_source(GC_SOURCE_FORCE_FLUX, md.poles[1],
                                                cfg.polar_ejection_force, b)
    return t

// ref_harness.cpp -- "reference" oracle: the reference's own micromd engine,
// compiled UNMODIFIED from /root/reference, behind the gc_oracle.h interface.
//
// TEST INFRASTRUCTURE ONLY (see gc_oracle.h).  This file contains no physics
// of its own: every force/energy/neighbor/integrator evaluation is done by
// <md.hpp> (micromd @ a886d68) and by the reference's contact_map.cpp and
// kinetochore_fiber_forcefield.cpp, which oracle/Makefile compiles in place.
// What is restated here is only driver glue that cannot be compiled because it
// pulls in HDF5:
//   - the potential lambdas of the stage drivers
//       SIM/stage_interphase/simulation_driver_forcefield.cpp:20-244
//       SIM/stage_anatelophase/simulation_driver.cpp:77-189
//       SIM/stage_prometaphase/simulation_driver.cpp:51-182
//   - the interphase per-step callback
//       SIM/stage_interphase/simulation_driver_interphase.cpp:16-53,67-90
//   - brownian_simulator::simulate_step with an injectable noise source
//       MD/simulation/detail/brownian_simulator.hpp:44-76
//     (proven bit-identical to md::simulate_brownian_dynamics in
//      tests/test_oracle.py::test_restated_stepper_matches_reference).

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <functional>
#include <memory>
#include <utility>
#include <vector>

#include <md.hpp>

#include "common/forcefield/kinetochore_fiber_forcefield.hpp"
#include "common/particle_data.hpp"
#include "common/potentials/force_flux_potential.hpp"
#include "stage_interphase/contact_map.hpp"
#include <spline.hpp>   // submodules/github.com/snsinfu/cxx-spline/include (reference, unmodified)

#include "gc_oracle.h"

namespace {

struct term_entry {
    std::vector<std::shared_ptr<md::forcefield>> ffs;
    bool enabled = true;
    // pair terms: how to recompute the neighbor distance (for gco_verlet_pairs)
    std::function<md::scalar()> neighbor_distance;
    std::vector<md::index> targets;
};

template<typename FF>
std::shared_ptr<md::forcefield> share(FF const &ff)
{
    return std::make_shared<FF>(ff);
}

template<typename FF>
std::shared_ptr<FF> share_typed(FF const &ff)
{
    return std::make_shared<FF>(ff);
}

} // namespace

struct gco_sys {
    md::system system;
    md::index n = 0;
    gc_context ctx{};
    gc_schedule sched{};
    gc_wall_params wall{};
    bool has_wall = false;
    std::vector<term_entry> terms;
    std::function<md::vector()> packing_reaction;
    std::function<md::scalar()> reaction_force;

    contact_map contacts;
    double contact_base = 0;
    int64_t contact_interval = 0;
    bool contact_scaled = false;

    double core() const { return ctx.core_scale; }
    double bond() const { return ctx.bond_scale; }

    int push(term_entry t)
    {
        for (auto &ff : t.ffs) {
            system.add_forcefield(ff);
        }
        terms.push_back(std::move(t));
        return int(terms.size()) - 1;
    }
};

extern "C" {

const char *gco_kind(void) { return "reference"; }

gco_sys *gco_create(int32_t n_beads)
{
    auto *s = new gco_sys;
    s->n = md::index(n_beads);
    s->system.add_attribute(particle_data_attribute);
    for (md::index i = 0; i < s->n; i++) {
        s->system.add_particle();
    }
    s->ctx.core_scale = 1;
    s->ctx.bond_scale = 1;
    return s;
}

void gco_destroy(gco_sys *s) { delete s; }

void gco_set_bead_attrs(gco_sys *s, const double *mobility, const double *a, const double *b)
{
    auto mob = s->system.view_mobilities();
    auto data = s->system.view(particle_data_attribute);
    for (md::index i = 0; i < s->n; i++) {
        mob[i] = mobility[i];
        data[i].a_factor = a ? a[i] : 0;
        data[i].b_factor = b ? b[i] : 0;
    }
}

void gco_set_positions(gco_sys *s, const double *xyz)
{
    auto pos = s->system.view_positions();
    for (md::index i = 0; i < s->n; i++) {
        pos[i] = {xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]};
    }
}

void gco_get_positions(gco_sys *s, double *xyz)
{
    auto pos = s->system.view_positions();
    for (md::index i = 0; i < s->n; i++) {
        xyz[3 * i] = pos[i].x;
        xyz[3 * i + 1] = pos[i].y;
        xyz[3 * i + 2] = pos[i].z;
    }
}

// SIM/stage_interphase/simulation_driver_forcefield.cpp:20-53
int gco_add_pair_softcore_ab(gco_sys *s, double eps_a, double sigma_a, double eps_b,
                             double sigma_b, int32_t scale_by_core)
{
    auto const max_diameter = std::max(sigma_a, sigma_b);
    auto scale = [=] { return scale_by_core ? s->core() : 1.0; };

    term_entry t;
    t.neighbor_distance = [=] { return max_diameter * scale(); };
    t.ffs.push_back(share(
        md::make_neighbor_pairwise_forcefield([=](md::index i, md::index j) {
            auto const data = s->system.view(particle_data_attribute);
            auto const a = 0.5 * (data[i].a_factor + data[j].a_factor);
            auto const b = 0.5 * (data[i].b_factor + data[j].b_factor);
            md::softcore_potential<2, 3> const a_potential{
                .energy = eps_a,
                .diameter = sigma_a * scale(),
            };
            md::softcore_potential<8, 3> const b_potential{
                .energy = eps_b,
                .diameter = sigma_b * scale(),
            };
            return a * a_potential + b * b_potential;
        }).set_neighbor_distance(t.neighbor_distance)));
    return s->push(std::move(t));
}

// SIM/stage_anatelophase/simulation_driver.cpp:77-91,
// SIM/stage_prometaphase/simulation_driver.cpp:51-63
int gco_add_pair_softcore(gco_sys *s, double eps, double sigma)
{
    term_entry t;
    t.neighbor_distance = [=] { return sigma; };
    t.ffs.push_back(share(md::make_neighbor_pairwise_forcefield(md::softcore_potential<2, 3>{
                                                                    .energy = eps,
                                                                    .diameter = sigma,
                                                                })
                              .set_neighbor_distance(sigma)));
    return s->push(std::move(t));
}

// SIM/stage_interphase/simulation_driver_forcefield.cpp:162-185
int gco_add_pair_softwell(gco_sys *s, double eps, double decay, double cutoff,
                          int32_t n_targets, const int32_t *targets)
{
    term_entry t;
    t.neighbor_distance = [=] { return cutoff; };
    t.targets.assign(targets, targets + n_targets);
    t.ffs.push_back(share(md::make_neighbor_pairwise_forcefield(
                              md::apply_cutoff(
                                  md::softwell_potential<6>{
                                      .energy = eps,
                                      .decay_distance = decay,
                                  },
                                  cutoff))
                              .set_neighbor_distance(cutoff)
                              .set_neighbor_targets(t.targets)));
    return s->push(std::move(t));
}

// SIM/stage_interphase/simulation_driver_forcefield.cpp:56-160 (interphase chain,
// loop and NOR bonds), SIM/stage_anatelophase/simulation_driver.cpp:99-117,
// SIM/stage_prometaphase/simulation_driver.cpp:69-118 (mitotic and sister bonds)
int gco_add_bonds(gco_sys *s, const gc_bond_params *params, int32_t n, const int32_t *pairs)
{
    gc_bond_params const p = *params;

    // Mixed/scaled spring constant and length exactly as the driver lambdas.
    auto mixed = [=](md::index i, md::index j) {
        md::scalar K = p.K_a;
        md::scalar l = p.l_a;
        if (p.mix_ab) {
            auto const data = s->system.view(particle_data_attribute);
            auto const a = 0.5 * (data[i].a_factor + data[j].a_factor);
            auto const b = 0.5 * (data[i].b_factor + data[j].b_factor);
            K = a * p.K_a + b * p.K_b;
            l = a * p.l_a + b * p.l_b;
        }
        if (p.scale) {
            auto const sc = s->bond();
            K = K * (1 / (sc * sc));
            l = l * sc;
        }
        return std::make_pair(K, l);
    };

    auto fill = [&](auto ff) {
        for (int32_t k = 0; k < n; k++) {
            ff.add_bonded_pair(md::index(pairs[2 * k]), md::index(pairs[2 * k + 1]));
        }
        return share(ff);
    };

    term_entry t;
    switch (p.kind) {
    case GC_BOND_SEMISPRING:
        t.ffs.push_back(fill(md::make_bonded_pairwise_forcefield([=](md::index i, md::index j) {
            auto const [K, l] = mixed(i, j);
            return md::semispring_potential{
                .spring_constant = K,
                .equilibrium_distance = l,
            };
        })));
        break;
    case GC_BOND_HARMONIC:
        t.ffs.push_back(fill(md::make_bonded_pairwise_forcefield([=](md::index i, md::index j) {
            auto const [K, l] = mixed(i, j);
            (void) l;
            return md::harmonic_potential{
                .spring_constant = K,
            };
        })));
        break;
    case GC_BOND_SPRING:
        t.ffs.push_back(fill(md::make_bonded_pairwise_forcefield([=](md::index i, md::index j) {
            auto const [K, l] = mixed(i, j);
            return md::spring_potential{
                .spring_constant = K,
                .equilibrium_distance = l,
            };
        })));
        break;
    default:
        return -1;
    }
    return s->push(std::move(t));
}

// SIM/stage_anatelophase/simulation_driver.cpp:119-141
int gco_add_triples(gco_sys *s, double bending_energy, int32_t n, const int32_t *triples)
{
    auto ff = md::make_bonded_triplewise_forcefield(md::cosine_bending_potential{
        .bending_energy = bending_energy,
    });
    for (int32_t k = 0; k < n; k++) {
        ff.add_bonded_triple(md::index(triples[3 * k]), md::index(triples[3 * k + 1]),
                             md::index(triples[3 * k + 2]));
    }
    term_entry t;
    t.ffs.push_back(share(ff));
    return s->push(std::move(t));
}

// SIM/stage_interphase/simulation_driver_forcefield.cpp:189-244
int gco_set_wall(gco_sys *s, const gc_wall_params *params)
{
    s->wall = *params;
    s->has_wall = true;
    s->ctx.semiaxes[0] = params->semiaxes[0];
    s->ctx.semiaxes[1] = params->semiaxes[1];
    s->ctx.semiaxes[2] = params->semiaxes[2];
    gc_wall_params const w = *params;

    auto set_ellipsoid = [s] {
        return md::ellipsoid{
            .center = {0, 0, 0},
            .semiaxis_x = s->ctx.semiaxes[0],
            .semiaxis_y = s->ctx.semiaxes[1],
            .semiaxis_z = s->ctx.semiaxes[2],
        };
    };
    auto scale = [=] { return w.scale_by_core ? s->core() : 1.0; };

    // inward: A/B-mixed soft core with HALF diameters (189-228)
    auto inward = share_typed(md::make_ellipsoid_inward_forcefield([=](md::index i) {
                                  auto const data = s->system.view(particle_data_attribute);
                                  auto const a = 0.5 * (data[i].a_factor + w.wall_a);
                                  auto const b = 0.5 * (data[i].b_factor + w.wall_b);
                                  md::softcore_potential<2, 3> const a_potential{
                                      .energy = w.eps_a,
                                      .diameter = w.sigma_a / 2 * scale(),
                                  };
                                  md::softcore_potential<8, 3> const b_potential{
                                      .energy = w.eps_b,
                                      .diameter = w.sigma_b / 2 * scale(),
                                  };
                                  return a * a_potential + b * b_potential;
                              }).set_ellipsoid(set_ellipsoid));

    // outward: harmonic (230-237)
    auto outward = share_typed(md::make_ellipsoid_outward_forcefield(md::harmonic_potential{
                                                                         .spring_constant = w.outward_spring,
                                                                     })
                                   .set_ellipsoid(set_ellipsoid));

    s->packing_reaction = [=] {
        auto const in = inward->stats.axial_reaction;
        auto const out = outward->stats.axial_reaction;
        return in + out;
    };
    s->reaction_force = [=] {
        return inward->stats.reaction_force + outward->stats.reaction_force;
    };

    term_entry t;
    t.ffs.push_back(inward);
    t.ffs.push_back(outward);
    return s->push(std::move(t));
}

// SIM/stage_anatelophase/simulation_driver.cpp:175-189 (packing),
// SIM/stage_prometaphase/simulation_driver.cpp:162-182 (polar ejection)
int gco_add_point_source(gco_sys *s, int32_t kind, const double point[3], double p0, double p1)
{
    md::point const pt = {point[0], point[1], point[2]};
    term_entry t;
    if (kind == GC_SOURCE_SEMISPRING) {
        t.ffs.push_back(share(md::make_point_source_forcefield(md::semispring_potential{
                                                                   .spring_constant = p0,
                                                                   .equilibrium_distance = p1,
                                                               })
                                  .set_point_source(pt)));
    } else if (kind == GC_SOURCE_FORCE_FLUX) {
        t.ffs.push_back(share(md::make_point_source_forcefield(force_flux_potential{
                                                                   .constant_force = p0,
                                                                   .reactive_distance = p1,
                                                               })
                                  .set_point_source(pt)));
    } else {
        return -1;
    }
    return s->push(std::move(t));
}

// SIM/stage_anatelophase/simulation_driver.cpp:149-172: the driver passes
// (mobility, decay_rate, length); the spring is decay_rate / mobility
// (kinetochore_fiber_forcefield.cpp:46-53).  We hand the reference class
// mobility = 1 / spring and decay_rate = 1 so that it evaluates K = spring.
int gco_add_kfiber(gco_sys *s, const double pole[3], int32_t n, const int32_t *beads,
                   const double *spring, const double *length)
{
    kinetochore_fiber_forcefield ff;
    ff.set_pole_position({pole[0], pole[1], pole[2]});
    for (int32_t k = 0; k < n; k++) {
        ff.add_kinetochore({
            .particle_index = md::index(beads[k]),
            .mobility = 1.0,
            .decay_rate = spring[k],
            .stationary_length = length[k],
        });
    }
    term_entry t;
    t.ffs.push_back(share(ff));
    return s->push(std::move(t));
}

// system.remove_forcefield / add_forcefield as
// SIM/stage_anatelophase/simulation_driver.cpp:244,277-285
void gco_set_term_enabled(gco_sys *s, int32_t term, int32_t enabled)
{
    auto &t = s->terms.at(size_t(term));
    if (bool(enabled) == t.enabled) {
        return;
    }
    for (auto &ff : t.ffs) {
        if (enabled) {
            s->system.add_forcefield(ff);
        } else {
            s->system.remove_forcefield(ff);
        }
    }
    t.enabled = bool(enabled);
}

void gco_set_schedule(gco_sys *s, const gc_schedule *schedule) { s->sched = *schedule; }

void gco_set_context(gco_sys *s, const gc_context *in)
{
    s->ctx.time = in->time;
    s->ctx.core_scale = in->core_scale;
    s->ctx.bond_scale = in->bond_scale;
    s->ctx.contact_distance = in->contact_distance;
    for (int k = 0; k < 3; k++) {
        s->ctx.semiaxes[k] = in->semiaxes[k];
    }
    s->contacts.set_contact_distance(in->contact_distance);
}

void gco_get_context(gco_sys *s, gc_context *out)
{
    if (s->packing_reaction) {
        auto const r = s->packing_reaction();
        s->ctx.axial_reaction[0] = r.x;
        s->ctx.axial_reaction[1] = r.y;
        s->ctx.axial_reaction[2] = r.z;
        s->ctx.reaction_force = s->reaction_force();
    }
    s->ctx.contact_distance = s->contacts.contact_distance();
    *out = s->ctx;
}

// ---------------------------------------------------------------------------
// Dynamics
// ---------------------------------------------------------------------------

} // extern "C"

namespace {

// SIM/stage_interphase/simulation_driver_interphase.cpp:67-76
void update_core_scale(gco_sys *s)
{
    if (s->sched.enabled) {
        auto const t_bead = s->ctx.time / s->sched.core_tau;
        auto const t_bond = s->ctx.time / s->sched.bond_tau;
        s->ctx.core_scale = 1 - (1 - s->sched.core_init) * std::exp(-t_bead);
        s->ctx.bond_scale = 1 - (1 - s->sched.bond_init) * std::exp(-t_bond);
    }
    if (s->contact_scaled) {
        s->contacts.set_contact_distance(s->contact_base * s->ctx.core_scale);
    }
}

// SIM/stage_interphase/simulation_driver_interphase.cpp:79-90
void update_wall_semiaxes(gco_sys *s)
{
    if (!s->has_wall || !s->wall.ode_enabled) {
        return;
    }
    md::vector semiaxes = {s->ctx.semiaxes[0], s->ctx.semiaxes[1], s->ctx.semiaxes[2]};
    md::vector const spring = {s->wall.ode_spring[0], s->wall.ode_spring[1],
                               s->wall.ode_spring[2]};
    md::vector net_force;
    net_force += s->packing_reaction();
    net_force -= spring.hadamard(semiaxes);
    semiaxes += s->ctx.last_timestep * s->wall.ode_mobility * net_force;
    s->ctx.semiaxes[0] = semiaxes.x;
    s->ctx.semiaxes[1] = semiaxes.y;
    s->ctx.semiaxes[2] = semiaxes.z;
}

// MD/simulation/detail/brownian_simulator.hpp:44-76 with the normal source
// as a parameter.  Timesteppers are the reference's own classes.
template<typename Normal>
md::scalar restated_step(gco_sys *s, md::detail::brownian_timestepper &timestepper,
                         md::scalar temperature, std::vector<md::vector> &forces_,
                         std::vector<md::vector> &weiners_, Normal &&normal)
{
    md::array_view<md::scalar const> mobilities = s->system.view_mobilities();
    md::array_view<md::point> positions = s->system.view_positions();
    md::array_view<md::vector> forces = forces_;
    md::array_view<md::vector> weiners = weiners_;

    s->system.compute_force(forces);

    md::scalar const timestep = timestepper.determine_timestep(mobilities, forces, temperature);

    for (md::index i = 0; i < s->system.particle_count(); i++) {
        md::scalar const mu_dt = timestep * mobilities[i];
        md::scalar const sigma = std::sqrt(2 * temperature * mu_dt);
        md::vector const weiner = {sigma * normal(), sigma * normal(), sigma * normal()};

        positions[i] += mu_dt * forces[i];
        positions[i] += 0.5 * (weiner + weiners[i]);
        weiners[i] = weiner;
    }
    return timestep;
}

} // namespace

extern "C" {

void gco_run(gco_sys *s, gco_run_config *c)
{
    md::index const n = s->n;
    md::step done = 0;

    // The interphase driver's wall ODE multiplies by _config.timestep
    // (simulation_driver_interphase.cpp:89), i.e. the configured dt.
    s->ctx.last_timestep = c->timestep;

    auto callback = [&](md::step step) {
        done = step;
        s->ctx.time = md::scalar(c->step_offset + step) * c->timestep;
        s->ctx.step = c->step_offset + step;
        auto const abs_step = c->step_offset + step;

        if (c->energy_interval && abs_step % c->energy_interval == 0) {
            c->last_energy = s->system.compute_energy() / md::scalar(n);
        }
        if (c->traj_out) {
            gco_get_positions(s, c->traj_out + size_t(step - 1) * n * 3);
        }
        if (s->contact_interval && abs_step % s->contact_interval == 0) {
            s->contacts.update(s->system.view_positions());
        }
        if (c->interphase_callback) {
            update_core_scale(s);
            update_wall_semiaxes(s);
        }
        if (c->ctx_out) {
            double *o = c->ctx_out + size_t(step - 1) * 8;
            o[0] = s->ctx.time;
            o[1] = s->ctx.semiaxes[0];
            o[2] = s->ctx.semiaxes[1];
            o[3] = s->ctx.semiaxes[2];
            o[4] = s->ctx.core_scale;
            o[5] = s->ctx.bond_scale;
            o[6] = s->ctx.last_timestep;
            o[7] = 0;
        }
    };

    if (c->stepper == 0 && c->noise == nullptr && c->force_out == nullptr) {
        md::simulate_brownian_dynamics(s->system, {
            .temperature = c->temperature,
            .timestep    = c->timestep,
            .spacestep   = c->spacestep,
            .steps       = c->steps,
            .seed        = c->seed,
            .callback    = callback,
        });
        return;
    }

    // Restated loop: MD/simulation/brownian_dynamics.hpp:49-72
    std::unique_ptr<md::detail::brownian_timestepper> timestepper;
    if (c->spacestep != 0) {
        timestepper = std::make_unique<md::detail::adaptive_brownian_timestepper>(
            c->timestep, c->spacestep);
    } else {
        timestepper = std::make_unique<md::detail::monotonic_brownian_timestepper>(c->timestep);
    }
    std::vector<md::vector> forces(n), weiners(n);
    md::random_engine random;
    random.seed(c->seed);
    md::normal_distribution<md::scalar> normal;
    double const *noise = c->noise;

    for (md::step step = 1; step <= c->steps; step++) {
        md::scalar dt;
        if (noise) {
            double const *z = noise + size_t(step - 1) * n * 3;
            size_t k = 0;
            dt = restated_step(s, *timestepper, c->temperature, forces, weiners,
                               [&] { return z[k++]; });
        } else {
            dt = restated_step(s, *timestepper, c->temperature, forces, weiners,
                               [&] { return normal(random); });
        }
        if (c->force_out) {
            double *o = c->force_out + size_t(step - 1) * n * 3;
            for (md::index i = 0; i < n; i++) {
                o[3 * i] = forces[i].x;
                o[3 * i + 1] = forces[i].y;
                o[3 * i + 2] = forces[i].z;
            }
        }
        callback(step);
        if (c->ctx_out) {
            c->ctx_out[size_t(step - 1) * 8 + 6] = dt;
        }
    }
}

void gco_compute_force(gco_sys *s, uint32_t term_mask, double *out)
{
    std::vector<md::vector> forces(s->n);
    for (size_t t = 0; t < s->terms.size(); t++) {
        if (!(term_mask & (1u << t)) || !s->terms[t].enabled) {
            continue;
        }
        for (auto &ff : s->terms[t].ffs) {
            ff->compute_force(s->system, forces);
        }
    }
    for (md::index i = 0; i < s->n; i++) {
        out[3 * i] = forces[i].x;
        out[3 * i + 1] = forces[i].y;
        out[3 * i + 2] = forces[i].z;
    }
}

double gco_compute_energy(gco_sys *s, uint32_t term_mask)
{
    double sum = 0;
    for (size_t t = 0; t < s->terms.size(); t++) {
        if (!(term_mask & (1u << t)) || !s->terms[t].enabled) {
            continue;
        }
        for (auto &ff : s->terms[t].ffs) {
            sum += ff->compute_energy(s->system);
        }
    }
    return sum;
}

} // extern "C"

namespace {

size_t emit_sorted(std::vector<std::pair<md::index, md::index>> &found, uint32_t *pairs,
                   size_t capacity)
{
    std::sort(found.begin(), found.end());
    for (size_t k = 0; k < found.size() && k < capacity; k++) {
        pairs[2 * k] = uint32_t(found[k].first);
        pairs[2 * k + 1] = uint32_t(found[k].second);
    }
    return found.size();
}

} // namespace

extern "C" {

// MD/misc/neighbor_searcher.hpp:22-131 with the open_box hint the Verlet list
// gives it (MD/forcefield/detail/neighbor_list_heuristics.hpp:86-92)
size_t gco_neighbor_pairs(gco_sys *s, double radius, uint32_t *pairs, size_t capacity)
{
    md::open_box box;
    box.particle_count = s->n;
    md::neighbor_searcher<md::open_box> searcher(box, radius);
    searcher.set_points(s->system.view_positions());
    std::vector<std::pair<md::index, md::index>> found;
    searcher.search(std::back_inserter(found));
    return emit_sorted(found, pairs, capacity);
}

// MD/forcefield/detail/neighbor_list.hpp:62-240 on a fresh list
size_t gco_verlet_pairs(gco_sys *s, int32_t term, uint32_t *pairs, size_t capacity)
{
    auto &t = s->terms.at(size_t(term));
    md::neighbor_list<md::open_box> list;
    if (!t.targets.empty()) {
        list.set_targets(t.targets);
    }
    list.update(s->system.view_positions(), t.neighbor_distance(), md::open_box{});
    std::vector<std::pair<md::index, md::index>> found(list.begin(), list.end());
    for (auto &pr : found) {
        if (pr.first > pr.second) {
            std::swap(pr.first, pr.second);
        }
    }
    return emit_sorted(found, pairs, capacity);
}

void gco_set_contact_map(gco_sys *s, double distance, int64_t interval, int32_t scale_by_core)
{
    s->contact_base = distance;
    s->contact_interval = interval;
    s->contact_scaled = scale_by_core != 0;
    s->contacts.set_contact_distance(scale_by_core ? distance * s->ctx.core_scale : distance);
}

// SIM/stage_interphase/contact_map.cpp:33-63 (the reference's own object code)
void gco_contacts_update(gco_sys *s) { s->contacts.update(s->system.view_positions()); }

// SIM/stage_interphase/contact_map.cpp:66-85
size_t gco_contacts_fetch(gco_sys *s, uint32_t *ijv, size_t capacity)
{
    auto const triples = s->contacts.accumulate();
    for (size_t k = 0; k < triples.size() && k < capacity; k++) {
        ijv[3 * k] = triples[k][0];
        ijv[3 * k + 1] = triples[k][1];
        ijv[3 * k + 2] = triples[k][2];
    }
    return triples.size();
}

void gco_contacts_clear(gco_sys *s) { s->contacts.clear(); }

size_t gco_reference_normals(uint64_t seed, size_t n, double *out)
{
    md::random_engine random;
    random.seed(seed);
    md::normal_distribution<md::scalar> normal;
    for (size_t k = 0; k < n; k++) {
        out[k] = normal(random);
    }
    return n;
}

double gco_potential_probe(int32_t kind, double p0, double p1, double p2, const double r[3],
                           double f[3])
{
    md::vector const rv = {r[0], r[1], r[2]};
    md::vector fv;
    double e = 0;
    auto eval = [&](auto const &pot) {
        e = pot.evaluate_energy(rv);
        fv = pot.evaluate_force(rv);
    };
    switch (kind) {
    case 0: eval(md::softcore_potential<2, 3>{.energy = p0, .diameter = p1}); break;
    case 1: eval(md::softcore_potential<8, 3>{.energy = p0, .diameter = p1}); break;
    case 2: eval(md::softwell_potential<6>{.energy = p0, .decay_distance = p1}); break;
    case 3: eval(md::semispring_potential{.spring_constant = p0, .equilibrium_distance = p1}); break;
    case 4: eval(md::harmonic_potential{.spring_constant = p0}); break;
    case 5: eval(md::spring_potential{.spring_constant = p0, .equilibrium_distance = p1}); break;
    case 6: eval(force_flux_potential{.constant_force = p0, .reactive_distance = p1}); break;
    case 7:
        eval(md::apply_cutoff(md::softwell_potential<6>{.energy = p0, .decay_distance = p1}, p2));
        break;
    case 8:
        switch (int(p2)) {
        case 21: eval(md::softcore_potential<2, 1>{.energy = p0, .diameter = p1}); break;
        case 22: eval(md::softcore_potential<2, 2>{.energy = p0, .diameter = p1}); break;
        case 23: eval(md::softcore_potential<2, 3>{.energy = p0, .diameter = p1}); break;
        case 33: eval(md::softcore_potential<3, 3>{.energy = p0, .diameter = p1}); break;
        case 44: eval(md::softcore_potential<4, 4>{.energy = p0, .diameter = p1}); break;
        case 55: eval(md::softcore_potential<5, 5>{.energy = p0, .diameter = p1}); break;
        case 83: eval(md::softcore_potential<8, 3>{.energy = p0, .diameter = p1}); break;
        default: break;
        }
        break;
    default: break;
    }
    f[0] = fv.x;
    f[1] = fv.y;
    f[2] = fv.z;
    return e;
}

double gco_bending_probe(double bending_energy, const double rij[3], const double rjk[3],
                         double f[9])
{
    md::cosine_bending_potential const pot{.bending_energy = bending_energy};
    md::vector const a = {rij[0], rij[1], rij[2]};
    md::vector const b = {rjk[0], rjk[1], rjk[2]};
    auto const [fi, fj, fk] = pot.evaluate_force(a, b);
    md::vector const fs[3] = {fi, fj, fk};
    for (int k = 0; k < 3; k++) {
        f[3 * k] = fs[k].x;
        f[3 * k + 1] = fs[k].y;
        f[3 * k + 2] = fs[k].z;
    }
    return pot.evaluate_energy(a, b);
}

void gco_ellipsoid_probe(const double semiaxes[3], const double pt[3], double out[8])
{
    md::ellipsoid const el{
        .center = {0, 0, 0},
        .semiaxis_x = semiaxes[0],
        .semiaxis_y = semiaxes[1],
        .semiaxis_z = semiaxes[2],
    };
    auto const ev = md::detail::evaluate_point(el, {pt[0], pt[1], pt[2]});
    for (int k = 0; k < 8; k++) out[k] = 0;
    out[0] = ev.undefined ? 1 : 0;
    if (ev.undefined) return;
    out[1] = ev.delta.x; out[2] = ev.delta.y; out[3] = ev.delta.z;
    out[4] = ev.strain.x; out[5] = ev.strain.y; out[6] = ev.strain.z;
    out[7] = ev.implicit;
}


// ---- inter-stage transitions (SIM/stage_transition/*.cpp cannot be compiled here: they
// include simulation_store.hpp -> <h5.hpp> -> <hdf5.h>, absent; their bodies are restated
// around the reference's OWN spline implementation, <spline.hpp>, compiled unmodified) ------
void gco_resample_chain(int32_t n, const double *chain, int32_t new_n, double *out)
{
    // transition_interphase.cpp:16-41
    constexpr auto spline_bc = cubic_spline::not_a_knot;
    std::vector<double> ts, xs, ys, zs;
    for (int32_t k = 0; k < n; k++) {
        auto const t = (0.5 + double(ts.size())) / double(n);
        ts.push_back(t);
        xs.push_back(chain[3 * k]);
        ys.push_back(chain[3 * k + 1]);
        zs.push_back(chain[3 * k + 2]);
    }
    cubic_spline x_spline(ts, xs, spline_bc);
    cubic_spline y_spline(ts, ys, spline_bc);
    cubic_spline z_spline(ts, zs, spline_bc);
    for (int32_t i = 0; i < new_n; i++) {
        auto const t = (0.5 + double(i)) / double(new_n);
        out[3 * i] = x_spline(t);
        out[3 * i + 1] = y_spline(t);
        out[3 * i + 2] = z_spline(t);
    }
}

void gco_coarse_grain(const double *positions, int64_t source_start, int64_t source_len, int64_t cg,
                      int64_t coarse_len, const double disp[3], double *target, double *sister)
{
    // transition_prometaphase.cpp:24-41 (compute_centroid) and 81-98, with md::point / md::vector
    md::vector const sister_displacement = {disp[0], disp[1], disp[2]};
    for (int64_t offset = 0; offset < coarse_len; offset++) {
        auto const start = source_start + cg * offset;
        auto const end = std::min(start + cg, start + source_len);
        md::point const origin = {};
        md::vector coords = {};
        md::scalar count = 0;
        for (int64_t k = start; k < end; k++) {
            md::point const point = {positions[3 * k], positions[3 * k + 1], positions[3 * k + 2]};
            coords += point - origin;
            count++;
        }
        coords /= count;
        md::point const centroid = origin + coords;
        md::point const sis = centroid + sister_displacement;
        target[3 * offset] = centroid.x; target[3 * offset + 1] = centroid.y; target[3 * offset + 2] = centroid.z;
        sister[3 * offset] = sis.x; sister[3 * offset + 1] = sis.y; sister[3 * offset + 2] = sis.z;
    }
}

void gco_quantize_f32(size_t n, const double *in, int32_t bits, float *out)
{
    // simulation_store.cpp:22-32 (quantize) + 199-207 (the float cast of save_positions)
    for (size_t k = 0; k < n; k++) {
        int exp;
        auto const mant = std::frexp(in[k], &exp);
        auto const scaled_mant = std::nearbyint(std::ldexp(mant, bits));
        out[k] = float(std::ldexp(scaled_mant, exp - bits));
    }
}

} // extern "C"

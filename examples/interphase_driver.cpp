// interphase_driver.cpp -- the reference's interphase stage driver, re-hosted on
// the B200 path through include/gc_md.hpp.
//
// Structure follows SIM/stage_interphase/ (SIM = src/simulation of the reference):
//   setup_particles   simulation_driver_particles.cpp:8-35
//   setup_forcefield  simulation_driver_forcefield.cpp:8-244
//   setup_context     simulation_driver.cpp:34-43
//   run_relaxation    simulation_driver_relaxation.cpp:8-56
//   run_simulation    simulation_driver_interphase.cpp:8-90
// The HDF5 store is replaced by two flat binary files (the trajectory I/O stays
// outside the hot path):
//   input : "GCD1" n n_chains n_nor | mobility[n] a[n] b[n] (f64) |
//           chains[n_chains][2] nor_bonds[n_nor][2] (i32) | positions[n][3] (f64)
//   output: "GCO1" n n_contacts | positions[n][3] (f64) | context: time, semiaxes[3],
//           core_scale, bond_scale, mean_energy (f64) | contacts[n_contacts][3] (u32)
//
//   interphase_driver <in> <out> <relaxation_steps> <steps> <seed> [sampling_interval]
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <ctime>
#include <iostream>
#include <string>
#include <vector>

#include "../include/gc_md.hpp"

namespace {

// interphase_config defaults (SIM/common/simulation_config.hpp:61-115)
struct interphase_config {
    double temperature = 1, timestep = 1e-5;
    std::int64_t steps = 700000, sampling_interval = 1000, logging_interval = 1000;
    double relaxation_spacestep = 0.001;
    std::int64_t relaxation_steps = 10000;
    double contactmap_distance = 0.24;
    std::int64_t contactmap_update_interval = 20, contactmap_output_window = 10;
    double a_core_diameter = 0.30, b_core_diameter = 0.24, a_core_repulsion = 2.5, b_core_repulsion = 2.5;
    double a_core_bond_spring = 100, b_core_bond_spring = 50, a_core_bond_length = 0, b_core_bond_length = 0;
    double a_core_2nd_bond_spring = 0, b_core_2nd_bond_spring = 0;   // absent from the JSON traits
    double core_scale_init = 0.5, core_scale_tau = 0.5, bond_scale_init = 0.5, bond_scale_tau = 0.5;
    double nucleolus_bond_spring = 10, nucleolus_bond_length = 0;
    double nucleolus_droplet_energy = 0.3, nucleolus_droplet_decay = 0.2, nucleolus_droplet_cutoff = 0.4;
    double wall_semiaxes_init[3] = {2, 2, 2}, wall_semiaxes_spring[3] = {3e4, 3e4, 3e4};
    double wall_packing_spring = 1000, wall_ab_factor[2] = {0, 10}, wall_mobility = 2e-4;
};

struct design {
    std::int32_t n = 0, n_chains = 0, n_nor = 0;
    std::vector<double> mobility, a, b, positions;
    std::vector<std::int32_t> chains, nor_bonds;
};

template<typename T>
void read_array(FILE* f, std::vector<T>& v, std::size_t count)
{
    v.resize(count);
    if (count && fread(v.data(), sizeof(T), count, f) != count) throw std::runtime_error("short input file");
}

design load_design(std::string const& path)
{
    FILE* f = fopen(path.c_str(), "rb");
    if (!f) throw std::runtime_error("cannot open " + path);
    char magic[4];
    design d;
    if (fread(magic, 1, 4, f) != 4 || memcmp(magic, "GCD1", 4) != 0) throw std::runtime_error("bad magic");
    if (fread(&d.n, 4, 1, f) != 1 || fread(&d.n_chains, 4, 1, f) != 1 || fread(&d.n_nor, 4, 1, f) != 1)
        throw std::runtime_error("short header");
    read_array(f, d.mobility, d.n);
    read_array(f, d.a, d.n);
    read_array(f, d.b, d.n);
    read_array(f, d.chains, 2 * (std::size_t) d.n_chains);
    read_array(f, d.nor_bonds, 2 * (std::size_t) d.n_nor);
    read_array(f, d.positions, 3 * (std::size_t) d.n);
    fclose(f);
    return d;
}

class simulation_driver {
public:
    simulation_driver(design const& d, interphase_config const& c, std::uint64_t seed)
        : _design(d), _config(c), _system(0, 1), _seed(seed)
    {
        setup_particles();
        setup_forcefield();
        setup_context();
    }

    void run()
    {
        run_relaxation();
        run_simulation();
    }

    void save(std::string const& path)
    {
        FILE* f = fopen(path.c_str(), "wb");
        if (!f) throw std::runtime_error("cannot write " + path);
        auto const positions = _system.view_positions();
        auto const contacts = _system.contact_map_accumulate();
        std::int32_t n = _design.n, m = (std::int32_t) (contacts.size() / 3);
        fwrite("GCO1", 1, 4, f);
        fwrite(&n, 4, 1, f);
        fwrite(&m, 4, 1, f);
        fwrite(positions.begin(), sizeof(gcmd::point), positions.size(), f);
        auto const ctx = _system.context();
        double tail[7] = {ctx.time, ctx.semiaxes[0], ctx.semiaxes[1], ctx.semiaxes[2], ctx.core_scale,
                          ctx.bond_scale, _mean_energy};
        fwrite(tail, sizeof(double), 7, f);
        fwrite(contacts.data(), sizeof(std::uint32_t), contacts.size(), f);
        fclose(f);
    }

private:
    void setup_particles()
    {
        for (std::int32_t i = 0; i < _design.n; i++) {
            _system.add_particle({.mobility = _design.mobility[i],
                                  .position = {_design.positions[3 * i], _design.positions[3 * i + 1],
                                               _design.positions[3 * i + 2]},
                                  .a_factor = _design.a[i],
                                  .b_factor = _design.b[i]});
        }
    }

    void setup_forcefield()
    {
        auto const& c = _config;
        // setup_repulsive_forcefield
        _system.add_forcefield(gcmd::neighbor_ab_softcore{c.a_core_repulsion, c.a_core_diameter,
                                                          c.b_core_repulsion, c.b_core_diameter, true});
        // setup_connectivity_forcefield
        auto chain = gcmd::ab_mixed_bonds(GC_BOND_SEMISPRING, c.a_core_bond_spring, c.b_core_bond_spring,
                                          c.a_core_bond_length, c.b_core_bond_length);
        // setup_loop_forcefield: (i, i + 2)
        auto loops = gcmd::ab_mixed_bonds(GC_BOND_HARMONIC, c.a_core_2nd_bond_spring,
                                          c.b_core_2nd_bond_spring, 0, 0);
        for (std::int32_t k = 0; k < _design.n_chains; k++) {
            auto const s = _design.chains[2 * k], e = _design.chains[2 * k + 1];
            chain.add_bonded_range(s, e);
            for (std::int32_t i = s; i + 2 < e; i++) loops.add_bonded_pair(i, i + 2);
        }
        _system.add_forcefield(chain);
        _system.add_forcefield(loops);
        // setup_nucleolus_forcefield
        auto nor = gcmd::semispring_bonds(c.nucleolus_bond_spring, c.nucleolus_bond_length);
        nor.params.scale = 1;
        gcmd::neighbor_softwell droplet{c.nucleolus_droplet_energy, c.nucleolus_droplet_decay,
                                        c.nucleolus_droplet_cutoff, {}};
        for (std::int32_t k = 0; k < _design.n_nor; k++) {
            nor.add_bonded_pair(_design.nor_bonds[2 * k], _design.nor_bonds[2 * k + 1]);
            droplet.targets.push_back(_design.nor_bonds[2 * k + 1]);
        }
        _system.add_forcefield(nor);
        if (c.nucleolus_droplet_energy != 0) _system.add_forcefield(droplet);
        // setup_membrane_forcefield (+ the wall ODE of update_wall_semiaxes)
        gcmd::ellipsoid_wall wall{};
        for (int k = 0; k < 3; k++) {
            wall.semiaxes[k] = c.wall_semiaxes_init[k];
            wall.ode_spring[k] = c.wall_semiaxes_spring[k];
        }
        wall.eps_a = c.a_core_repulsion; wall.sigma_a = c.a_core_diameter;
        wall.eps_b = c.b_core_repulsion; wall.sigma_b = c.b_core_diameter;
        wall.wall_a = c.wall_ab_factor[0]; wall.wall_b = c.wall_ab_factor[1];
        wall.outward_spring = c.wall_packing_spring;
        wall.scale_by_core = 1;
        wall.ode_enabled = 1;
        wall.ode_mobility = c.wall_mobility;
        _system.add_forcefield(wall);
    }

    void setup_context()
    {
        gc_context ctx{};
        ctx.time = 0;
        for (int k = 0; k < 3; k++) ctx.semiaxes[k] = _config.wall_semiaxes_init[k];
        ctx.core_scale = _config.core_scale_init;
        ctx.bond_scale = _config.bond_scale_init;
        _system.set_context(ctx);
        _system.set_seeds({_seed});
    }

    void run_relaxation()
    {
        // scales and wall frozen, no contact map (simulation_driver_relaxation.cpp:25-45)
        _system.set_schedule({0, 0, _config.core_scale_init, _config.core_scale_tau, _config.bond_scale_init,
                              _config.bond_scale_tau});
        _system.contact_map_configure(_config.contactmap_distance, 0, true);
        gcmd::simulate_brownian_dynamics(_system, {
            .temperature = _config.temperature,
            .timestep = _config.timestep,
            .spacestep = _config.relaxation_spacestep,
            .steps = _config.relaxation_steps,
            .callback = [this](gcmd::step step) { print_progress("relaxation", step); },
            .callback_interval = _config.logging_interval,
        });
    }

    void run_simulation()
    {
        _system.set_schedule({1, 0, _config.core_scale_init, _config.core_scale_tau, _config.bond_scale_init,
                              _config.bond_scale_tau});
        _system.contact_map_configure(_config.contactmap_distance, _config.contactmap_update_interval, true);
        // callback(0): update_core_scale + update_wall_semiaxes with zero reaction
        // (simulation_driver_interphase.cpp:55,67-90)
        auto ctx = _system.context();
        ctx.time = 0;
        ctx.core_scale = 1 - (1 - _config.core_scale_init) * std::exp(-0.0 / _config.core_scale_tau);
        ctx.bond_scale = 1 - (1 - _config.bond_scale_init) * std::exp(-0.0 / _config.bond_scale_tau);
        for (int k = 0; k < 3; k++) {
            ctx.semiaxes[k] += _config.timestep * _config.wall_mobility *
                               (0.0 - _config.wall_semiaxes_spring[k] * ctx.semiaxes[k]);
        }
        ctx.contact_distance = _config.contactmap_distance * ctx.core_scale;
        _system.set_context(ctx);

        gcmd::simulate_brownian_dynamics(_system, {
            .temperature = _config.temperature,
            .timestep = _config.timestep,
            .steps = _config.steps,
            .callback = [this](gcmd::step step) { print_progress("interphase", step); },
            .callback_interval = _config.sampling_interval,
        });
    }

    void print_progress(std::string const& phase, gcmd::step step)
    {
        _mean_energy = _system.compute_energy() / (double) _system.particle_count();
        auto const ctx = _system.context();
        auto const radius = std::cbrt(ctx.semiaxes[0] * ctx.semiaxes[1] * ctx.semiaxes[2]);
        std::clog << "[" << phase << "] " << std::time(nullptr) << "\t" << step << "\tt: " << ctx.time
                  << "\tR: " << radius << "\tE: " << _mean_energy << "\n";
    }

    design _design;
    interphase_config _config;
    gcmd::system _system;
    std::uint64_t _seed;
    double _mean_energy = 0;
};

} // namespace

int main(int argc, char** argv)
{
    if (argc < 6) {
        std::cerr << "usage: interphase_driver <in> <out> <relaxation_steps> <steps> <seed> [interval]\n";
        return 1;
    }
    try {
        interphase_config config;
        config.relaxation_steps = std::stoll(argv[3]);
        config.steps = std::stoll(argv[4]);
        if (argc > 6) config.sampling_interval = config.logging_interval = std::stoll(argv[6]);
        simulation_driver driver(load_design(argv[1]), config, std::stoull(argv[5]));
        driver.run();
        driver.save(argv[2]);
    } catch (std::exception const& e) {
        std::cerr << "error: " << e.what() << "\n";
        return 1;
    }
    return 0;
}

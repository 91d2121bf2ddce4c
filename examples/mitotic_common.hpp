// mitotic_common.hpp -- what the two mitotic stage drivers of examples/ share: the
// configuration defaults (SIM/common/simulation_config.hpp:15-58; SIM = src/simulation of the
// reference), the coarse-grained design, and the flat binary files that stand in for the
// HDF5 store (trajectory I/O stays outside the hot path):
//   input : "GCM1" n n_chains n_sisters | chains[n_chains][2] kinetochores[n_chains]
//           sisters[n_sisters][2] (i32) | poles[2][3] positions[n][3] (f64)
//   output: "GCO1" n 0 | positions[n][3] (f64) | time, 0, 0, 0, 0, 0, mean_energy (f64)
#pragma once

#include <cmath>
#include <cstdint>
#include <cstdio>
#include <ctime>
#include <iostream>
#include <stdexcept>
#include <string>
#include <vector>

#include "../include/gc_md.hpp"

namespace mitotic {

struct config {
    double temperature = 1, timestep = 1e-4;
    std::int64_t anaphase_steps = 200000, telophase_steps = 50000, prometaphase_steps = 400000;
    std::int64_t sampling_interval = 1000, logging_interval = 10000;
    double core_diameter = 0.3, core_repulsion = 2, bond_length = 0.3, bond_spring = 1000;
    double bending_energy = 1;
    bool penalize_centromere_bending = false;
    double core_mobility = 0.1, sister_separation = 0.3;
    double kfiber_decay_rate_prometaphase = 1, kfiber_decay_rate_anaphase = 1;
    double kfiber_length_prometaphase = 0, kfiber_length_anaphase = 0;
    double polar_ejection_force = 0, polar_ejection_cross_section = 0;
    double anaphase_spindle_shift[3] = {0, 2, 0};
    double telophase_packing_radius = 1.5, telophase_packing_spring = 100;
    double telophase_bond_spring_multiplier = 1, telophase_bending_energy_multiplier = 1;
};

struct design {
    std::int32_t n = 0, n_chains = 0, n_sisters = 0;
    std::vector<std::int32_t> chains, kinetochores, sisters;
    double poles[2][3] = {};
    std::vector<double> positions;
};

template<typename T>
void read_array(FILE* f, T* out, std::size_t count)
{
    if (count && fread(out, sizeof(T), count, f) != count) throw std::runtime_error("short input file");
}

inline design load_design(std::string const& path)
{
    FILE* f = fopen(path.c_str(), "rb");
    if (!f) throw std::runtime_error("cannot open " + path);
    design d;
    char magic[4];
    std::int32_t head[3];
    try {
        read_array(f, magic, 4);
        if (std::string(magic, 4) != "GCM1") throw std::runtime_error("not a GCM1 file");
        read_array(f, head, 3);
        d.n = head[0];
        d.n_chains = head[1];
        d.n_sisters = head[2];
        d.chains.resize(2 * (std::size_t) d.n_chains);
        d.kinetochores.resize((std::size_t) d.n_chains);
        d.sisters.resize(2 * (std::size_t) d.n_sisters);
        d.positions.resize(3 * (std::size_t) d.n);
        read_array(f, d.chains.data(), d.chains.size());
        read_array(f, d.kinetochores.data(), d.kinetochores.size());
        read_array(f, d.sisters.data(), d.sisters.size());
        read_array(f, &d.poles[0][0], 6);
        read_array(f, d.positions.data(), d.positions.size());
    } catch (...) {
        fclose(f);
        throw;
    }
    fclose(f);
    return d;
}

inline void save_state(std::string const& path, gcmd::system& sys, double mean_energy)
{
    auto pos = sys.view_positions();
    auto const ctx = sys.context();
    FILE* f = fopen(path.c_str(), "wb");
    if (!f) throw std::runtime_error("cannot write " + path);
    std::int32_t head[2] = {(std::int32_t) pos.size(), 0};
    double tail[7] = {ctx.time, 0, 0, 0, 0, 0, mean_energy};
    fwrite("GCO1", 1, 4, f);
    fwrite(head, sizeof(head), 1, f);
    fwrite(&pos.begin()->x, sizeof(double), 3 * pos.size(), f);
    fwrite(tail, sizeof(tail), 1, f);
    fclose(f);
}

// particles: uniform mobility, positions of the previous stage
// (SIM/stage_anatelophase/simulation_driver.cpp:62-75, stage_prometaphase/...:60-73)
inline void setup_particles(gcmd::system& sys, design const& d, config const& c)
{
    for (std::int32_t i = 0; i < d.n; i++) {
        sys.add_particle({.mobility = c.core_mobility,
                          .position = {d.positions[3 * i], d.positions[3 * i + 1], d.positions[3 * i + 2]}});
    }
}

// chain bonds over every chain (add_bonded_range per chain)
inline gcmd::bonded_pairs chain_bonds(design const& d, double spring, double length)
{
    auto b = gcmd::semispring_bonds(spring, length);
    for (std::int32_t c = 0; c < d.n_chains; c++) b.add_bonded_range(d.chains[2 * c], d.chains[2 * c + 1]);
    return b;
}

// bending over every chain; the kinetochore bead splits its chain unless
// penalize_centromere_bending (SIM/stage_anatelophase/simulation_driver.cpp:125-135)
inline gcmd::bonded_triples chain_bending(design const& d, config const& c, double energy)
{
    gcmd::bonded_triples t;
    t.bending_energy = energy;
    for (std::int32_t k = 0; k < d.n_chains; k++) {
        const std::int32_t s = d.chains[2 * k], e = d.chains[2 * k + 1], kin = d.kinetochores[k];
        if (c.penalize_centromere_bending) {
            t.add_bonded_range(s, e);
        } else {
            t.add_bonded_range(s, kin);
            t.add_bonded_range(kin + 1, e);
        }
    }
    return t;
}

// spring of the fibre of chain k: decay_rate / (core_mobility / chain length)
// (SIM/stage_anatelophase/simulation_driver.cpp:156-170,
//  SIM/common/forcefield/kinetochore_fiber_forcefield.cpp:46-53)
inline double kfiber_spring(design const& d, config const& c, std::int32_t k, double decay_rate)
{
    return decay_rate / (c.core_mobility / (double) (d.chains[2 * k + 1] - d.chains[2 * k]));
}

inline double print_progress(gcmd::system& sys, std::string const& phase, gcmd::step step)
{
    const double mean_energy = sys.compute_energy() / (double) sys.particle_count();
    std::clog << "[" << phase << "] " << std::time(nullptr) << "\t" << step << "\tE: " << mean_energy << "\n";
    return mean_energy;
}

} // namespace mitotic

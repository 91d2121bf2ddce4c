// prometaphase_driver.cpp -- the reference's prometaphase stage driver (sister chromatids
// congressing between two poles), re-hosted on the B200 path through include/gc_md.hpp.
//
// Structure follows SIM/stage_prometaphase/simulation_driver.cpp (SIM = src/simulation of the
// reference):
//   setup_particles / setup_forcefield   :60-183   repulsion, chain bonds + bending, sister
//                                                  bonds between the kinetochores (spring =
//                                                  bond_spring, :106), one k-fibre force field
//                                                  per pole, polar ejection (force flux) per pole
//   run                                  :185-237  prometaphase steps, progress + frame cadence
// Files: mitotic_common.hpp.
//
//   prometaphase_driver <in> <out> <steps> <seed> [interval] [polar_ejection_force cross_section]
#include "mitotic_common.hpp"

namespace {

class simulation_driver {
public:
    simulation_driver(mitotic::design d, mitotic::config c, std::uint64_t seed)
        : _design(std::move(d)), _config(c), _seed(seed)
    {
        mitotic::setup_particles(_system, _design, _config);
        setup_forcefield();
    }

    void run()
    {
        _system.set_seeds({_seed});
        gcmd::simulate_brownian_dynamics(_system, {
            .temperature = _config.temperature,
            .timestep = _config.timestep,
            .steps = _config.prometaphase_steps,
            .callback = [this](gcmd::step step) {
                _mean_energy = mitotic::print_progress(_system, "prometaphase", step);
            },
            .callback_interval = _config.sampling_interval,
        });
    }

    void save(std::string const& path) { mitotic::save_state(path, _system, _mean_energy); }

private:
    void setup_forcefield()
    {
        auto const& c = _config;
        auto const& d = _design;
        _system.add_forcefield(gcmd::neighbor_softcore{c.core_repulsion, c.core_diameter});
        _system.add_forcefield(mitotic::chain_bonds(d, c.bond_spring, c.bond_length));
        _system.add_forcefield(mitotic::chain_bending(d, c, c.bending_energy));

        // sister kinetochores: semispring of the CHAIN spring at the sister separation
        auto sisters = gcmd::semispring_bonds(c.bond_spring, c.sister_separation);
        for (std::int32_t s = 0; s < d.n_sisters; s++) {
            sisters.add_bonded_pair(d.kinetochores[d.sisters[2 * s]], d.kinetochores[d.sisters[2 * s + 1]]);
        }
        _system.add_forcefield(sisters);

        // target chromatids are pulled to pole 0, their sisters to pole 1
        for (int side = 0; side < 2; side++) {
            gcmd::kinetochore_fibers fibers;
            fibers.pole = {d.poles[side][0], d.poles[side][1], d.poles[side][2]};
            for (std::int32_t s = 0; s < d.n_sisters; s++) {
                const std::int32_t chain = d.sisters[2 * s + side];
                fibers.add_kinetochore(d.kinetochores[chain],
                                       mitotic::kfiber_spring(d, c, chain, c.kfiber_decay_rate_prometaphase),
                                       c.kfiber_length_prometaphase);
            }
            _system.add_forcefield(fibers);
        }
        const double b = std::sqrt(c.polar_ejection_cross_section);
        for (int side = 0; side < 2; side++) {
            _system.add_forcefield(gcmd::point_source{GC_SOURCE_FORCE_FLUX,
                                                      {d.poles[side][0], d.poles[side][1], d.poles[side][2]},
                                                      c.polar_ejection_force, b});
        }
    }

    mitotic::design _design;
    mitotic::config _config;
    gcmd::system _system;
    std::uint64_t _seed;
    double _mean_energy = 0;
};

} // namespace

int main(int argc, char** argv)
{
    if (argc < 5) {
        std::cerr << "usage: prometaphase_driver <in> <out> <steps> <seed> [interval] "
                     "[polar_ejection_force cross_section]\n";
        return 1;
    }
    try {
        mitotic::config config;
        config.prometaphase_steps = std::stoll(argv[3]);
        if (argc > 5) config.sampling_interval = config.logging_interval = std::stoll(argv[5]);
        if (argc > 7) {
            config.polar_ejection_force = std::stod(argv[6]);
            config.polar_ejection_cross_section = std::stod(argv[7]);
        }
        simulation_driver driver(mitotic::load_design(argv[1]), config, std::stoull(argv[4]));
        driver.run();
        driver.save(argv[2]);
    } catch (std::exception const& e) {
        std::cerr << "error: " << e.what() << "\n";
        return 1;
    }
    return 0;
}

// anatelophase_driver.cpp -- the reference's anaphase / telophase stage driver, re-hosted on
// the B200 path through include/gc_md.hpp.
//
// Structure follows SIM/stage_anatelophase/simulation_driver.cpp (SIM = src/simulation of the
// reference):
//   setup_particles / setup_forcefield   :62-190   repulsion, chain bonds + bending (an anaphase
//                                                  and a telophase set), dragging k-fibres,
//                                                  telophase packing source
//   run_dragging_stage                   :240-262  add the dragging force field, anaphase steps
//   run_packing_stage                    :264-310  swap dragging -> packing and the anaphase
//                                                  bonded terms -> the telophase ones
// Force fields are registered once; "add / remove" is handle.enable() / disable()
// (gc_set_term_enabled), which is what md::system::add_forcefield / remove_forcefield amount to
// on a device-resident term table.  Files: mitotic_common.hpp.
//
//   anatelophase_driver <in> <out> <anaphase_steps> <telophase_steps> <seed_ana> <seed_telo> [interval]
#include "mitotic_common.hpp"

namespace {

class simulation_driver {
public:
    simulation_driver(mitotic::design d, mitotic::config c, std::uint64_t seed_ana, std::uint64_t seed_telo)
        : _design(std::move(d)), _config(c), _seed_ana(seed_ana), _seed_telo(seed_telo)
    {
        mitotic::setup_particles(_system, _design, _config);
        setup_forcefield();
    }

    void run()
    {
        run_dragging_stage();
        run_packing_stage();
    }

    void save(std::string const& path) { mitotic::save_state(path, _system, _mean_energy); }

private:
    void setup_forcefield()
    {
        auto const& c = _config;
        _system.add_forcefield(gcmd::neighbor_softcore{c.core_repulsion, c.core_diameter});
        _ana_bonds = _system.add_forcefield(mitotic::chain_bonds(_design, c.bond_spring, c.bond_length));
        _ana_bend = _system.add_forcefield(mitotic::chain_bending(_design, c, c.bending_energy));
        _telo_bonds = _system.add_forcefield(
            mitotic::chain_bonds(_design, c.bond_spring * c.telophase_bond_spring_multiplier, c.bond_length));
        _telo_bend = _system.add_forcefield(
            mitotic::chain_bending(_design, c, c.bending_energy * c.telophase_bending_energy_multiplier));

        gcmd::kinetochore_fibers fibers;
        fibers.pole = {c.anaphase_spindle_shift[0], c.anaphase_spindle_shift[1], c.anaphase_spindle_shift[2]};
        for (std::int32_t k = 0; k < _design.n_chains; k++) {
            fibers.add_kinetochore(_design.kinetochores[k],
                                   mitotic::kfiber_spring(_design, c, k, c.kfiber_decay_rate_anaphase),
                                   c.kfiber_length_anaphase);
        }
        _dragging = _system.add_forcefield(fibers);
        _packing = _system.add_forcefield(gcmd::point_source{GC_SOURCE_SEMISPRING, {0, 0, 0},
                                                             c.telophase_packing_spring,
                                                             c.telophase_packing_radius});
        _telo_bonds.disable();
        _telo_bend.disable();
        _dragging.disable();
        _packing.disable();
    }

    void run_dragging_stage()
    {
        _dragging.enable();
        _system.set_seeds({_seed_ana});
        simulate("anaphase", _config.anaphase_steps);
    }

    void run_packing_stage()
    {
        _dragging.disable();
        _packing.enable();
        _ana_bonds.disable();
        _ana_bend.disable();
        _telo_bonds.enable();
        _telo_bend.enable();
        _system.set_seeds({_seed_telo});
        simulate("telophase", _config.telophase_steps);
    }

    void simulate(std::string const& phase, gcmd::step steps)
    {
        gcmd::simulate_brownian_dynamics(_system, {
            .temperature = _config.temperature,
            .timestep = _config.timestep,
            .steps = steps,
            .callback = [this, phase](gcmd::step step) {
                _mean_energy = mitotic::print_progress(_system, phase, step);
            },
            .callback_interval = _config.sampling_interval,
        });
    }

    mitotic::design _design;
    mitotic::config _config;
    gcmd::system _system;
    gcmd::forcefield_handle _ana_bonds, _ana_bend, _telo_bonds, _telo_bend, _dragging, _packing;
    std::uint64_t _seed_ana, _seed_telo;
    double _mean_energy = 0;
};

} // namespace

int main(int argc, char** argv)
{
    if (argc < 7) {
        std::cerr << "usage: anatelophase_driver <in> <out> <anaphase_steps> <telophase_steps> "
                     "<seed_ana> <seed_telo> [interval]\n";
        return 1;
    }
    try {
        mitotic::config config;
        config.anaphase_steps = std::stoll(argv[3]);
        config.telophase_steps = std::stoll(argv[4]);
        if (argc > 7) config.sampling_interval = config.logging_interval = std::stoll(argv[7]);
        simulation_driver driver(mitotic::load_design(argv[1]), config, std::stoull(argv[5]),
                                 std::stoull(argv[6]));
        driver.run();
        driver.save(argv[2]);
    } catch (std::exception const& e) {
        std::cerr << "error: " << e.what() << "\n";
        return 1;
    }
    return 0;
}

// gc_md.hpp -- header-only C++ host adapter over the C-ABI (gc_b200.h).
//
// It mirrors the slice of micromd's API that 3d-genome-cycle's stage drivers
// program against ("calls the drivers actually make"), so that a
// driver written like src/simulation/stage_*/simulation_driver*.cpp keeps its
// structure and links against libgc_b200.so instead of evaluating forces on the
// host.  Names follow md:: (MD/ = submodules/github.com/snsinfu/micromd/include/md/):
//
//   md::system                       -> gcmd::system            (MD/system.hpp:60-272)
//   system.add_particle / view_*     -> same names
//   md::make_*_forcefield(potential) -> system.add_forcefield(gcmd::<kind>{...})
//                                       returning a handle with enable()/disable()
//                                       (system.add_forcefield / remove_forcefield,
//                                        MD/system.hpp:200-228)
//   md::simulate_brownian_dynamics   -> gcmd::simulate_brownian_dynamics
//                                       (MD/simulation/brownian_dynamics.hpp:24-73)
//   contact_map                      -> system.contact_map_*   (SIM/stage_interphase/contact_map.hpp)
//
// Two deliberate differences, both forced by keeping positions resident on the GPU:
//   * potentials are an enumerated closed set with parameter blocks, not C++
//     lambdas (a lambda cannot cross a C ABI);
//   * the per-step part of the interphase callback (core/bond scale schedule, wall
//     ODE, contact-map cadence) is DECLARED on the system (set_schedule, wall ODE
//     parameters, contact_map_configure) and executed on device every step; the
//     host callback runs every `callback_interval` steps (the logging / sampling
//     cadence), with the same 1-based step numbers the reference passes.
// Errors: the reference throws C++ exceptions; so does this adapter
// (gcmd::error, derived from std::runtime_error) from the C-ABI status codes.
#ifndef GC_MD_HPP
#define GC_MD_HPP

#include <cstddef>
#include <cstdint>
#include <functional>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "gc_b200.h"

namespace gcmd {

using scalar = double;
using index = std::size_t;
using step = std::int64_t;

struct vector {
    scalar x = 0, y = 0, z = 0;
};
using point = vector;

struct error : std::runtime_error {
    int code;
    error(int c, std::string const& msg) : std::runtime_error(msg), code(c) {}
};

template<typename T>
struct array_view {
    T* ptr = nullptr;
    index count = 0;
    T* begin() const { return ptr; }
    T* end() const { return ptr + count; }
    T& operator[](index i) const { return ptr[i]; }
    index size() const { return count; }
};

struct basic_particle_data {
    scalar mobility = 1;
    point position = {};
    scalar a_factor = 0;   // particle_data_attribute (SIM/common/particle_data.hpp:6-13)
    scalar b_factor = 0;
};

// ---- force-field descriptions (the closed set of SURVEY.md 8b) ----------------

// a * softcore<2,3>(eps_a, sigma_a s) + b * softcore<8,3>(eps_b, sigma_b s)
// (SIM/stage_interphase/simulation_driver_forcefield.cpp:20-53)
struct neighbor_ab_softcore {
    scalar eps_a, sigma_a, eps_b, sigma_b;
    bool scale_by_core = true;
};
// softcore<2,3>{eps, sigma} (SIM/stage_anatelophase/simulation_driver.cpp:77-91)
struct neighbor_softcore {
    scalar eps, sigma;
};
// apply_cutoff(softwell<6>{eps, decay}, cutoff) on set_neighbor_targets(targets)
struct neighbor_softwell {
    scalar eps, decay, cutoff;
    std::vector<std::int32_t> targets;
};
// make_bonded_pairwise_forcefield + add_bonded_pair / add_bonded_range
struct bonded_pairs {
    gc_bond_params params{};
    std::vector<std::int32_t> pairs;   // flat (i, j)
    bonded_pairs& add_bonded_pair(index i, index j)
    {
        pairs.push_back((std::int32_t) i);
        pairs.push_back((std::int32_t) j);
        return *this;
    }
    // (i, i + 1) for i in [start, end - 1)  (MD/forcefield/bonded_pairwise_forcefield.hpp:45-52)
    bonded_pairs& add_bonded_range(index start, index end)
    {
        for (index i = start; i + 1 < end; i++) add_bonded_pair(i, i + 1);
        return *this;
    }
};
inline bonded_pairs semispring_bonds(scalar K, scalar l)
{
    bonded_pairs b;
    b.params = {GC_BOND_SEMISPRING, 0, 0, 0, K, 0, l, 0};
    return b;
}
// K = (a K_a + b K_b) / s^2, l = (a l_a + b l_b) s with s = bond_scale(t)
inline bonded_pairs ab_mixed_bonds(gc_bond_kind kind, scalar K_a, scalar K_b, scalar l_a, scalar l_b,
                                   bool scale = true)
{
    bonded_pairs b;
    b.params = {(std::int32_t) kind, 1, scale ? 1 : 0, 0, K_a, K_b, l_a, l_b};
    return b;
}
// make_bonded_triplewise_forcefield(cosine_bending_potential{energy})
struct bonded_triples {
    scalar bending_energy = 0;
    std::vector<std::int32_t> triples;   // flat (i, j, k)
    bonded_triples& add_bonded_range(index start, index end)
    {
        for (index i = start; i + 2 < end; i++) {
            triples.push_back((std::int32_t) i);
            triples.push_back((std::int32_t) (i + 1));
            triples.push_back((std::int32_t) (i + 2));
        }
        return *this;
    }
};
// make_ellipsoid_inward_forcefield + make_ellipsoid_outward_forcefield + wall ODE
using ellipsoid_wall = gc_wall_params;
struct point_source {
    gc_source_kind kind;
    point source;
    scalar p0, p1;   // semispring: K, l;  force flux: f, b
};
// kinetochore_fiber_forcefield (SIM/common/forcefield/kinetochore_fiber_forcefield.hpp:8-30)
struct kinetochore_fibers {
    point pole;
    std::vector<std::int32_t> beads;
    std::vector<scalar> spring, length;
    void add_kinetochore(index bead, scalar K, scalar l)
    {
        beads.push_back((std::int32_t) bead);
        spring.push_back(K);
        length.push_back(l);
    }
};

class system;

// What md::system::add_forcefield returns (a shared_ptr there): enough to
// remove / re-add the component at run time.
class forcefield_handle {
public:
    forcefield_handle() = default;
    int term() const { return term_; }
    void enable(bool on = true);
    void disable() { enable(false); }

private:
    friend class system;
    forcefield_handle(system* s, int t) : sys_(s), term_(t) {}
    system* sys_ = nullptr;
    int term_ = -1;
};

class system {
public:
    explicit system(int device = 0, int replicas = 1) : device_(device), replicas_(replicas) {}
    system(system const&) = delete;
    system& operator=(system const&) = delete;
    ~system()
    {
        if (ctx_) gc_destroy(ctx_);
    }

    // ---- particles (before the first force field / run) -----------------------
    index add_particle(basic_particle_data const& data = {})
    {
        if (ctx_) throw error(GC_ERR_STATE, "add_particle after the device context exists");
        particles_.push_back(data);
        positions_.push_back(data.position);
        return particles_.size() - 1;
    }
    index particle_count() const { return particles_.size(); }
    int replica_count() const { return replicas_; }

    // Mutable host view; writing through it makes the host copy authoritative
    // (uploaded before the next device operation), as drivers std::copy initial
    // structures into view_positions().
    array_view<point> view_positions(int replica = 0)
    {
        pull(replica);
        host_dirty_ = true;
        host_replica_ = replica;
        return {positions_.data(), positions_.size()};
    }
    array_view<point const> view_positions(int replica = 0) const
    {
        const_cast<system*>(this)->pull(replica);
        return {positions_.data(), positions_.size()};
    }

    // ---- force fields -----------------------------------------------------------
    forcefield_handle add_forcefield(neighbor_ab_softcore const& f)
    {
        return wrap(gc_add_pair_softcore_ab(ctx(), f.eps_a, f.sigma_a, f.eps_b, f.sigma_b,
                                            f.scale_by_core));
    }
    forcefield_handle add_forcefield(neighbor_softcore const& f)
    {
        return wrap(gc_add_pair_softcore(ctx(), f.eps, f.sigma));
    }
    forcefield_handle add_forcefield(neighbor_softwell const& f)
    {
        return wrap(gc_add_pair_softwell(ctx(), f.eps, f.decay, f.cutoff, (std::int32_t) f.targets.size(),
                                         f.targets.data()));
    }
    forcefield_handle add_forcefield(bonded_pairs const& f)
    {
        return wrap(gc_add_bonds(ctx(), &f.params, (std::int32_t) (f.pairs.size() / 2), f.pairs.data()));
    }
    forcefield_handle add_forcefield(bonded_triples const& f)
    {
        return wrap(gc_add_triples(ctx(), f.bending_energy, (std::int32_t) (f.triples.size() / 3),
                                   f.triples.data()));
    }
    forcefield_handle add_forcefield(ellipsoid_wall const& f) { return wrap(gc_set_wall(ctx(), &f)); }
    forcefield_handle add_forcefield(point_source const& f)
    {
        double pt[3] = {f.source.x, f.source.y, f.source.z};
        return wrap(gc_add_point_source(ctx(), f.kind, pt, f.p0, f.p1));
    }
    forcefield_handle add_forcefield(kinetochore_fibers const& f)
    {
        double pole[3] = {f.pole.x, f.pole.y, f.pole.z};
        return wrap(gc_add_kfiber(ctx(), pole, (std::int32_t) f.beads.size(), f.beads.data(),
                                  f.spring.data(), f.length.data()));
    }

    // ---- per-step driver state executed on device -----------------------------------
    void set_schedule(gc_schedule const& s) { check(gc_set_schedule(ctx(), &s)); }
    gc_context context(int replica = 0)
    {
        gc_context c{};
        check(gc_get_context(ctx(), replica, &c));
        return c;
    }
    void set_context(gc_context const& c, int replica = -1) { check(gc_set_context(ctx(), replica, &c)); }
    void set_seeds(std::vector<std::uint64_t> const& seeds)
    {
        if ((int) seeds.size() != replicas_) throw error(GC_ERR_INVALID, "one seed per replica");
        check(gc_set_seeds(ctx(), seeds.data()));
    }
    void contact_map_configure(scalar distance, step interval, bool scale_by_core, std::size_t capacity = 0)
    {
        check(gc_set_contact_map(ctx(), distance, interval, scale_by_core, capacity));
    }
    void contact_map_update() { push(); check(gc_contacts_update(ctx())); }
    void contact_map_clear() { check(gc_contacts_clear(ctx())); }
    // contact_map::accumulate(): (i, j, count) sorted by (i, j); replica -1 = ensemble sum
    std::vector<std::uint32_t> contact_map_accumulate(int replica = 0)
    {
        std::size_t count = 0;
        check(gc_contacts_fetch(ctx(), replica, nullptr, 0, &count));
        std::vector<std::uint32_t> ijv(3 * count);
        if (count) check(gc_contacts_fetch(ctx(), replica, ijv.data(), count, &count));
        return ijv;
    }

    // simulation_store::save_positions' payload: the frame rounded to `fraction_bits`
    // significant bits, packed f32 [n][3] (SIM/common/simulation_store.cpp:192-216)
    std::vector<float> quantized_frame(int replica = 0, int fraction_bits = 16)
    {
        push();
        std::vector<float> xyz(3 * static_cast<std::size_t>(particle_count()));
        check(gc_get_frame(ctx(), replica, fraction_bits, xyz.data()));
        return xyz;
    }

    // ---- evaluation -------------------------------------------------------------------
    scalar compute_potential_energy(int replica = 0, std::uint32_t mask = GC_MASK_ALL)
    {
        push();
        scalar e = 0;
        check(gc_compute_energy(ctx(), replica, mask, &e));
        return e;
    }
    scalar compute_energy(int replica = 0) { return compute_potential_energy(replica); }   // no kinetic part
    void compute_force(array_view<vector> forces, int replica = 0, std::uint32_t mask = GC_MASK_ALL)
    {
        push();
        if (forces.size() != particle_count()) throw error(GC_ERR_INVALID, "force view size");
        static_assert(sizeof(vector) == 3 * sizeof(double), "vector is three doubles");
        check(gc_compute_force(ctx(), replica, mask, &forces.ptr->x));
    }

    gc_ctx* handle() { return ctx(); }
    void check(int rc)
    {
        if (rc < 0) throw error(rc, std::string("gc_b200: ") + gc_last_error(ctx_));
    }

private:
    friend class forcefield_handle;
    friend void simulate_brownian_dynamics(system&, struct brownian_dynamics_config const&);

    gc_ctx* ctx()
    {
        if (!ctx_) {
            int rc = gc_create((std::int32_t) particles_.size(), replicas_, device_, &ctx_);
            if (rc != GC_OK) {
                std::string msg = ctx_ ? gc_last_error(ctx_) : "no CUDA device (there is no CPU fallback)";
                throw error(rc, "gc_create: " + msg);
            }
            std::vector<double> mob(particles_.size()), a(particles_.size()), b(particles_.size());
            for (index i = 0; i < particles_.size(); i++) {
                mob[i] = particles_[i].mobility;
                a[i] = particles_[i].a_factor;
                b[i] = particles_[i].b_factor;
            }
            check(gc_set_bead_attrs(ctx_, mob.data(), a.data(), b.data()));
            host_dirty_ = true;
            host_replica_ = -1;          // initial structure goes to every replica
        }
        return ctx_;
    }
    forcefield_handle wrap(int term)
    {
        check(term);
        return forcefield_handle(this, term);
    }
    void push()   // host -> device if the host copy was written
    {
        ctx();
        if (host_dirty_) {
            check(gc_set_positions(ctx_, host_replica_, &positions_.data()->x));
            host_dirty_ = false;
            host_valid_ = host_replica_ >= 0 ? host_replica_ : 0;
        }
    }
    void pull(int replica)   // device -> host if the device moved on
    {
        if (!ctx_ || host_dirty_) return;
        if (host_valid_ != replica) {
            check(gc_get_positions(ctx_, replica, &positions_.data()->x));
            host_valid_ = replica;
        }
    }
    void device_advanced() { host_valid_ = -2; }

    int device_, replicas_;
    gc_ctx* ctx_ = nullptr;
    std::vector<basic_particle_data> particles_;
    std::vector<point> positions_;
    bool host_dirty_ = false;
    int host_replica_ = -1;
    int host_valid_ = -2;   // which replica the host copy mirrors (-2: none)
};

inline void forcefield_handle::enable(bool on)
{
    if (!sys_) throw error(GC_ERR_STATE, "empty force-field handle");
    sys_->check(gc_set_term_enabled(sys_->ctx(), term_, on));
}

// md::brownian_dynamics_config (MD/simulation/brownian_dynamics.hpp:24-47) plus the
// host-visit cadence.
struct brownian_dynamics_config {
    scalar temperature = 1;
    scalar timestep = 1;
    scalar spacestep = 0;         // != 0 selects the adaptive time stepper
    step steps = 1;
    std::function<void(step)> callback;   // called with 1-based step numbers
    step callback_interval = 0;   // 0: once, after the last step
};

// Synchronous like the reference: returns when all steps are done.  The Philox key
// of each replica is set with system::set_seeds (the reference seeds one sfc64
// stream per call with config.seed).
inline void simulate_brownian_dynamics(system& sys, brownian_dynamics_config const& config)
{
    sys.push();
    step done = 0;
    const step every = config.callback_interval > 0 ? config.callback_interval : config.steps;
    while (done < config.steps) {
        step chunk = every - done % every;
        if (chunk > config.steps - done) chunk = config.steps - done;
        gc_run_config rc{};
        rc.steps = chunk;
        rc.temperature = config.temperature;
        rc.timestep = config.timestep;
        rc.spacestep = config.spacestep;
        rc.restart = done == 0;
        sys.check(gc_run(sys.handle(), &rc));
        sys.device_advanced();
        done += chunk;
        if (config.callback && (done % every == 0 || done == config.steps)) config.callback(done);
    }
}

} // namespace gcmd

#endif // GC_MD_HPP

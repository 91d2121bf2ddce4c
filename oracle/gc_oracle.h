/*
 * gc_oracle.h -- C interface shared by the two CPU oracles.  TEST
 * INFRASTRUCTURE ONLY: nothing outside tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may link or load it.
 *
 *   oracle/_ref/libgc_ref.so   ("reference"): micromd headers compiled
 *       UNMODIFIED from /root/reference (ref_harness.cpp + the reference's own
 *       contact_map.cpp and kinetochore_fiber_forcefield.cpp compiled in place).
 *   oracle/libgc_port.so       ("port"): gc_oracle_port.cpp, an FP64
 *       restatement with no dependency on the reference tree.
 *
 * Both export exactly these symbols so every parity test can be pointed at
 * either.  The shapes mirror include/gc_b200.h (single replica, FP64).
 */
#ifndef GC_ORACLE_H
#define GC_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#include "../include/gc_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct gco_sys gco_sys;

typedef struct gco_run_config {
    int64_t steps;
    double temperature;
    double timestep;
    double spacestep;
    uint64_t seed;        /* sfc64 seed when noise == NULL */
    const double *noise;  /* [steps][n][3] injected standard normals or NULL */
    /* interphase driver callback (schedule, wall ODE, contact cadence) after
       every step: SIM/stage_interphase/simulation_driver_interphase.cpp:16-53 */
    int32_t interphase_callback;
    /* also evaluate the energy every `energy_interval` steps as the driver's
       logging does (0 = never); value returned through last_energy */
    int32_t energy_interval;
    /* optional dumps, NULL to skip */
    double *traj_out;     /* [steps][n][3] positions after each step */
    double *ctx_out;      /* [steps][8]: time, sx, sy, sz, core, bond, dt, reserved */
    double *force_out;    /* [steps][n][3] forces evaluated at the start of each step */
    int64_t step_offset;  /* callback time = (step_offset + step) * timestep */
    /* 0 = unmodified md::simulate_brownian_dynamics (needs noise == NULL and no
       force dump); 1 = restated simulate_step (injectable noise, force dump).
       The port has only its own stepper and ignores this. */
    int32_t stepper;
    int32_t reserved;
    double last_energy;   /* out: mean energy per bead at the last energy step */
} gco_run_config;

const char *gco_kind(void); /* "reference" or "port" */

gco_sys *gco_create(int32_t n_beads);
void gco_destroy(gco_sys *s);

void gco_set_bead_attrs(gco_sys *s, const double *mobility, const double *a, const double *b);
void gco_set_positions(gco_sys *s, const double *xyz);
void gco_get_positions(gco_sys *s, double *xyz);

int gco_add_pair_softcore_ab(gco_sys *s, double eps_a, double sigma_a, double eps_b,
                             double sigma_b, int32_t scale_by_core);
int gco_add_pair_softcore(gco_sys *s, double eps, double sigma);
int gco_add_pair_softwell(gco_sys *s, double eps, double decay, double cutoff,
                          int32_t n_targets, const int32_t *targets);
int gco_add_bonds(gco_sys *s, const gc_bond_params *params, int32_t n, const int32_t *pairs);
int gco_add_triples(gco_sys *s, double bending_energy, int32_t n, const int32_t *triples);
int gco_set_wall(gco_sys *s, const gc_wall_params *params);
int gco_add_point_source(gco_sys *s, int32_t kind, const double point[3], double p0, double p1);
int gco_add_kfiber(gco_sys *s, const double pole[3], int32_t n, const int32_t *beads,
                   const double *spring, const double *length);
void gco_set_term_enabled(gco_sys *s, int32_t term, int32_t enabled);

void gco_set_schedule(gco_sys *s, const gc_schedule *schedule);
void gco_set_context(gco_sys *s, const gc_context *in);
void gco_get_context(gco_sys *s, gc_context *out);

void gco_run(gco_sys *s, gco_run_config *config);
void gco_compute_force(gco_sys *s, uint32_t term_mask, double *out);
double gco_compute_energy(gco_sys *s, uint32_t term_mask);
/* returns the number of pairs; writes at most `capacity` of them, sorted */
size_t gco_neighbor_pairs(gco_sys *s, double radius, uint32_t *pairs, size_t capacity);
/* pairs currently held by the Verlet list of pair term `term` (reference
   only; the port returns the exact set at 1.5 * dcut) */
size_t gco_verlet_pairs(gco_sys *s, int32_t term, uint32_t *pairs, size_t capacity);

void gco_set_contact_map(gco_sys *s, double distance, int64_t interval, int32_t scale_by_core);
void gco_contacts_update(gco_sys *s);
size_t gco_contacts_fetch(gco_sys *s, uint32_t *ijv, size_t capacity);
void gco_contacts_clear(gco_sys *s);

/* n standard normals from md::sfc64(seed) + md::ziggurat_normal_distribution
   in the integrator's draw order (reference only; port returns 0). */
size_t gco_reference_normals(uint64_t seed, size_t n, double *out);
/* single-potential probes for the known-answer tests:
   kind: 0 softcore<2,3> 1 softcore<8,3> 2 softwell<6> 3 semispring 4 harmonic
         5 spring 6 force_flux 7 softwell<6>+cutoff(p2)
         8 softcore<P,Q> with P = (int) p2 / 10, Q = (int) p2 % 10 (the (P,Q)
           pairs of MDT/potential/test_softcore_potential.cc:55-76)
   returns energy; force written to f[3] */
double gco_potential_probe(int32_t kind, double p0, double p1, double p2, const double r[3],
                           double f[3]);
/* cosine bending: forces on i, j, k written to f[9]; returns energy */
double gco_bending_probe(double bending_energy, const double rij[3], const double rjk[3],
                         double f[9]);

/* detail::evaluate_point (MD/forcefield/ellipsoid_surface_forcefield.hpp:55-90):
   out = {undefined, delta xyz, strain xyz, implicit} */
void gco_ellipsoid_probe(const double semiaxes[3], const double pt[3], double out[8]);

/* ---- inter-stage transitions and the stored-frame rounding ----------
   resample_chain of SIM/stage_transition/transition_interphase.cpp:13-41: not-a-knot cubic
   spline through the n points of `chain` at t = (k + 1/2) / n, evaluated at
   t = (i + 1/2) / new_n.  The reference build uses the reference's own
   submodules/github.com/snsinfu/cxx-spline/include/spline.hpp; the port restates it. */
void gco_resample_chain(int32_t n, const double *chain, int32_t new_n, double *out);
/* the centroid loop of SIM/stage_transition/transition_prometaphase.cpp:81-98 for ONE
   chromosome: coarse_len windows of `cg` beads starting at positions + 3 * source_start
   (window end = min(start + cg, start + source_len), as written there); writes the target
   chromatid to target[coarse_len][3] and the sister copy, displaced by disp, to sister. */
void gco_coarse_grain(const double *positions, int64_t source_start, int64_t source_len,
                      int64_t cg, int64_t coarse_len, const double disp[3], double *target,
                      double *sister);
/* quantize() of SIM/common/simulation_store.cpp:22-32 followed by the f32 store of
   save_positions (192-216): out[k] = (float) quantize(in[k], bits). */
void gco_quantize_f32(size_t n, const double *in, int32_t bits, float *out);

#ifdef __cplusplus
}
#endif
#endif

/*
 * gc_b200.h -- C-ABI of the B200-native force-evaluation + overdamped-Langevin
 * hot path that sits behind 3d-genome-cycle's src/simulation stage drivers.
 *
 * Every entry point names the reference interface it replaces.  Citations are
 * relative to the reference tree; MD/ = submodules/github.com/snsinfu/micromd/
 * include/md/, SIM/ = src/simulation/.
 *
 * Conventions
 *   - plain pointers and sizes, no C++/torch types; host arrays are copied in/out
 *   - every function returns GC_OK (0) or a negative gc_status; the message is
 *     available from gc_last_error().  Nothing throws across this boundary.
 *     (The reference signals errors with C++ exceptions)
 *   - one gc_ctx = one CUDA device + one stream + R replicas that share bead
 *     attributes and topology and differ in positions / seed / context.  All work is
 *     ordered on that stream (gc_stream); with R > 1 the replayed step graph forks
 *     internally into replica-group branches and joins before the launch returns to it.
 *   - term-adding functions return a term id >= 0 usable with
 *     gc_set_term_enabled() and as bit (1u << id) in force/energy masks.
 */
#ifndef GC_B200_H
#define GC_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct gc_ctx gc_ctx;

typedef enum gc_status {
    GC_OK = 0,
    GC_ERR_INVALID = -1,   /* bad argument */
    GC_ERR_CUDA = -2,      /* CUDA runtime failure (message has the string) */
    GC_ERR_STATE = -3,     /* call out of order (e.g. run before attrs) */
    GC_ERR_NUMERIC = -4,   /* NaN/Inf position detected on device */
    GC_ERR_CAPACITY = -5   /* contact table / output buffer too small */
} gc_status;

#define GC_MAX_TERMS 32
#define GC_MASK_ALL 0xffffffffu

/* Bond potentials (MD/potential/{semispring,harmonic,spring}_potential.hpp). */
typedef enum gc_bond_kind {
    GC_BOND_SEMISPRING = 0, /* F = (K l / r - K) r_vec for r > l, else 0 */
    GC_BOND_HARMONIC = 1,   /* F = -K r_vec */
    GC_BOND_SPRING = 2      /* F = (K l / r - K) r_vec, 0 at r == 0 */
} gc_bond_kind;

/*
 * Parameters of one bonded-pair force field.
 *   mix_ab = 1: K = a K_a + b K_b, l = a l_a + b l_b with a, b the arithmetic
 *               means of the two beads' A/B factors
 *               (SIM/stage_interphase/simulation_driver_forcefield.cpp:56-96);
 *   mix_ab = 0: K = K_a, l = l_a.
 *   scale  = 1: K <- K / s^2, l <- l s with s = bond_scale(t) (same file, 83-85).
 */
typedef struct gc_bond_params {
    int32_t kind;
    int32_t mix_ab;
    int32_t scale;
    int32_t reserved;
    double K_a, K_b, l_a, l_b;
} gc_bond_params;

/* Point-source potentials (MD/forcefield/point_source_forcefield.hpp:34-147). */
typedef enum gc_source_kind {
    GC_SOURCE_SEMISPRING = 0, /* p0 = K, p1 = l  (telophase packing) */
    GC_SOURCE_FORCE_FLUX = 1  /* p0 = f, p1 = b  (polar ejection,
                                 SIM/common/potentials/force_flux_potential.hpp:19-38) */
} gc_source_kind;

/*
 * Ellipsoidal nuclear wall: inward A/B-mixed soft core + outward harmonic,
 * optionally moving under the wall ODE
 * (SIM/stage_interphase/simulation_driver_forcefield.cpp:189-244,
 *  SIM/stage_interphase/simulation_driver_interphase.cpp:79-90).
 * The inward soft-core diameters are HALF of sigma_a / sigma_b and scale with
 * core_scale(t) when scale_by_core is set.
 */
typedef struct gc_wall_params {
    double semiaxes[3];
    double eps_a, sigma_a, eps_b, sigma_b; /* full bead diameters */
    double wall_a, wall_b;                 /* wall_ab_factor */
    double outward_spring;                 /* wall_packing_spring */
    int32_t scale_by_core;
    int32_t ode_enabled;                   /* move semiaxes each step */
    double ode_mobility;                   /* wall_mobility */
    double ode_spring[3];                  /* wall_semiaxes_spring */
} gc_wall_params;

/*
 * Per-step parameter schedule evaluated on device after every step
 * (SIM/stage_interphase/simulation_driver_interphase.cpp:67-76):
 *   core_scale = 1 - (1 - core_init) exp(-t / core_tau), same for bond_scale.
 */
typedef struct gc_schedule {
    int32_t enabled;
    int32_t reserved;
    double core_init, core_tau, bond_init, bond_tau;
} gc_schedule;

/* Mutable per-replica global state (SIM/common/simulation_context.hpp:7-18). */
typedef struct gc_context {
    double time;
    double semiaxes[3];
    double core_scale;
    double bond_scale;
    double contact_distance;   /* distance used by the NEXT contact update */
    double axial_reaction[3];  /* sum of inward+outward stats of last force call */
    double reaction_force;
    double last_timestep;      /* dt used by the last step (adaptive mode) */
    int64_t step;              /* steps done in the current logical run */
    int64_t rebuilds;          /* cell-list rebuilds so far */
} gc_context;

/*
 * One call of md::simulate_brownian_dynamics
 * (MD/simulation/brownian_dynamics.hpp:24-73), possibly split in chunks.
 */
typedef struct gc_run_config {
    int64_t steps;        /* steps to run in this call */
    double temperature;
    double timestep;      /* dt, or dt_max when spacestep != 0 */
    double spacestep;     /* != 0 selects the adaptive time stepper */
    int32_t restart;      /* 1: first chunk of a logical run: zero the Wiener
                             history and the step counter (the reference's
                             integrator state lives for one call only) */
    int32_t reserved;
    /* Test mode: injected standard normals, [steps][n_beads][3] doubles for
       replica 0 .. R-1 consecutively ([R][steps][N][3]); NULL = Philox. */
    const double *noise;
} gc_run_config;

/* ---- lifetime ---------------------------------------------------------- */

/* Replaces md::system construction + add_particle (MD/system.hpp:60-140). */
int gc_create(int32_t n_beads, int32_t n_replicas, int32_t device, gc_ctx **out);
int gc_destroy(gc_ctx *ctx);
const char *gc_last_error(const gc_ctx *ctx);
/* Library build info ("gc_b200 <ver> sm_100a"). */
const char *gc_version(void);

/* ---- beads ------------------------------------------------------------- */

/* view_mobilities() + view(particle_data_attribute)
   (SIM/stage_interphase/simulation_driver_particles.cpp:8-35).
   a_factor / b_factor may be NULL (all zero). Shared by all replicas. */
int gc_set_bead_attrs(gc_ctx *ctx, const double *mobility, const double *a_factor,
                      const double *b_factor);
/* view_positions() write / read, xyz = [n_beads][3]. replica = -1 on set
   broadcasts to every replica. */
int gc_set_positions(gc_ctx *ctx, int32_t replica, const double *xyz);
int gc_get_positions(gc_ctx *ctx, int32_t replica, double *xyz);
int gc_set_positions_f32(gc_ctx *ctx, int32_t replica, const float *xyz);
int gc_get_positions_f32(gc_ctx *ctx, int32_t replica, float *xyz);
/* The same for the whole ensemble in one call: xyz = [R][n_beads][3] f32.  One
   asynchronous copy per direction (the float3 <-> float4 repack runs on the device).
   Buffers from gc_host_alloc (page-locked) are transferred in place; any other host
   pointer goes through the context's page-locked staging buffer.  set returns once the
   copy is queued (a page-locked source must stay untouched until the next synchronising
   call, e.g. gc_run); get returns when xyz is filled. */
int gc_set_positions_all_f32(gc_ctx *ctx, const float *xyz);
int gc_get_positions_all_f32(gc_ctx *ctx, float *xyz);
/* Page-locked host memory for the transfers above (cudaMallocHost / cudaFreeHost). */
int gc_host_alloc(size_t bytes, void **out);
int gc_host_free(void *p);
/* Trajectory frame as simulation_store::save_positions stores it
   (SIM/common/simulation_store.cpp:22-32,192-216): every coordinate rounded to
   fraction_bits significant binary digits (16 in the reference) on the device and
   returned as packed f32 [n][3]; replica = -1 returns all replicas, [R][n][3]. */
int gc_get_frame(gc_ctx *ctx, int32_t replica, int32_t fraction_bits, float *xyz);
/* The same frame WITHOUT stalling the step loop (the reference writes frames
   synchronously, SIM/common/simulation_store.cpp:192-216): gc_frame_begin quantises on
   the launch stream and starts the device-to-host copy on a second stream into a
   page-locked buffer; steps queued afterwards overlap it.  gc_frame_end waits for that
   copy only and hands the frame out.  One frame in flight per context. */
int gc_frame_begin(gc_ctx *ctx, int32_t replica, int32_t fraction_bits);
int gc_frame_end(gc_ctx *ctx, float *xyz);
/* Philox key per replica (the drivers draw it from mt19937_64(design.seed),
   SIM/stage_interphase/simulation_driver_interphase.cpp:61). */
int gc_set_seeds(gc_ctx *ctx, const uint64_t *seeds);

/* ---- force-field terms ------------------------------------------------- */

/* make_neighbor_pairwise_forcefield(a*softcore<2,3> + b*softcore<8,3>)
   (SIM/stage_interphase/simulation_driver_forcefield.cpp:20-53). */
int gc_add_pair_softcore_ab(gc_ctx *ctx, double eps_a, double sigma_a, double eps_b,
                            double sigma_b, int32_t scale_by_core);
/* make_neighbor_pairwise_forcefield(softcore<2,3>{eps, sigma})
   (SIM/stage_anatelophase/simulation_driver.cpp:77-91). */
int gc_add_pair_softcore(gc_ctx *ctx, double eps, double sigma);
/* apply_cutoff(softwell<6>) on set_neighbor_targets(targets)
   (SIM/stage_interphase/simulation_driver_forcefield.cpp:162-185). */
int gc_add_pair_softwell(gc_ctx *ctx, double eps, double decay, double cutoff,
                         int32_t n_targets, const int32_t *targets);
/* make_bonded_pairwise_forcefield + add_bonded_pair/add_bonded_range
   (MD/forcefield/bonded_pairwise_forcefield.hpp:35-132); pairs = [n][2]. */
int gc_add_bonds(gc_ctx *ctx, const gc_bond_params *params, int32_t n,
                 const int32_t *pairs);
/* make_bonded_triplewise_forcefield(cosine_bending_potential)
   (MD/forcefield/bonded_triplewise_forcefield.hpp:35-149); triples = [n][3]. */
int gc_add_triples(gc_ctx *ctx, double bending_energy, int32_t n, const int32_t *triples);
/* make_ellipsoid_inward_forcefield + make_ellipsoid_outward_forcefield. */
int gc_set_wall(gc_ctx *ctx, const gc_wall_params *params);
/* make_point_source_forcefield(pot).set_point_source(pt), all beads. */
int gc_add_point_source(gc_ctx *ctx, int32_t kind, const double point[3], double p0,
                        double p1);
/* kinetochore_fiber_forcefield
   (SIM/common/forcefield/kinetochore_fiber_forcefield.cpp:8-84):
   spring K[k] = decay_rate / mobility, rest length[k], bead[k] -> pole. */
int gc_add_kfiber(gc_ctx *ctx, const double pole[3], int32_t n, const int32_t *beads,
                  const double *spring, const double *length);
/* system.add_forcefield / remove_forcefield at run time
   (SIM/stage_anatelophase/simulation_driver.cpp:244,277-285). */
int gc_set_term_enabled(gc_ctx *ctx, int32_t term, int32_t enabled);

/* ---- schedule / context ------------------------------------------------ */

int gc_set_schedule(gc_ctx *ctx, const gc_schedule *schedule);
/* replica = -1 on set applies to all. Only time, semiaxes, core_scale,
   bond_scale and contact_distance are taken from `in`. */
int gc_set_context(gc_ctx *ctx, int32_t replica, const gc_context *in);
int gc_get_context(gc_ctx *ctx, int32_t replica, gc_context *out);

/* Resume a logical run at `step` (replica = -1: all): the next gc_run with restart = 0
   continues with step + 1, i.e. schedule time (step + 1) dt and Philox counter step + 1.
   The reference cannot resume inside a stage (its checkpoint is the stage's last frame,
   SURVEY section 5); here a saved frame + context + step is a mid-stage checkpoint. */
int gc_set_step(gc_ctx *ctx, int32_t replica, int64_t step);
/* Contexts of all replicas in one copy: out = [R]. */
int gc_get_contexts(gc_ctx *ctx, gc_context *out);

/* ---- dynamics ---------------------------------------------------------- */

/* md::simulate_brownian_dynamics incl. the per-step driver callback body
   (schedule, wall ODE, contact-map cadence) executed on device. */
int gc_run(gc_ctx *ctx, const gc_run_config *config);
/* Cell-list skin as a fraction of the pair cutoff (0 = rebuild every step). */
int gc_set_skin(gc_ctx *ctx, double skin_fraction);
/* Standard normals the device would draw for `step` (1-based) of replica:
   out = [n_beads][3] floats.  Diagnostic for the Philox stream. */
int gc_debug_noise(gc_ctx *ctx, int32_t replica, int64_t step, float *out);

/* md::system::compute_force (MD/system.hpp:264-272); out = [n_beads][3].
   Also refreshes the wall statistics returned by gc_get_context. */
int gc_compute_force(gc_ctx *ctx, int32_t replica, uint32_t term_mask, double *out);
/* md::system::compute_potential_energy (MD/system.hpp:249-252). */
int gc_compute_energy(gc_ctx *ctx, int32_t replica, uint32_t term_mask, double *out);

/* The same for every replica with ONE cell-list rebuild and one copy: out = [R]
   (the drivers log the energy of every cell at sampling cadence,
   SIM/stage_interphase/simulation_driver_interphase.cpp:24-27). */
int gc_compute_energy_all(gc_ctx *ctx, uint32_t term_mask, double *out);

/* md::neighbor_searcher<open_box>(box, radius).set_points().search()
   (MD/misc/neighbor_searcher.hpp:22-131): unique pairs i < j with
   |xi - xj|^2 <= radius^2 (inclusive), sorted by (i, j); pairs = [cap][2].
   Diagnostic: radius < 0 searches |radius| through the cell list exactly as
   the last step left it (no re-binning), to test the skin invariant. */
int gc_neighbor_pairs(gc_ctx *ctx, int32_t replica, double radius, uint32_t *pairs,
                      size_t capacity, size_t *count);

/* ---- contact map (SIM/stage_interphase/contact_map.cpp:12-85) ---------- */

/* interval = update cadence in steps inside gc_run (0 = never automatic);
   distance scales with core_scale(t) when scale_by_core is set (interphase
   callback: contact distance = contactmap_distance * core_scale);
   table_capacity = slots per replica of the OVERFLOW table (0 = default 64 N): every
   bead first fills a private bucket of partners (128 slots = 1 KiB per bead; environment
   GC_CONTACT_BUCKET overrides), only the excess spills. */
int gc_set_contact_map(gc_ctx *ctx, double distance, int64_t interval,
                       int32_t scale_by_core, size_t table_capacity);
/* contact_map::update on the current positions of every replica. */
int gc_contacts_update(gc_ctx *ctx);
/* contact_map::accumulate: (i, j, count) sorted by (i, j); replica = -1 sums
   over replicas.  ijv = [cap][3]; *count receives the number of triples even
   when it exceeds capacity (then GC_ERR_CAPACITY). ijv may be NULL to query. */
int gc_contacts_fetch(gc_ctx *ctx, int32_t replica, uint32_t *ijv, size_t capacity,
                      size_t *count);
int gc_contacts_clear(gc_ctx *ctx);
/* The same reduction left ON THE DEVICE for an NCCL gather across GPUs
   (the reference merges ensembles offline, src/cool.py:80-114): keys = i << 32 | j
   ascending, counts parallel to them.  The buffers belong to the context and stay
   valid until its next contact query. */
int gc_contacts_reduce_device(gc_ctx *ctx, int32_t replica, uint64_t **keys, uint32_t **counts,
                              size_t *n);
/* Sorts and sums n_in (key, count) device pairs gathered from several ranks; inputs
   may alias a previous result of this context. */
int gc_contacts_merge_device(gc_ctx *ctx, const uint64_t *keys_in, const uint32_t *counts_in,
                             size_t n_in, uint64_t **keys, uint32_t **counts, size_t *n);

/* ---- spatial decomposition of one large replica over several GPUs --------
   (BASELINE configs[4]; not in the reference, which is one process per cell.)
   SLAB + HALO: every rank creates a context with n_replicas = 1 over ALL beads, calls
   gc_decomp_configure(rank, world), gc_set_positions with all N beads and
   gc_run(steps = 0, restart = 1).  Rank k then owns the z layers [z_k, z_k+1) of the cell grid
   (equal-volume slabs of the bounding ellipsoid, re-derived every step from the global
   bounding box) and keeps current copies of the beads within `layers` cell layers of its
   slab.  One step is
       gc_decomp_step_compute   cell list of the slab window, pair forces and the move of the
                                owned beads, this rank's 128-byte header
       all-gather               header -> headers_all over all ranks        (NCCL, gc_stream)
       gc_decomp_step_pack      global reductions + the per-step callback body (next grid);
                                moved beads near a slab face -> send_lo / send_hi
       send/recv                send_hi -> rank + 1's recv_lo, send_lo -> rank - 1's recv_hi
       gc_decomp_step_commit    received halo beads -> positions
   Traffic per step and rank: two halo buffers out, two in (surface, not volume).  The
   trajectory is bit-identical to the single-GPU one for every world size.  Fixed timestep,
   no skin, no contact map.  Positions on one rank are only partly current afterwards: merge
   the ranks with gc_decomp_owned + gc_get_positions.  gc_run(steps = 0, restart = 0)
   refreshes parameters mid-run without touching the grid.  Resuming a checkpoint: all N
   positions, gc_run(steps = 0, restart = 1), gc_set_context, gc_set_step (re-stamps every bead
   as current), gc_run(steps = 0, restart = 0) (re-plans the grid for the loaded context).  world = 1 switches it off.
   Errors (reported by the next synchronising call): GC_ERR_STATE when a bonded partner, or a
   bead after one step, lies outside the halo or a slab is thinner than the halo;
   GC_ERR_CAPACITY when a halo buffer overflows. */
int gc_decomp_configure(gc_ctx *ctx, int32_t rank, int32_t world);
/* Position halo in cell layers (default 2; bonded partners must lie within it) and beads per
   halo buffer (0 keeps the current one; default min(N, 2 ceil(N / world))). */
int gc_decomp_set_halo(gc_ctx *ctx, int32_t layers, int32_t capacity);
/* Upper bound on the beads one rank owns (default, and 0: n_beads); the per-bead kernel's grid
   is sized by it.  A rank that owns more reports GC_ERR_CAPACITY. */
int gc_decomp_set_owned_bound(gc_ctx *ctx, int32_t beads);
/* Device buffers: four halo buffers of halo_elems float4 -- element 0 carries the bead count
   in its first 4 bytes, then (x, y, z, bead) -- and the headers: header = 16 doubles,
   headers_all = [world][16]. */
int gc_decomp_buffers(gc_ctx *ctx, void **send_lo, void **send_hi, void **recv_lo, void **recv_hi,
                      void **header, void **headers_all, size_t *halo_elems);
int gc_decomp_step_compute(gc_ctx *ctx);
int gc_decomp_step_pack(gc_ctx *ctx);
int gc_decomp_step_commit(gc_ctx *ctx);

/* PEER MODE: the halo leaves inside gc_decomp_step_compute, written by the pack kernel
   straight into the neighbours' receive buffers over NVLink (peer-mapped memory), so the
   header all-gather is the ONLY collective of a step and orders those writes before the
   neighbour reads them:
       gc_decomp_step_compute -> all-gather(header) -> gc_decomp_step_pack [-> _commit: no-op]
   The receive buffers are double-buffered by step parity.  The halo is chosen under the grid
   of the step just taken with two layers more than `layers`; the next step checks under the
   new grid that nothing a neighbour needs was left out (GC_ERR_STATE otherwise).
   Set-up, before gc_run(restart = 1): every rank exports its four receive buffers
   (gc_decomp_peer_export: 4 CUDA IPC handles = 4 x 64 bytes: recv_lo[0], recv_lo[1],
   recv_hi[0], recv_hi[1]), the blobs travel by any means, and each rank attaches the blob of
   rank - 1 and of rank + 1 (NULL where there is none).  Ranks that live in one process attach
   raw pointers instead (gc_decomp_recv_buffers of the neighbours -> gc_decomp_peer_attach).
   gc_decomp_peer_detach closes the mappings (before a neighbour resizes or frees its
   buffers) and returns to the send / receive scheme. */
#define GC_DECOMP_IPC_HANDLE_BYTES 64
int gc_decomp_recv_buffers(gc_ctx *ctx, void **out4);
int gc_decomp_peer_export(gc_ctx *ctx, void *blob /* 4 x GC_DECOMP_IPC_HANDLE_BYTES */);
int gc_decomp_peer_attach_ipc(gc_ctx *ctx, const void *lower_blob, const void *upper_blob);
int gc_decomp_peer_attach(gc_ctx *ctx, void *lower_recv_hi0, void *lower_recv_hi1, void *upper_recv_lo0,
                          void *upper_recv_lo1);
int gc_decomp_peer_detach(gc_ctx *ctx);
/* The three step calls only enqueue kernels, so a block of steps -- with the caller's
   collectives between them -- can be captured into a CUDA graph on gc_stream and replayed.
   The library keeps a host-side mirror of the step count and of the position buffer in use:
   a capture advances it without running anything, a replay runs without advancing it.
   gc_decomp_host_advance(ctx, k) corrects the mirror by k steps (k < 0 after a capture of
   -k steps, k > 0 after every replay). */
int gc_decomp_host_advance(gc_ctx *ctx, int64_t steps);
/* mask[n_beads] (host): 1 for the beads this rank owns now; over all ranks every bead is
   owned exactly once. */
int gc_decomp_owned(gc_ctx *ctx, uint8_t *mask);

/* ---- inter-stage transitions on the device ---------------------------------
   The reference runs `transition {interphase|prometaphase|cycle}` as a separate program
   between stage binaries, through the trajectory file.  Here the positions move between two
   contexts of the SAME device without leaving it (a multi-generation lineage stays on the
   GPU).  Ranges are [n_chains][2] = [start, end) bead indices; arithmetic is FP64 as in the
   reference; fraction_bits > 0 reproduces the file hand-off (the source is rounded as the
   stored last frame would be and the result as frame 0 of the next stage,
   SIM/common/simulation_store.cpp:22-32; the reference uses 16), 0 skips the rounding.
   Each call returns when the destination positions are written. */

/* telophase -> relaxation (SIM/stage_transition/transition_interphase.cpp:13-84): every
   coarse chain is fitted with a not-a-knot cubic spline at t = (k + 1/2) / n
   (cxx-spline) and resampled along its interphase chain at t = (i + 1/2) / m; then
   dst[copy_pairs[k][1]] = dst[copy_pairs[k][0]] (nucleolar beads start on their NOR bead).
   Source chains of 3 beads, where the reference's spline is singular, use the parabola. */
int gc_transition_interphase(gc_ctx *src, int32_t src_replica, gc_ctx *dst, int32_t dst_replica,
                             int32_t n_chains, const int32_t *src_ranges, const int32_t *dst_ranges,
                             int32_t n_copies, const int32_t *copy_pairs, int32_t fraction_bits);
/* interphase -> prometaphase (transition_prometaphase.cpp:24-105): bead k of target chain c =
   centroid of source beads [s + cg k, min(s + cg k + cg, s + cg k + len_c)), its sister copy
   = centroid + displacement (-sister_separation along the spindle axis).  sister_ranges may
   be NULL (coarse-graining only). */
int gc_transition_prometaphase(gc_ctx *src, int32_t src_replica, gc_ctx *dst, int32_t dst_replica,
                               int32_t n_chains, const int32_t *src_ranges, const int32_t *target_ranges,
                               const int32_t *sister_ranges, int32_t coarse_graining,
                               const double displacement[3], int32_t fraction_bits);
/* prometaphase -> next anaphase (transition_cycle.cpp:25-76): dst chain = src chain (the
   target chromatid) + displacement (-spindle_axis). */
int gc_transition_cycle(gc_ctx *src, int32_t src_replica, gc_ctx *dst, int32_t dst_replica,
                        int32_t n_chains, const int32_t *src_ranges, const int32_t *dst_ranges,
                        const double displacement[3], int32_t fraction_bits);

/* ---- observables ------------------------------------------------------- */

/* Squared radius of gyration per chain range [start, end) (harness-defined
   observable): out = [n_ranges]. */
int gc_radius_of_gyration(gc_ctx *ctx, int32_t replica, int32_t n_ranges,
                          const int32_t *ranges, double *out_rg2);

/* The same WITHOUT any allocation, for callers that cannot call cudaMalloc (code under CUDA
   graph capture, tight serving loops, harnesses that forbid it).  No counterpart in the
   reference.  The kernel's scratch (the ranges
   and the per-range results) lives in a static device array of the library, so n_ranges is
   limited to GC_RG_NOALLOC_MAX_RANGES (4096; 64 KiB per GPU that loads the library, held until
   exit); more returns GC_ERR_CAPACITY.  Calls no cudaMalloc / cudaFree and does not touch the
   context's topology.  All contexts and calls on one GPU share the array: calls that overlap
   on the GPU (different contexts / streams / host threads) race -- serialise them. */
#define GC_RG_NOALLOC_MAX_RANGES 4096
int gc_radius_of_gyration_noalloc(gc_ctx *ctx, int32_t replica, int32_t n_ranges,
                                  const int32_t *ranges, double *out_rg2);

/* Diagnostic view of the cell list of one replica: out = {dims x,y,z, ncell,
   origin x,y,z, cell_size, rc_valid, max displacement since rebuild,
   need_rebuild, parity, dt_cur, 0, 0, 0}. */
int gc_debug_grid(gc_ctx *ctx, int32_t replica, double out[16]);

/* ---- contact-map observables on the device (gc_obs.cu) ----------------------
   The reference computes these offline in Python from the trajectory file; here the ensemble
   contact map is already on the GPU as sorted (key, count) runs, so the binned matrix and
   everything derived from it are built there.  One gc_obs = one n_bins x n_bins float32 matrix
   on one device.  chrom_ranges = [n_chroms][2] bin ranges [start, end), ascending. */
typedef struct gc_obs gc_obs;
int gc_obs_create(int32_t device, int32_t n_bins, int32_t n_chroms, const int32_t *chrom_ranges,
                  gc_obs **out);
int gc_obs_destroy(gc_obs *obs);
const char *gc_obs_last_error(const gc_obs *obs);
/* bead -> bin (< 0 drops the bead) and, optionally, bead -> chain (for gc_obs_ps_runs_device). */
int gc_obs_set_bead_map(gc_obs *obs, int32_t n_beads, const int32_t *bead_bin, const int32_t *bead_chain);
int gc_obs_clear(gc_obs *obs);
/* Adds contact pixels to the symmetric bin matrix as cooler's matrix() shows the pixels that
   src/cool.py:80-110 ingests (mirrored, the diagonal counted once): DEVICE runs as returned by
   gc_contacts_reduce_device / gc_contacts_merge_device, or host (i, j, count) triples. */
int gc_obs_add_runs_device(gc_obs *obs, const uint64_t *keys, const uint32_t *counts, size_t n);
int gc_obs_add_triples(gc_obs *obs, const uint32_t *ijv, size_t n);
int gc_obs_get_matrix(gc_obs *obs, float *out);          /* [n_bins][n_bins] */
int gc_obs_set_matrix(gc_obs *obs, const float *in);
/* P(s) at bead resolution straight from device runs: hist[s] = sum of counts of intra-chain
   pairs with j - i = s, s = 0..max_s (hist = [max_s + 1], host). */
int gc_obs_ps_runs_device(gc_obs *obs, const uint64_t *keys, const uint32_t *counts, size_t n,
                          int32_t max_s, uint64_t *hist);
/* estimate_expected_contacts (src/pc1.py:156-183): intra[s] = mean finite contact on the s-th
   diagonal over all chromosomes (the cis decay profile), inter = mean finite trans contact.
   capacity = length of intra (gc_obs_max_separation gives the longest chromosome). */
int gc_obs_expected(gc_obs *obs, double *intra, int32_t capacity, double *inter);
int gc_obs_max_separation(gc_obs *obs, int32_t *out);
/* src/pc1.py:36-101 + src/pkg/pc1/math.py:13-37: O/E transform, column selection (coverage >
   0) and standardisation, first right singular vector by power iteration until max |dv| <
   tolerance, PC1 = data @ v scaled so that var(PC1) is the explained variance ratio.
   ev1 / pc1 = [n_bins] (ev1 NaN on dropped bins). */
int gc_obs_pc1(gc_obs *obs, int32_t mask_intra, int32_t use_covariance, double tolerance,
               int32_t max_steps, float *ev1, float *pc1, double *explained_variance_ratio,
               int32_t *steps);
/* compute_nci (src/pkg/nci/nci.py:9-35): out[i] = M[i][i+1] / sqrt(M[i][i] M[i+1][i+1]) inside a
   chromosome, NaN on every chromosome's last bin; out = [n_bins]. */
int gc_obs_nci(gc_obs *obs, float *out);
/* Compartment strength (harness-defined): mean finite contact between bins of type
   a and b (bin_type in [0, n_types), < 0 ignores the bin) over pairs i < j that are trans or
   at least min_separation apart; mean_contact = [n_types][n_types], symmetric. */
int gc_obs_type_enrichment(gc_obs *obs, const int32_t *bin_type, int32_t n_types,
                           int32_t min_separation, double *mean_contact);

/* ---- measurement (bench.py; no reference counterpart) ------------------- */

/* Per-kernel CUDA-event timing on the context's stream.  While enabled every
   launch is bracketed by an event pair (slows the step slightly, so it is used
   in a separate short pass, never inside a timed region).  Enabling resets the
   record.  gc_profile_read(index) returns GC_ERR_INVALID past the last kernel. */
int gc_profile_enable(gc_ctx *ctx, int32_t enabled);
int gc_profile_read(gc_ctx *ctx, int32_t index, char *name, size_t name_capacity,
                    int64_t *launches, double *total_ms);
/* Measured non-tensor FP32 peak of `device` (dependent-free FFMA chains on every
   SM, best of 5), the denominator of the FP32 roofline fraction. */
int gc_measure_fp32_peak(int32_t device, double *tflops);

/* Number of kernels launched by this context since creation (bench claim). */
int gc_launch_count(gc_ctx *ctx, int64_t *out);
/* Block until all queued device work of the context has finished. */
int gc_synchronize(gc_ctx *ctx);
/* CUDA stream handle (cudaStream_t as void*) for event timing by the caller. */
int gc_stream(gc_ctx *ctx, void **out);

#ifdef __cplusplus
}
#endif
#endif /* GC_B200_H */

// gc_types.cuh -- device-visible data structures of the B200 hot path.
//
// Layout in HBM (R replicas x N beads, one context per GPU):
//   pos [2][R][N] float4   positions in ORIGINAL bead order, ping-pong by step parity
//   spos[2][R][N] float4   positions in CELL-SORTED order (w = original index bits)
//   sref   [R][N] float4   sorted positions at the last rebuild (skin > 0 only)
//   scell  [R][N] int32    cell of each sorted slot (query kernels)
//   sab    [R][N] float2   (a, b) factors gathered into sorted order
//   punit  [3][R][N/2+1] float4  the same beads pair-interleaved: planes (x0 x1 y0 y1),
//                           (z0 z1 a0 a1), (b0 b1 - -): what k_pair_cells stages, two beads at a time
//   cell_count / cell_start [R][ncell_max(+1)] int32
//   cell_list [R][N] int4   non-empty cells (the pair kernel's work queue)
//   force  [R][N] float4   only for the adaptive stepper / gc_compute_force
//   pforce [R][N] float4   nonbonded force in sorted order (pair kernel -> k_finish)
//   st     [R]    RepState per-replica scalars (schedule, wall, reductions)
// Bead attributes and bonded topology are shared by all replicas.
#pragma once

#include <cstdint>

#include <cuda_runtime.h>

#define GC_MAX_BOND_TERMS 8
#define GC_MAX_TRIPLE_TERMS 4
#define GC_MAX_SOURCES 4
#define GC_MAX_POLES 4

// fixed-point scale of the wall-reaction accumulators (deterministic sums)
#define GC_FX_SCALE 4294967296.0 /* 2^32 */

// Programmatic dependent launch: every kernel waits for its predecessor to complete and
// flush (a no-op for launches without the attribute), which lets the driver stage the
// launch while the predecessor still runs.  Triggering the successor EARLY
// (griddepcontrol.launch_dependents at entry, GC_PDL_TRIGGER = 1) was measured slower:
// its blocks take SM slots from the draining grid (profiles/r1_notes.md).
#ifndef GC_PDL_TRIGGER
#define GC_PDL_TRIGGER 0
#endif
#define GC_PDL_ENTRY()                                                                \
    do {                                                                              \
        if (GC_PDL_TRIGGER) asm volatile("griddepcontrol.launch_dependents;");        \
        asm volatile("griddepcontrol.wait;" ::: "memory");                            \
    } while (0)

namespace gc {

enum PairMode : int { PAIR_NONE = 0, PAIR_AB = 1, PAIR_PLAIN = 2 };

struct PairParams {
    int mode;           // PairMode
    int term;           // term id (enable bit)
    int scale_by_core;
    int pad;
    double eps_a, sigma_a, eps_b, sigma_b;
};

struct WellParams {     // nucleolar droplet, softwell<6> + cutoff on targets
    int term;           // -1 = absent
    int n_targets;
    const int *targets; // device
    double eps, decay, cutoff;
};

struct BondTerm {
    int kind, mix_ab, scale, term;
    double K_a, K_b, l_a, l_b;
};

struct TripleTerm {
    int term, pad;
    double energy;
};

struct SourceTerm {
    int kind, term;
    double point[3];
    double p0, p1;
};

struct WallParams {
    int term;           // -1 = absent
    int scale_by_core;
    int ode_enabled;
    int pad;
    double eps_a, sigma_a, eps_b, sigma_b, wall_a, wall_b, outward_spring;
    double ode_mobility;
    double ode_spring[3];
};

struct Schedule {
    int enabled, pad;
    double core_init, core_tau, bond_init, bond_tau;
};

// Contact map (SIM/stage_interphase/contact_map.hpp:11-68).  Every bead i owns a
// small open-addressing bucket of (j << 32 | count) words for its partners j > i,
// so all inserts of a home bead land in two cache lines; pairs that do not fit
// spill into a conventional (i << 32 | j) -> count hash table.
struct ContactParams {
    double base_distance;
    long long interval;
    int scale_by_core;
    int enabled;
    unsigned long long *bucket;   // [R][N][bucket_slots]
    int bucket_slots;             // power of two
    int pad;
    unsigned long long *keys;     // spill table [R][capacity]
    unsigned int *vals;           // [R][capacity]
    unsigned long long capacity;  // power of two, per replica
    unsigned int *used;           // [R] distinct pairs stored (buckets + spill)
};

// Per-bead bonded adjacency (gather form: every bead sums its own terms, no
// atomics, deterministic).
struct BondEntry { int partner; int bterm; };          // bterm indexes Params::bonds
struct TripleEntry { int i, j, k; int role_term; };    // role (0,1,2) | tterm << 2
struct TetherEntry { int pole; float spring; float length; int term; };

// FP32 constants of one step, derived from the FP64 context by the last block
// of the previous step (refresh_consts) so that no per-bead FP64 work is needed.
struct StepConst {
    double qd[3];                                 // 1 / semiaxis^2 (wall geometry is FP64)
    float kA2, kB2, cA, cB, eA, eB, rc2;          // pair soft cores (PairCoef)
    float w_kA2, w_kB2, w_cA, w_cB, w_eA, w_eB;   // wall inward soft cores (half diameters)
    float q[3];                                   // 1 / semiaxis^2
    float wall_a, wall_b, wall_K;
    float bK_a[GC_MAX_BOND_TERMS], bK_b[GC_MAX_BOND_TERMS];
    float bl_a[GC_MAX_BOND_TERMS], bl_b[GC_MAX_BOND_TERMS];
    float tE[GC_MAX_TRIPLE_TERMS];
    float src_pt[GC_MAX_SOURCES][3], src_p0[GC_MAX_SOURCES], src_p1[GC_MAX_SOURCES];
    float pole[GC_MAX_POLES][3];
    float well_eps, well_k2, well_cut2;
    float dt_cur;                                 // dt of the step in flight
    float pad;
};

// Per-replica mutable scalars.  Written only by the last block of a step (or
// by the host between runs); read by every block at kernel start.
constexpr int GC_MAX_RANKS = 16;          // ranks of a decomposed run

struct RepState {
    // --- context (double, as the reference's interphase_context) ---
    double time;
    double core_scale, bond_scale;
    double semiaxes[3];
    double contact_distance;       // used by the next explicit update
    double contact_distance_prev;  // value in force during the last callback
    double axial[3], reaction;     // wall statistics of the last force pass
    double dt_last;                // dt used by the last step
    double dt_prev;                // dt of the step before (Wiener regeneration)
    long long step;                // steps done in this logical run
    long long rebuilds;
    unsigned long long seed;
    // --- per-step reductions (reset by the last block) ---
    long long axial_fx[3], reaction_fx;
    unsigned int maxdisp2_bits;    // max |x - x_ref|^2 since rebuild (float bits)
    unsigned int dtmin_bits;       // adaptive dt candidate (float bits)
    int bbox_min[3], bbox_max[3];  // order-preserving int encoding of floats
    int scan_ticket;               // block order of the chained cell scan
    int n_used_cells;              // heavy entries of Params::cell_list (from the front)
    int cell_cursor;               // work-queue head of the pair kernel, reset every step
    // --- cell grid ---
    int need_rebuild;
    int parity;                    // which ping-pong buffer holds current positions
    int dims[3];
    int ncell;
    float origin[3];
    float cell_size, inv_cell;
    float rc_valid;                // interaction radius the current cells guarantee
    int error;                     // sticky: 1 = NaN/Inf seen
    int n_light_cells;             // lightly populated cells, queued from the BACK of cell_list
    int contact_cursor;            // work-queue head of k_contact_cells
    int epoch;                     // logical runs since the seeds were set (Philox counter word 3)
    // z-window of a decomposed run: dims[2] counts the LOCAL layers [zoff, zoff + dims[2]) of the
    // global grid (gdimz layers); the local layers [zown_lo, zown_hi) are owned by this rank
    int zoff, gdimz, zown_lo, zown_hi;
    // slab boundaries of all ranks: dfrac[k] = volume fraction of the bounding ellipsoid below
    // rank k's slab (k / world at restart, then nudged every step towards equal bead counts,
    // identically on every rank: decomp_reduce), dz[k] = its first layer in the current grid
    double dfrac[GC_MAX_RANKS + 1];
    int dz[GC_MAX_RANKS + 1];
    // the beads current on this rank, as a work list for the next cell-list build: the sorted
    // slots [dprev_lo, dprev_hi) it moved in the last step (bead index in spos[].w) plus the
    // two received halo buffers; dfull = 1: all N beads (after gc_run(restart = 1))
    int dfull, dprev_lo, dprev_hi;
    // peer mode packs under the grid of the step just taken (GC_DECOMP_PEER_MARGIN layers more, so that
    // the header all-gather is the only collective of a step); the next cell-list build
    // checks, under the new grid, that everything a neighbour needs was in that set
    float dold_oz, dold_inv;
    int dold_gdimz, dold_z0, dold_z1;
    StepConst sc;
};

struct Params {
    int n, n_rep;
    int ncell_max;
    int cells_per_cut;             // M: cell edge = r_list / M
    float skin;                    // fraction of the cutoff
    float kT;
    float dt;                      // configured dt (dt_max in adaptive mode)
    double dt_cfg;                 // the same in FP64 (time, wall ODE)
    float spacestep;
    unsigned int enabled;          // term enable bits

    // shared attributes
    const float *mob;              // [N]
    const float2 *ab;              // [N]
    const int *flags;              // [N] bit0 triples, bit1 tethers, bit2 droplet target
    const int *bond_off;           // [N+1]
    const BondEntry *bond_ent;
    const int *triple_off;         // [N+1]
    const TripleEntry *triple_ent;
    const int *tether_off;         // [N+1]
    const TetherEntry *tether_ent;

    // per-replica state
    float4 *pos[2];
    float4 *spos[2];
    float4 *sref;
    int *scell;
    float2 *sab;
    // pair-interleaved copy of the sorted beads for k_pair_cells (gc_cells.cuh): unit u =
    // sorted slots 2u, 2u + 1 in three planes of [R][n_units] float4,
    // (x0 x1 y0 y1), (z0 z1 a0 a1), (b0 b1 - -); plane k starts at punit + k * unit_plane
    float4 *punit;
    size_t unit_plane;
    int n_units;                   // (n + 1) / 2 + 1 + max_layers (layer-relative pairing: unit_slot)
    int max_layers;                // bound on the z layers of the cell grid (plan_grid coarsens beyond)
    int *cell_count, *cell_start, *members, *cell_of;
    unsigned long long *scan_status;   // [R][scan_blocks] look-back words of k_scan
    int scan_blocks;               // ceil(ncell_max / SCAN_TILE)
    int4 *cell_list;               // [R][N] non-empty cells {cell, first slot, beads, 0}
    float4 *force;                 // total force, ORIGINAL order (adaptive stepper, gc_compute_force)
    float4 *pforce;                // nonbonded force, SORTED order (k_pair_cells -> k_finish)
    RepState *st;
    // spatial decomposition of ONE large replica over several GPUs (gc_decomp.cuh): z-slabs of
    // the cell grid with a halo exchange between neighbouring ranks
    int decomp;                    // 0: off
    int drank, dworld;             // this rank, ranks
    int dhalo;                     // position halo in cell layers (>= 1; bonded partners must lie inside)
    int dcap;                      // beads per halo buffer
    int down_bound;                // bound on the beads a rank owns (sizes k_finish's grid)
    int *dstamp;                   // [N] step at which pos[] of a bead was last written on this rank
    float4 *dsend_lo, *dsend_hi;   // [dcap + 1] (x, y, z, bead) for rank - 1 / rank + 1; [0].x = count bits
    // the same, received from rank - 1 / rank + 1.  Peer mode (dpeer = 1) double-buffers them by
    // step parity and the neighbours write them DIRECTLY over NVLink: dpeer_lo[b] is rank - 1's
    // drecv_hi[b], dpeer_hi[b] is rank + 1's drecv_lo[b] (peer-mapped)
    const float4 *drecv_lo[2], *drecv_hi[2];
    float4 *dpeer_lo[2], *dpeer_hi[2];
    int dpeer;
    double *dhdr;                  // [16] this rank's share of the per-step reductions
    const double *dhdr_all;        // [dworld][16] every rank's header, gathered
    const float *noise;            // injected normals [R][steps][N][3] or null
    long long noise_steps;         // steps covered by `noise`
    long long noise_step0;         // st->step value that maps to noise row 0

    // terms
    PairParams pair;
    WellParams well;
    WallParams wall;
    Schedule sched;
    ContactParams contact;
    int n_bond_terms, n_triple_terms, n_sources, n_poles;
    BondTerm bonds[GC_MAX_BOND_TERMS];
    TripleTerm triples[GC_MAX_TRIPLE_TERMS];
    SourceTerm sources[GC_MAX_SOURCES];
    double poles[GC_MAX_POLES][3];
};

// first float of sorted slot `slot` of replica r in plane `plane` of Params::punit (the
// second value of the same bead sits two floats further)
__device__ __forceinline__ float *unit_half(const Params &P, int r, int slot, int plane)
{
    return reinterpret_cast<float *>(P.punit + (size_t) plane * P.unit_plane + (size_t) r * P.n_units +
                                     (size_t) (slot >> 1)) + (slot & 1);
}

// Peer mode of a decomposed run ships the halo under the grid of the step just taken.  Between
// two plans a slab boundary moves by at most one layer (the grid gains or loses one) and a
// bead's layer by at most one (its own motion plus the shift of the origin are far below a
// cell edge), so two extra layers cover what the new grid can ask for; k_count verifies.
constexpr int GC_DECOMP_PEER_MARGIN = 2;

// Position of sorted slot s (a bead of z layer L, whose first slot is layer_start) in the
// pair-interleaved copy.  Beads are paired RELATIVE TO THEIR LAYER -- the first bead of a
// layer always opens a unit -- so the units the pair kernel stages for a home cell, and with
// them the order of every force sum, depend on the three layers around the cell only: a
// decomposed run, which bins a window of layers, reproduces the whole-grid sums bit for bit.
// Layer L occupies units from (layer_start >> 1) + L, which leaves 1-3 unused halves between
// consecutive layers; the first of them is kept a dummy (gather_slot).
__device__ __forceinline__ int unit_slot(int s, int L, int layer_start)
{
    return s + 2 * L - (layer_start & 1);
}

static_assert(sizeof(Params) <= 4000, "Params is passed by value as a kernel argument");

} // namespace gc

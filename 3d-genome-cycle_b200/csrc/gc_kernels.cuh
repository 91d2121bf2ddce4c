// gc_kernels.cuh -- sm_100a kernels of the hot path (part 1: helpers, cell
// list construction).  One thread block grid is (ceil(N / BS), R): blockIdx.y
// selects the replica.
//
// Neighbor search replaces md::neighbor_searcher<open_box> + md::neighbor_list
// (MD/misc/neighbor_searcher.hpp:22-131, MD/forcefield/detail/neighbor_list.hpp:
// 62-240): instead of a hashed grid + stored Verlet pair list on the host, beads
// are counting-sorted into a dense x-fastest cell grid in HBM and every step
// walks the (2M+1)^2 contiguous x-rows of neighbouring cells directly (no pair
// list is ever stored).  The pair SET that results is identical (tests).
#pragma once

#include "gc_math.cuh"
#include "gc_types.cuh"

namespace gc {

constexpr int BS = 128;          // threads per block of the per-bead kernels
constexpr int SCAN_BS = 1024;
#ifndef GC_HEAVY_CELL
#define GC_HEAVY_CELL 12
#endif

// ---------------------------------------------------------------------------
// small helpers
// ---------------------------------------------------------------------------

__device__ __forceinline__ double pair_cutoff(const Params &P, const RepState &st)
{
    if (P.pair.mode == PAIR_NONE) return 0.0;
    double s = P.pair.scale_by_core ? st.core_scale : 1.0;
    if (P.pair.mode == PAIR_PLAIN) return P.pair.sigma_a * s;
    return fmax(P.pair.sigma_a, P.pair.sigma_b) * s;
}

// Radius the cell list has to resolve for this replica right now.
__device__ __forceinline__ double interaction_radius(const Params &P, const RepState &st)
{
    double r = pair_cutoff(P, st);
    if (P.contact.enabled) {
        r = fmax(r, fmax(st.contact_distance, st.contact_distance_prev));
    }
    return r;
}

__device__ __forceinline__ PairCoef make_pair_coef(double eps_a, double sigma_a, double eps_b,
                                                   double sigma_b, double s, int mode)
{
    PairCoef c;
    double dA = sigma_a * s, dB = sigma_b * s;
    double kA2 = 1.0 / (dA * dA);
    double kB2 = mode == PAIR_PLAIN ? 0.0 : 1.0 / (dB * dB);
    c.kA2 = (float) kA2;
    c.kB2 = (float) kB2;
    c.cA = (float) (6.0 * eps_a * kA2);
    c.cB = (float) (24.0 * eps_b * kB2);
    c.eA = (float) eps_a;
    c.eB = mode == PAIR_PLAIN ? 0.0f : (float) eps_b;
    double rc = mode == PAIR_PLAIN ? dA : fmax(dA, dB);
    c.rc2 = (float) (rc * rc);
    return c;
}

// The grid carries a halo of `halo` (= cells_per_cut) always-empty cells on every
// side, so neighbourhood walks need no bounds checks.
__device__ __forceinline__ int cell_coord(float x, float origin, float inv, int dim, int halo)
{
    int c = (int) floorf((x - origin) * inv);
    return min(max(c, halo), dim - 1 - halo);
}

// z layer of a position in the GLOBAL grid (identical arithmetic on every rank of a
// decomposed run)
__device__ __forceinline__ int cell_layer(const RepState &st, float z, int halo)
{
    return cell_coord(z, st.origin[2], st.inv_cell, st.gdimz, halo);
}

// Cell of a position, or -1 when its layer lies outside this context's z-window (a decomposed
// run bins only its slab plus one layer on either side; otherwise the window is the grid).
__device__ __forceinline__ int cell_index(const RepState &st, float4 p, int halo)
{
    int cx = cell_coord(p.x, st.origin[0], st.inv_cell, st.dims[0], halo);
    int cy = cell_coord(p.y, st.origin[1], st.inv_cell, st.dims[1], halo);
    int cz = cell_layer(st, p.z, halo) - st.zoff;
    if (cz < 0 || cz >= st.dims[2]) return -1;
    return cx + st.dims[0] * (cy + st.dims[1] * cz);
}

// First layer of a slab that starts at volume fraction f of the ellipsoid the bounding box
// encloses, in a grid of gdimz layers whose outer `halo` layers on either side are the
// always-empty margin.  The volume fraction of an ellipsoid below height u in [-1, 1] is
// (2 + 3u - u^3) / 4, inverted in closed form -- so equal fractions are slabs of equal volume,
// and the ranks of a roughly uniformly filled nucleus start out with similar bead counts.
__device__ __host__ inline int slab_layer(double f, int gdimz, int halo)
{
    const int nint = gdimz - 2 * halo;
    if (f <= 0.0) return halo;
    if (f >= 1.0) return halo + nint;
    const double u = 2.0 * cos((acos(1.0 - 2.0 * f) + 4.0 * 3.14159265358979323846) / 3.0);
    const int z = halo + (int) floor(0.5 * (u + 1.0) * (double) nint + 0.5);
    return z < halo ? halo : (z > halo + nint ? halo + nint : z);
}

// first layer of rank k's slab in the current grid (k outside [0, world] clamps)
__device__ __forceinline__ int slab_start(const Params &P, const RepState &st, int k)
{
    return st.dz[k < 0 ? 0 : (k > P.dworld ? P.dworld : k)];
}

// z-window of this rank: its slab plus the pair halo (cells_per_cut layers) on either side
__device__ inline void decomp_window(const Params &P, RepState &st)
{
    const int h = P.cells_per_cut, W = P.dworld;
    // a slab must be at least as thick as the position halo (plus peer mode's margin): its far
    // neighbour would otherwise need beads from beyond the near one
    const int tmin = P.dhalo + (P.dpeer ? GC_DECOMP_PEER_MARGIN : 0);
    int z[GC_MAX_RANKS + 1];
    for (int k = 0; k <= W; k++) {
        z[k] = slab_layer(k == 0 ? 0.0 : (k == W ? 1.0 : st.dfrac[k]), st.gdimz, h);
        // an inner boundary moves by at most one layer between two plans (the halo logic,
        // peer mode's margin in particular, relies on it); it catches up over the next steps
        if (!st.dfull && k > 0 && k < W) z[k] = min(max(z[k], st.dz[k] - 1), st.dz[k] + 1);
    }
    // (slab k - 1 lies between z[k - 1] and z[k]; the two end slabs have no far neighbour and
    // may be as thin as one layer)
    for (int k = 1; k < W; k++) z[k] = max(z[k], z[k - 1] + (k - 1 >= 1 ? tmin : 1));
    for (int k = W - 1; k > 0; k--) z[k] = min(z[k], z[k + 1] - (k <= W - 2 ? tmin : 1));
    for (int k = 0; k <= W; k++) {
        st.dz[k] = z[k];
        const bool inner = k - 1 >= 1 && k - 1 <= W - 2;
        if (k > 0 && z[k] - z[k - 1] < (inner ? tmin : 1)) st.error = 33;      // too many ranks for this grid
    }
    const int z0 = st.dz[P.drank], z1 = st.dz[P.drank + 1];
    st.zoff = z0 - h;
    st.dims[2] = (z1 - z0) + 2 * h;
    st.zown_lo = h;
    st.zown_hi = h + (z1 - z0);
}

// Chooses origin / cell size / dims from the bounding box for list radius
// `rlist`, plus a halo of cells_per_cut empty cells on every side.  Beads
// outside the box are folded into the outermost interior cells by clamping,
// which never separates two beads further than their true cell distance, so no
// pair within rlist can be missed.
__device__ inline void plan_grid(const Params &P, RepState &st, double rlist)
{
    float lo[3], hi[3];
    for (int k = 0; k < 3; k++) {
        lo[k] = ordered_to_float(st.bbox_min[k]);
        hi[k] = ordered_to_float(st.bbox_max[k]);
        if (!(hi[k] >= lo[k]) || isinf(lo[k]) || isinf(hi[k])) {   // NaN/Inf or empty
            if (P.n > 0) st.error = 1;
            lo[k] = 0.0f;
            hi[k] = 0.0f;
        }
    }
    if (!(rlist > 0.0)) rlist = 1.0;
    // a hair above rlist / M: the FP32 image of the edge (inv_cell) may round the other
    // way, and M cells must span at least rlist for every pair at exactly that distance
    double cs = rlist * (1.0 + 1.0e-6) / P.cells_per_cut;
    int dims[3];
    for (;;) {
        double tot = 1.0;
        for (int k = 0; k < 3; k++) {
            double e = (double) hi[k] - (double) lo[k];
            double d = floor(e / cs) + 1.0 + 2.0 * P.cells_per_cut;
            if (d > 1.0e6) d = 1.0e6;
            dims[k] = (int) d;
            tot *= d;
        }
        if (tot <= (double) P.ncell_max && dims[2] <= P.max_layers) break;
        cs *= 1.2599210498948732;   // 2^(1/3): halves the cell count per round
    }
    for (int k = 0; k < 3; k++) {
        st.dims[k] = dims[k];
        st.origin[k] = (float) ((double) lo[k] - cs * P.cells_per_cut);
    }
    st.gdimz = dims[2];
    st.zoff = 0;
    st.zown_lo = 0;
    st.zown_hi = dims[2];
    if (P.decomp) decomp_window(P, st);          // dims[2] becomes this rank's layer count
    st.ncell = st.dims[0] * st.dims[1] * st.dims[2];
    st.cell_size = (float) cs;
    st.inv_cell = (float) (1.0 / cs);
    // A pair is guaranteed to be seen while (distance now) + 2 max displacement
    // <= M * cell edge.  The float cell edge may round below cs, so keep a
    // hair of margin.
    st.rc_valid = (float) (cs * P.cells_per_cut * (1.0 - 1.0e-6));
}

// ---------------------------------------------------------------------------
// bounding box + grid planning (run start, explicit force / pair queries)
// ---------------------------------------------------------------------------

__global__ void k_reset_reductions(const __grid_constant__ Params P)
{
    GC_PDL_ENTRY();
    int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= P.n_rep) return;
    RepState &st = P.st[r];
    for (int k = 0; k < 3; k++) {
        st.bbox_min[k] = 0x7fffffff;
        st.bbox_max[k] = (int) 0x80000000;
        st.axial_fx[k] = 0;
    }
    st.reaction_fx = 0;
    st.maxdisp2_bits = 0u;
    st.dtmin_bits = 0x7f800000u;
    st.cell_cursor = 0;
    st.contact_cursor = 0;
}

__global__ void k_bbox(const __grid_constant__ Params P)
{
    GC_PDL_ENTRY();
    int r = blockIdx.y;
    RepState &st = P.st[r];
    const float4 *pos = P.pos[st.parity] + (size_t) r * P.n;
    float lo[3] = {INFINITY, INFINITY, INFINITY};
    float hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    bool bad = false;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < P.n; i += gridDim.x * blockDim.x) {
        float4 p = pos[i];
        float v[3] = {p.x, p.y, p.z};
        for (int k = 0; k < 3; k++) {
            lo[k] = fminf(lo[k], v[k]);
            hi[k] = fmaxf(hi[k], v[k]);
            bad |= !(fabsf(v[k]) < INFINITY);
        }
    }
    for (int k = 0; k < 3; k++) {
        for (int o = 16; o > 0; o >>= 1) {
            lo[k] = fminf(lo[k], __shfl_xor_sync(0xffffffffu, lo[k], o));
            hi[k] = fmaxf(hi[k], __shfl_xor_sync(0xffffffffu, hi[k], o));
        }
    }
    bad = __any_sync(0xffffffffu, bad);
    if ((threadIdx.x & 31) == 0) {
        for (int k = 0; k < 3; k++) {
            atomicMin(&st.bbox_min[k], float_to_ordered(lo[k]));
            atomicMax(&st.bbox_max[k], float_to_ordered(hi[k]));
        }
        if (bad) st.error = 1;
    }
}

// ---------------------------------------------------------------------------
// cell list construction: count -> scan -> fill -> gather (deterministic:
// beads inside a cell end up ordered by original index)
// ---------------------------------------------------------------------------

// Decomposed runs: the beads current on this rank, enumerated without touching the other
// N - N / ranks beads (RepState::dfull / dprev_*).
__device__ __forceinline__ int decomp_recv_index(const Params &P, const RepState &st)
{
    return P.dpeer ? (int) ((st.step + 1) & 1) : 0;      // written during step st.step - 1
}

__device__ __forceinline__ int decomp_current_count(const Params &P, const RepState &st)
{
    if (st.dfull) return P.n;
    const int rb = decomp_recv_index(P, st);
    int w = st.dprev_hi - st.dprev_lo;
    if (P.drank > 0) w += min(reinterpret_cast<const int *>(P.drecv_lo[rb])[0], P.dcap);
    if (P.drank + 1 < P.dworld) w += min(reinterpret_cast<const int *>(P.drecv_hi[rb])[0], P.dcap);
    return w;
}

// (x, y, z, bead index bits) of work item w; *received = it came in a halo buffer
__device__ __forceinline__ float4 decomp_current_bead(const Params &P, const RepState &st, int w, bool *received)
{
    *received = false;
    if (st.dfull) {
        float4 p = P.pos[st.parity][w];
        p.w = __int_as_float(w);
        return p;
    }
    const int na = st.dprev_hi - st.dprev_lo;
    // k_finish left the moved beads in the slot order of the step just taken, in the sorted
    // buffer of the parity that is current now (this step's k_gather overwrites it later)
    if (w < na) return P.spos[st.parity][st.dprev_lo + w];
    *received = true;
    w -= na;
    const int rb = decomp_recv_index(P, st);
    const int nlo = P.drank > 0 ? min(reinterpret_cast<const int *>(P.drecv_lo[rb])[0], P.dcap) : 0;
    if (w < nlo) return P.drecv_lo[rb][1 + w];
    return P.drecv_hi[rb][1 + w - nlo];
}

__global__ void __launch_bounds__(256) k_count(const __grid_constant__ Params P)
{
    GC_PDL_ENTRY();
    int r = blockIdx.y;
    const RepState &st = P.st[r];
    if (!st.need_rebuild) return;
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    // reset the scan's look-back words, its ticket and the used-cell list
    if (i < P.scan_blocks) P.scan_status[(size_t) r * P.scan_blocks + i] = 0ull;
    if (i == 0) { P.st[r].scan_ticket = 0; P.st[r].n_used_cells = 0; P.st[r].n_light_cells = 0; }
    if (P.decomp) {
        // a decomposed run bins the beads whose position is current on this rank (grid-stride
        // over that list; the grid need not cover N)
        const int stride = gridDim.x * blockDim.x;
        for (int k = i + stride; k < P.scan_blocks; k += stride) P.scan_status[k] = 0ull;
        const int nw = decomp_current_count(P, st);
        const int h = P.cells_per_cut;
        __shared__ int zs[4];       // slab starts of ranks drank - 1 .. drank + 2
        if (threadIdx.x < 4) zs[threadIdx.x] = slab_start(P, st, P.drank - 1 + threadIdx.x);
        __syncthreads();
        for (int w = i; w < nw; w += stride) {
            bool received;
            const float4 p = decomp_current_bead(P, st, w, &received);
            const int b = __float_as_int(p.w);
            if (received) {
                // a halo copy: position by bead index (bonded partners are looked up there),
                // stamped current
                P.pos[st.parity][b] = make_float4(p.x, p.y, p.z, 0.0f);
                P.dstamp[b] = (int) st.step;
            } else if (P.dpeer && !st.dfull) {
                // moved here in the last step and shipped under THAT step's grid: everything a
                // neighbour needs under the new grid must have been in the shipped set
                const int ln = cell_layer(st, p.z, h);
                const int lo_ = cell_coord(p.z, st.dold_oz, st.dold_inv, st.dold_gdimz, h);
                const int reach = P.dhalo + GC_DECOMP_PEER_MARGIN;
                const bool sent_lo = lo_ < st.dold_z0 + reach, sent_hi = lo_ >= st.dold_z1 - reach;
                const bool need_lo = P.drank > 0 && ln < zs[1] + P.dhalo;
                const bool need_hi = P.drank + 1 < P.dworld && ln >= zs[2] - P.dhalo;
                const bool lost = (P.drank > 1 && ln < zs[0] + P.dhalo) ||
                                  (P.drank + 2 < P.dworld && ln >= zs[3] - P.dhalo);
                if ((need_lo && !sent_lo) || (need_hi && !sent_hi)) atomicExch(&P.st[r].error, 30);
                if (lost) atomicExch(&P.st[r].error, 32);
            }
            const int c = cell_index(st, p, h);
            P.cell_of[b] = c;
            if (c >= 0) atomicAdd(&P.cell_count[c], 1);
        }
        return;
    }
    if (i >= P.n) return;
    float4 p = P.pos[st.parity][(size_t) r * P.n + i];
    int c = cell_index(st, p, P.cells_per_cut);
    P.cell_of[(size_t) r * P.n + i] = c;
    atomicAdd(&P.cell_count[(size_t) r * P.ncell_max + c], 1);
}

// Exclusive scan of the cell counts: chained multi-block scan with decoupled
// look-back.  A block takes a ticket (so that every predecessor it may wait for
// has already been scheduled), scans SCAN_TILE cells, publishes its aggregate,
// looks back over its predecessors' (aggregate | inclusive prefix) words, then
// writes the cell starts.  Status words and the ticket are reset by k_count.
constexpr int SCAN_ITEMS = 4;
constexpr int SCAN_TILE = SCAN_BS * SCAN_ITEMS;
constexpr unsigned long long SCAN_AGG = 1ull << 62, SCAN_PREFIX = 2ull << 62;
constexpr unsigned long long SCAN_VALUE = (1ull << 62) - 1;

__global__ void __launch_bounds__(SCAN_BS) k_scan(const __grid_constant__ Params P)
{
    GC_PDL_ENTRY();
    const int r = blockIdx.y;
    RepState &st = P.st[r];
    if (!st.need_rebuild) return;
    // the grid is sized for ncell_max: only as many blocks as the grid has tiles take a ticket
    // (the others would only queue up on the ticket word)
    if ((int) blockIdx.x * SCAN_TILE >= st.ncell) return;
    __shared__ int warp_sums[SCAN_BS / 32];
    __shared__ int sh_bid, sh_excl;
    if (threadIdx.x == 0) sh_bid = atomicAdd(&st.scan_ticket, 1);
    __syncthreads();
    const int bid = sh_bid;
    const int ncell = st.ncell;
    const int base = bid * SCAN_TILE;
    if (base >= ncell) return;                               // block-uniform

    int *count = P.cell_count + (size_t) r * P.ncell_max;   // 16-byte aligned: ncell_max % 4 == 0
    int *start = P.cell_start + (size_t) r * (P.ncell_max + 1);
    unsigned long long *status = P.scan_status + (size_t) r * P.scan_blocks;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;

    const int c = base + threadIdx.x * SCAN_ITEMS;
    int4 v = make_int4(0, 0, 0, 0);
    if (c + 3 < ncell) {
        v = *reinterpret_cast<const int4 *>(count + c);
        *reinterpret_cast<int4 *>(count + c) = make_int4(0, 0, 0, 0);   // k_fill's cursors
    } else {
        if (c < ncell) { v.x = count[c]; count[c] = 0; }
        if (c + 1 < ncell) { v.y = count[c + 1]; count[c + 1] = 0; }
        if (c + 2 < ncell) { v.z = count[c + 2]; count[c + 2] = 0; }
    }
    const int tsum = v.x + v.y + v.z + v.w;
    int x = tsum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        int y = __shfl_up_sync(0xffffffffu, x, o);
        if (lane >= o) x += y;
    }
    if (lane == 31) warp_sums[wid] = x;
    __syncthreads();
    if (wid == 0) {
        int w = warp_sums[lane];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int y = __shfl_up_sync(0xffffffffu, w, o);
            if (lane >= o) w += y;
        }
        warp_sums[lane] = w;                                 // inclusive over warps
        const int aggregate = __shfl_sync(0xffffffffu, w, 31);

        // ---- publish, look back (warp 0) ---------------------------------------
        int excl = 0;
        if (bid == 0) {
            if (lane == 0) atomicExch(&status[0], SCAN_PREFIX | (unsigned long long) aggregate);
        } else {
            if (lane == 0) atomicExch(&status[bid], SCAN_AGG | (unsigned long long) aggregate);
            int pos = bid - 1;
            for (;;) {
                const int p = pos - lane;
                unsigned long long sw = SCAN_PREFIX;         // virtual predecessor of block 0
                if (p >= 0) sw = *(volatile unsigned long long *) &status[p];
                const unsigned ready = __ballot_sync(0xffffffffu, (sw >> 62) != 0);
                const unsigned pref = __ballot_sync(0xffffffffu, (sw >> 62) == 2);
                // lanes closer than the first inclusive prefix must all be published
                const int first = pref ? __ffs(pref) - 1 : 32;
                const unsigned need = first >= 31 ? 0xffffffffu : ((2u << first) - 1u);
                if ((ready & need) != need) continue;        // spin: predecessors are running
                int val = (lane <= first) ? (int) (sw & SCAN_VALUE) : 0;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) val += __shfl_xor_sync(0xffffffffu, val, o);
                excl += val;
                if (pref) break;
                pos -= 32;
            }
            if (lane == 0) {
                atomicExch(&status[bid], SCAN_PREFIX | (unsigned long long) (excl + aggregate));
            }
        }
        if (lane == 0) sh_excl = excl;
    }
    __syncthreads();
    const int warp_off = wid == 0 ? 0 : warp_sums[wid - 1];
    int e = sh_excl + warp_off + x - tsum;   // exclusive prefix of this thread's first cell
    if (c < ncell) start[c] = e;
    e += v.x;
    if (c + 1 < ncell) start[c + 1] = e;
    e += v.y;
    if (c + 2 < ncell) start[c + 2] = e;
    e += v.z;
    if (c + 3 < ncell) start[c + 3] = e;
    e += v.w;
    if (c <= ncell - 1 && ncell - 1 <= c + 3) start[ncell] = e;   // the thread owning the last cell
}

__global__ void __launch_bounds__(256) k_fill(const __grid_constant__ Params P)
{
    GC_PDL_ENTRY();
    int r = blockIdx.y;
    const RepState &st = P.st[r];
    if (!st.need_rebuild) return;
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (P.decomp) {
        const int nw = decomp_current_count(P, st);
        for (int w = i; w < nw; w += gridDim.x * blockDim.x) {
            bool received;
            const int b = __float_as_int(decomp_current_bead(P, st, w, &received).w);
            const int c = P.cell_of[b];
            if (c < 0) continue;
            P.members[P.cell_start[c] + atomicAdd(&P.cell_count[c], 1)] = b;
        }
        return;
    }
    if (i >= P.n) return;
    int c = P.cell_of[(size_t) r * P.n + i];
    int s = P.cell_start[(size_t) r * (P.ncell_max + 1) + c] +
            atomicAdd(&P.cell_count[(size_t) r * P.ncell_max + c], 1);
    P.members[(size_t) r * P.n + s] = i;
}

// Sorted slot of member entry `s`: rank inside the cell = ascending bead id, so the sorted
// order (and with it every force sum) does not depend on the order of k_fill's atomics.
__device__ __forceinline__ void gather_slot(const Params &P, RepState &st, int r, int s)
{
    size_t ro = (size_t) r * P.n;
    const int *members = P.members + ro;
    const int total = P.cell_start[(size_t) r * (P.ncell_max + 1) + st.ncell];   // beads binned
    if (s >= total) return;
    int idx = members[s];
    int c = P.cell_of[ro + idx];
    const int *start = P.cell_start + (size_t) r * (P.ncell_max + 1);
    int base = start[c], end = start[c + 1];
    int rank = 0;
    for (int k = base; k < end; k++) rank += members[k] < idx;
    if (rank == 0) {
        P.cell_count[(size_t) r * P.ncell_max + c] = 0;   // cursor back to 0
        // work queue of the pair kernel; its order is arbitrary but no result depends on it
        // (a decomposed run only evaluates the cells of the layers this rank owns)
        const int layer = c / (st.dims[0] * st.dims[1]);
        if (!P.decomp || (layer >= st.zown_lo && layer < st.zown_hi)) {
            // well populated cells are queued from the front, the rest from the back:
            // the pair kernel then starts on the heavy ones and fills its tail with
            // light ones (longest-processing-time-first)
            const int4 item = make_int4(c, base, end - base, layer);
            if (end - base >= GC_HEAVY_CELL) {
                P.cell_list[ro + atomicAdd(&st.n_used_cells, 1)] = item;
            } else {
                P.cell_list[ro + P.n - 1 - atomicAdd(&st.n_light_cells, 1)] = item;
            }
        }
    }
    int dst = base + rank;
    float4 p = P.pos[st.parity][ro + idx];
    p.w = __int_as_float(idx);
    P.spos[st.parity][ro + dst] = p;
    if (P.skin > 0.0f) P.sref[ro + dst] = p;
    P.scell[ro + dst] = c;
    const float2 ab = P.ab[idx];
    P.sab[ro + dst] = ab;
    {
        // pair-interleaved copy (k_pair_cells stages whole units), paired per layer
        const int dxy = st.dims[0] * st.dims[1];
        // (= c / dxy without the integer division: the layer cell_index used)
        const int L = cell_layer(st, p.z, P.cells_per_cut) - st.zoff;
        const int lstart = start[L * dxy], lend = start[(L + 1) * dxy];
        const int us = unit_slot(dst, L, lstart);
        float *u0 = unit_half(P, r, us, 0), *u1 = unit_half(P, r, us, 1);
        u0[0] = p.x;
        u0[2] = p.y;
        u1[0] = p.z;
        u1[2] = ab.x;
        unit_half(P, r, us, 2)[0] = ab.y;
        if (dst == lend - 1) {
            // the half behind the last bead of a layer (it shares a unit with that bead when
            // the layer holds an odd number): a dummy bead 64 cells off the grid with zero
            // mixing factors (finite r^2, force exactly 0)
            float *d0 = unit_half(P, r, us + 1, 0), *d1 = unit_half(P, r, us + 1, 1);
            d0[0] = st.origin[0] - 64.0f * st.cell_size;
            d0[2] = st.origin[1];
            d1[0] = st.origin[2];
            d1[2] = 0.0f;
            unit_half(P, r, us + 1, 2)[0] = 0.0f;
        }
    }
}

__global__ void __launch_bounds__(256) k_gather(const __grid_constant__ Params P)
{
    GC_PDL_ENTRY();
    int r = blockIdx.y;
    RepState &st = P.st[r];
    if (!st.need_rebuild) return;
    int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (P.decomp) {
        // the grid need not cover N: stride over the slots this rank binned
        const int total = P.cell_start[st.ncell];
        for (; s < total; s += gridDim.x * blockDim.x) gather_slot(P, st, r, s);
        return;
    }
    if (s >= P.n) return;
    gather_slot(P, st, r, s);
}

// ---------------------------------------------------------------------------
// Small systems (the ~600 / ~1,200-bead mitotic stages): the four rebuild kernels are
// pure launch latency there, so ONE block per replica does count -> scan -> fill ->
// gather with block barriers in between.  Same data structures, same sorted order
// (rank = ascending bead id), hence bit-identical forces.
// ---------------------------------------------------------------------------
constexpr int SMALL_BS = 1024;
constexpr int SMALL_MAX_BEADS = 4096;
constexpr int SMALL_MAX_CELLS = 32768;

// executed by all SMALL_BS threads of a block; sh: 32 ints of shared memory
__device__ inline void block_rebuild(const Params &P, int r, int *sh)
{
    RepState &st = P.st[r];
    if (!st.need_rebuild) return;                            // block-uniform
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const size_t ro = (size_t) r * P.n;
    int *count = P.cell_count + (size_t) r * P.ncell_max;
    int *start = P.cell_start + (size_t) r * (P.ncell_max + 1);
    const float4 *pos = P.pos[st.parity] + ro;
    const int ncell = st.ncell;
    if (tid == 0) { st.n_used_cells = 0; st.n_light_cells = 0; }

    // 1. cell of every bead, histogram (counts are zero on entry: gather_slot's invariant)
    for (int i = tid; i < P.n; i += SMALL_BS) {
        const int c = cell_index(st, pos[i], P.cells_per_cut);
        P.cell_of[ro + i] = c;
        atomicAdd(&count[c], 1);
    }
    __syncthreads();

    // 2. exclusive scan, counts back to 0: every thread owns `per` consecutive cells (a
    //    multiple of 4, read as int4; the second pass re-reads its own lines from L1), so
    //    the whole scan costs two block barriers whatever the cell count
    {
        const int per = (((ncell + SMALL_BS - 1) / SMALL_BS) + 3) & ~3;
        const int c0 = min(tid * per, ncell), c1 = min(c0 + per, ncell);
        int tsum = 0;
        for (int c = c0; c < c1; c += 4) {
            if (c + 3 < c1) {
                const int4 v = *reinterpret_cast<const int4 *>(count + c);
                tsum += v.x + v.y + v.z + v.w;
            } else {
                for (int k = c; k < c1; k++) tsum += count[k];
            }
        }
        int x = tsum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int y = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o) x += y;
        }
        if (lane == 31) sh[wid] = x;
        __syncthreads();
        if (wid == 0) {
            int w = sh[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int y = __shfl_up_sync(0xffffffffu, w, o);
                if (lane >= o) w += y;
            }
            sh[lane] = w;                                    // inclusive over warps
        }
        __syncthreads();
        int e = (wid ? sh[wid - 1] : 0) + x - tsum;
        for (int c = c0; c < c1; c += 4) {
            if (c + 3 < c1) {
                const int4 v = *reinterpret_cast<const int4 *>(count + c);
                *reinterpret_cast<int4 *>(count + c) = make_int4(0, 0, 0, 0);
                int4 o4;
                o4.x = e; e += v.x;
                o4.y = e; e += v.y;
                o4.z = e; e += v.z;
                o4.w = e; e += v.w;
                start[c] = o4.x; start[c + 1] = o4.y; start[c + 2] = o4.z; start[c + 3] = o4.w;
            } else {
                for (int k = c; k < c1; k++) {
                    const int v = count[k];
                    count[k] = 0;
                    start[k] = e;
                    e += v;
                }
            }
        }
        if (tid == SMALL_BS - 1) start[ncell] = e;           // = total: empty chunks carry it
    }
    __syncthreads();

    // 3. members of every cell (arbitrary order), 4. sorted copies (deterministic order)
    for (int i = tid; i < P.n; i += SMALL_BS) {
        const int c = P.cell_of[ro + i];
        P.members[ro + start[c] + atomicAdd(&count[c], 1)] = i;
    }
    __syncthreads();
    for (int s = tid; s < P.n; s += SMALL_BS) gather_slot(P, st, r, s);
}

__global__ void __launch_bounds__(SMALL_BS) k_rebuild_small(const __grid_constant__ Params P)
{
    GC_PDL_ENTRY();
    __shared__ int sh[32];
    block_rebuild(P, blockIdx.x, sh);
}

} // namespace gc

// gc_cells.cuh -- nonbonded soft-core forces, warp-per-cell (the production pair kernel).
//
// Replaces the pair loop of md::neighbor_pairwise_forcefield
// (MD/forcefield/neighbor_pairwise_forcefield.hpp:68-84) together with the
// Verlet list it walks (MD/forcefield/detail/neighbor_list.hpp:162-230): no pair
// list exists here.  Each WARP pops non-empty home cells from a per-replica work
// queue (Params::cell_list, filled by k_gather) and, privately, with no block
// barrier anywhere:
//   1. lanes 0..8 fetch the slot ranges of the 9 contiguous x-rows of 3 neighbouring
//      cells (the grid carries a one-cell halo, so there are no bounds checks) and turn
//      them into disjoint ranges of 48-byte candidate UNITS (two sorted slots each);
//   2. the warp stages the units into ITS OWN shared-memory tile, 32 units (64
//      candidates) per iteration: 3 LDG.128, one packed bounding-box cull of both beads,
//      ballot compaction, 3 STS.128;
//   3. lanes split into H home beads x S candidate slices (H * S <= 32, chosen per
//      cell; cells with more than 16 beads are served in groups of 16, each with its own
//      culled tile) and evaluate EVERY staged candidate branch-free: the clamped soft
//      cores vanish identically beyond their diameter, so there is no distance
//      test, no ballot and no compaction in the loop (packed FP32x2 math);
//   4. S - 1 shuffles per component fold the slices and lanes < H store the force
//      of their home bead (sorted-slot order, consumed by k_finish).
// Every bead gathers its own force (full neighbourhood, no atomics), so the
// result is bit-reproducible and independent of the queue order.
#pragma once

#include <cstddef>

#include "gc_step.cuh"

namespace gc {

constexpr int CW = 4;                    // warps per block
#ifndef GC_CCAP
#define GC_CCAP 320
#endif
#ifndef GC_CELL_MINBLOCKS
#define GC_CELL_MINBLOCKS 8
#endif
constexpr int CCAP = GC_CCAP;            // candidates per tile chunk of the contact kernel
// kRecip[d] = 65536 / d + 1: x / d == (x * kRecip[d]) >> 16 for x * d < 65536 (lane splits,
// slice counts and tile lengths are all far below); index 0 is unused
__constant__ unsigned int kRecip[33] = {0, 65537, 32769, 21846, 16385, 13108, 10923, 9363, 8193, 7282, 6554, 5958, 5462, 5042, 4682, 4370, 4097, 3856, 3641, 3450, 3277, 3121, 2979, 2850, 2731, 2622, 2521, 2428, 2341, 2260, 2185, 2115, 2049};
// Home beads per staged tile.  (A 32-bead tile evaluated in two 16-bead passes was
// measured against this: the second pass' live registers cost more than re-staging the
// few cells with more than 16 beads, and the per-group bounding box culls tighter.)
constexpr int CHMAX = 16;

// Per-warp shared-memory tile of the contact kernel (gc_contact.cuh): candidates staged
// component-wise, plus the row tables of its flattening walk.
template<int CAP>
struct CellWarpSmemT {
    static_assert(CAP % 2 == 0, "candidates are staged in pairs");
    float2 tx[CAP / 2], ty[CAP / 2], tz[CAP / 2];
    float2 tA[CAP / 2], tB[CAP / 2];     // bead index of the candidate / ring of queued hits
    int row_off[33];                     // first flattened index of each row (+ sentinel)
    int row_shift[32];                   // slot = flattened index + row_shift[row]
};
using CellWarpSmem = CellWarpSmemT<CCAP>;
// The pair kernel exists in two shapes: GC_CELL_MINBLOCKS = 8 resident blocks per SM (64
// registers) for launches that own the GPU, and 10 (48 registers) for the replica-group
// branches of the step graph, where several pair kernels share the SMs and the extra
// resident warps pay (profiles/r1_notes.md).  Both use the same tile size, hence the same
// chunking of long candidate lists and the same summation order: a trajectory does not
// depend on which shape ran a step.
#ifndef GC_DENSE_MINBLOCKS
#define GC_DENSE_MINBLOCKS 10
#endif
constexpr int MINBLOCKS_DENSE = GC_DENSE_MINBLOCKS;

// Sum over the S slices of each home bead (lane = slice * H + home bead), as a fixed binary
// tree over the slices: ceil(log2 S) rounds instead of S - 1; valid in lanes < H.
__device__ __forceinline__ void cell_fold(int H, int S, int sl, float &fx, float &fy, float &fz)
{
    // (sl = slice of this lane; lanes >= H * S carry zeros and never add)
    int step = 1;
    while (step < S) step <<= 1;
    for (step >>= 1; step > 0; step >>= 1) {
        const float gx = __shfl_down_sync(0xffffffffu, fx, step * H);
        const float gy = __shfl_down_sync(0xffffffffu, fy, step * H);
        const float gz = __shfl_down_sync(0xffffffffu, fz, step * H);
        if (sl < step && sl + step < S) { fx += gx; fy += gy; fz += gz; }
    }
}

// ---------------------------------------------------------------------------
// The production pair kernel (round 2).  Same algorithm as round 1's kernel; what changed
// is the data movement.  ncu on that kernel (profiles/r1_pair_cells_v8_sparse_ncu_summary.txt) and on
// two intermediate kernels (profiles/r2_pair_cells_v9_*, r2_pair_bulk_rejected_*): 43 % of the
// instructions were per-cell set-up, 34-45 % moved candidates one at a time through registers
// (flatten -> LDG -> cull -> ballot -> 5 st.shared each), 24 % were the pair loop.
//   * k_gather leaves, beside the sorted float4 positions, a PAIR-INTERLEAVED copy of the
//     sorted beads, one 48-byte unit per two slots (2u, 2u + 1):
//         (x0 x1 y0 y1) (z0 z1 a0 a1) (b0 b1 - -)
//     -- exactly the operand layout of the packed FP32x2 pair loop.  Staging moves UNITS:
//     3 LDG.128 in, one packed bounding-box test for both beads, 3 STS.128 out, for 64
//     candidates per warp iteration (was 32, at 2.5 x the instructions).
//   * the 9 candidate rows become disjoint unit ranges (rows may touch in slot space); the
//     flattened unit index is mapped to its row by a 4-step binary search over the row ends,
//     which stay in lanes 0..15 (shuffles, no shared-memory tables, no per-lane walk).
//     A unit may carry one bead of a cell OUTSIDE the 27 at the outer end of a run of rows;
//     it lies more than a cell edge >= cutoff from the home cell, where both clamped soft
//     cores are exactly zero (and it usually fails the cull anyway).
//   * the loop evaluates e' = x_j - x_i against the negated home position, the clamps are
//     min(g', 0) with g' = -g: no operand negation per pair; the sum is negated once.
//   * every slice runs `for (m = s; m < units; m += S)`: no reciprocal trip count, no scalar
//     tail loop (the unused half of an odd last unit is a far dummy written by k_gather).
//   * the home bounding box is reduced on the raw bits of (x - origin) >= 0: one REDUX per
//     bound.
// A bulk-async-copy variant (cp.async.bulk of whole rows onto an mbarrier, no culling) was
// built and measured slower: see profiles/r2_notes.md.
// ---------------------------------------------------------------------------
#ifndef GC_TILE_UNITS
#define GC_TILE_UNITS 112
#endif
constexpr int TILE_UNITS = GC_TILE_UNITS;        // candidate pairs per tile chunk (48 bytes each)

struct PairUnit {
    float4 xy, zA, B;
};
static_assert(sizeof(PairUnit) == 48, "unit layout");

// One home bead against one candidate unit (two candidates, packed).
template<bool PLAIN>
__device__ __forceinline__ void unit_eval(float4 a, float4 b, float2 bj, float nx, float ny, float nz,
                                          float Ai, float Bi, float2 kA, float2 kB, float2 &ax,
                                          float2 &ay, float2 &az)
{
    const float2 mone = make_float2(-1.0f, -1.0f);
    const float2 ex = __fadd2_rn(make_float2(a.x, a.y), make_float2(nx, nx));       // e' = x_j - x_i
    const float2 ey = __fadd2_rn(make_float2(a.z, a.w), make_float2(ny, ny));
    const float2 ez = __fadd2_rn(make_float2(b.x, b.y), make_float2(nz, nz));
    const float2 r2 = __ffma2_rn(ex, ex, __ffma2_rn(ey, ey, __fmul2_rn(ez, ez)));
    float2 gA = __ffma2_rn(r2, kA, mone);                           // -(1 - r2 kA2)
    gA = make_float2(fminf(gA.x, 0.0f), fminf(gA.y, 0.0f));
    float2 fac = __fmul2_rn(__fadd2_rn(make_float2(Ai, Ai), make_float2(b.z, b.w)), __fmul2_rn(gA, gA));
    if (!PLAIN) {
        const float2 uB = __fmul2_rn(r2, kB);
        const float2 uB2 = __fmul2_rn(uB, uB);
        float2 gB = __ffma2_rn(uB2, uB2, mone);                     // -(1 - uB^4)
        gB = make_float2(fminf(gB.x, 0.0f), fminf(gB.y, 0.0f));
        const float2 tB = __fmul2_rn(__fmul2_rn(gB, gB), __fmul2_rn(uB2, uB));
        fac = __ffma2_rn(__fadd2_rn(make_float2(Bi, Bi), bj), tB, fac);
    }
    ax = __ffma2_rn(fac, ex, ax);
    ay = __ffma2_rn(fac, ey, ay);
    az = __ffma2_rn(fac, ez, az);
}

// Home beads of one lane: TWO, so that every unit fetched from shared memory serves four pair
// evaluations.  (With one bead per lane the loop was bound by the shared-memory data pipe,
// not by issue: ncu l1tex__data_pipe_lsu_wavefronts 69 % of peak, issue 61 %; an LDS.128
// costs four wavefronts however many lanes share an address.)
#ifndef GC_HOME_PER_LANE
#define GC_HOME_PER_LANE 1
#endif
constexpr int HB = GC_HOME_PER_LANE;     // home beads per lane (1 or 2)
#ifndef GC_PASS_UNROLL
#define GC_PASS_UNROLL (HB == 1 ? 2 : 1)
#endif
constexpr int kPassUnroll = GC_PASS_UNROLL;
struct HomePair {
    float nx[HB], ny[HB], nz[HB], A[HB], B[HB];
    float2 ax[HB], ay[HB], az[HB];
};

// lanes = H2 home-bead groups (HB beads each) x S slices; slice s takes units s, s + S, ...
template<bool PLAIN>
__device__ __forceinline__ void unit_pass(const PairUnit *tile, int nu, int S, int s, float kA2, float kB2,
                                          HomePair &h)
{
    const float2 kA = make_float2(kA2, kA2), kB = make_float2(kB2, kB2);
#pragma unroll kPassUnroll
    for (int m = s; m < nu; m += S) {
        const float4 a = tile[m].xy;
        const float4 b = tile[m].zA;
        float2 bj = make_float2(0.f, 0.f);
        if (!PLAIN) bj = *reinterpret_cast<const float2 *>(&tile[m].B);
#pragma unroll
        for (int j = 0; j < HB; j++) {
            unit_eval<PLAIN>(a, b, bj, h.nx[j], h.ny[j], h.nz[j], h.A[j], h.B[j], kA, kB, h.ax[j], h.ay[j], h.az[j]);
        }
    }
}

#ifndef GC_TICKET
#define GC_TICKET 1
#endif
constexpr int TICKET = GC_TICKET;       // cells per work-queue claim (1 or 2)
static_assert(TICKET == 1 || TICKET == 2, "cells per claim");
constexpr int CH2MAX = 16 * HB;          // home beads per staged tile (16 lanes x HB)

// read-only path for the launch-per-step kernels; plain loads when the arrays are rewritten
// inside the same kernel (the persistent small-system kernel, where they live in shared memory)
template<bool NC, typename T>
__device__ __forceinline__ T pair_ld(const T *p)
{
    if (NC) return __ldg(p);
    return *p;
}

// All home cells of replica r that this warp gets hold of: from the shared work queue
// (STATIC = false: atomic claims) or a static share of it (STATIC = true: warp `slot` of
// `nslots` takes cells slot, slot + nslots, ...).  NC selects the read-only load path (false when
// the arrays are rewritten inside the same kernel).  The arithmetic per home
// cell -- staging order, culling, lane split, fold -- is the same either way.
template<bool PLAIN, bool STATIC, bool NC>
__device__ __forceinline__ void pair_cells_replica(const Params &P, RepState &st, int r, PairUnit *tile, int lane,
                                                   int slot, int nslots)
{
    const unsigned lt_mask = (1u << lane) - 1u;
    int next_cell = slot * TICKET;
    (void) next_cell;
    const StepConst &sc = st.sc;
    const size_t ro = (size_t) r * P.n;
    const float4 *__restrict__ sp = P.spos[st.parity] + ro;
    const float2 *__restrict__ sab = P.sab + ro;
    const float4 *__restrict__ uxy = P.punit + (size_t) r * P.n_units;
    const float4 *__restrict__ uzA = uxy + P.unit_plane;
    const float4 *__restrict__ uB = uzA + P.unit_plane;
    float4 *__restrict__ pforce = P.pforce + ro;
    const int *__restrict__ start = P.cell_start + (size_t) r * (P.ncell_max + 1);
    const int4 *__restrict__ cells = P.cell_list + ro;
    const int nheavy = st.n_used_cells, nused = nheavy + st.n_light_cells;
    const int dx = st.dims[0], dxy = st.dims[0] * st.dims[1];
    const float kA2 = sc.kA2, kB2 = PLAIN ? 0.0f : sc.kB2;
    const float hA = 0.5f * sc.cA, hB = PLAIN ? 0.0f : 0.5f * sc.cB;
    // squared cutoff of the pair term (the larger soft core), with a hair of margin
    const float cull2 = 1.0002f / (PLAIN ? kA2 : fminf(kA2, kB2));
    const float ox = st.origin[0], oy = st.origin[1], oz = st.origin[2];
    // row `lane` of the 9 neighbour rows (dy, dz in -1..1), relative to the home cell; the
    // grid has a one-cell halo (plan_grid), so every row exists
    const int row_rel = lane < 9 ? (lane % 3 - 1) * dx + (lane / 3 - 1) * dxy - 1 : 0;

    for (;;) {
        // Work queue: GC_TICKET cells per claim.  The dependent chain ticket -> cell
        // descriptor -> row starts (three L2 round trips) is paid once per claim, and the
        // loads of the second cell are in flight while the first is evaluated.
        int kk0 = 0;
        if (STATIC) {
            // one block owns the replica: warp `slot` of `nslots` takes cells slot, slot + nslots, ...
            kk0 = next_cell;
            next_cell += nslots * TICKET;
        } else {
            if (lane == 0) kk0 = atomicAdd(&st.cell_cursor, TICKET);
            kk0 = __shfl_sync(0xffffffffu, kk0, 0);
        }
        if (kk0 >= nused) break;
        // (scalars, not arrays: the second cell's five values stay in registers)
        const int kA_ = kk0, kB_ = min(kk0 + 1, nused - 1);
        const int4 cellA = pair_ld<NC>(&cells[kA_ < nheavy ? kA_ : P.n - 1 - (kA_ - nheavy)]);
        int4 cellB = cellA;
        if (TICKET == 2) cellB = pair_ld<NC>(&cells[kB_ < nheavy ? kB_ : P.n - 1 - (kB_ - nheavy)]);
        int loA = 0, u1A = 0, loB = 0, u1B = 0;
        if (lane < 9) {
            // slots -> the layer-relative unit space of the pair-interleaved copy (unit_slot)
            const int LA = cellA.w + lane / 3 - 1;
            const int sA = unit_slot(0, LA, pair_ld<NC>(&start[LA * dxy]));
            loA = pair_ld<NC>(&start[cellA.x + row_rel]) + sA;
            u1A = (pair_ld<NC>(&start[cellA.x + row_rel + 3]) + sA + 1) >> 1;
            if (TICKET == 2) {
                const int LB = cellB.w + lane / 3 - 1;
                const int sB = unit_slot(0, LB, pair_ld<NC>(&start[LB * dxy]));
                loB = pair_ld<NC>(&start[cellB.x + row_rel]) + sB;
                u1B = (pair_ld<NC>(&start[cellB.x + row_rel + 3]) + sB + 1) >> 1;
            }
        }
        int h0 = cellA.y, nhome = cellA.z, lo = loA, u1 = u1A;
#pragma unroll 1
        for (int t = 0; t < TICKET && kk0 + t < nused; t++) {
        if (t == 1) { h0 = cellB.y; nhome = cellB.z; lo = loB; u1 = u1B; }

        // ---- 1. candidate rows as DISJOINT unit ranges ------------------------------
        // Rows ascend in slot space and may touch (everything between two rows empty): a
        // row starts no earlier than its predecessor ended, so the union covers every unit
        // that holds a bead of the 27 cells exactly once.
        const int prev = __shfl_up_sync(0xffffffffu, u1, 1);
        const int u0 = max(lo >> 1, lane > 0 ? prev : 0);
        const int nu = lane < 9 ? max(u1 - u0, 0) : 0;
        int incl = nu;                                         // lanes 9..15 end up holding U
#pragma unroll
        for (int o = 1; o < 16; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += v;
        }
        const int U = __shfl_sync(0xffffffffu, incl, 8);
        const int shift = u0 - (incl - nu);                    // unit = flattened index + shift(row)
        const int last = U - 1;

        for (int g0 = 0; g0 < nhome; g0 += CH2MAX) {           // home beads, <= 32 per pass
            const int ng = min(CH2MAX, nhome - g0);
            // lanes = H2 home-bead pairs x S candidate slices
            const int H2 = (ng + HB - 1) / HB;
            const unsigned rh = kRecip[H2];
            const int S = (int) ((32u * rh) >> 16);
            const int sl = (int) (((unsigned) lane * rh) >> 16), hl = lane - sl * H2;

            // Bounding box of this group's home beads: a candidate further than the cutoff
            // from the box is further than the cutoff from every home bead, where the clamped
            // soft cores are exactly zero.  x - origin >= one cell edge > 0 for every bead
            // (plan_grid puts the origin a cell below the bounding box), so the raw bits of
            // the shifted coordinates order like the floats: one REDUX per bound.
            float2 blx, bly, blz, bhx, bhy, bhz;
            {
                const float4 hp = pair_ld<NC>(&sp[h0 + g0 + min(lane, ng - 1)]);
                const unsigned ux = __float_as_uint(hp.x - ox), uy = __float_as_uint(hp.y - oy);
                const unsigned uz = __float_as_uint(hp.z - oz);
                float v;
                v = __uint_as_float(__reduce_min_sync(0xffffffffu, ux)) + ox; blx = make_float2(v, v);
                v = __uint_as_float(__reduce_min_sync(0xffffffffu, uy)) + oy; bly = make_float2(v, v);
                v = __uint_as_float(__reduce_min_sync(0xffffffffu, uz)) + oz; blz = make_float2(v, v);
                v = -(__uint_as_float(__reduce_max_sync(0xffffffffu, ux)) + ox); bhx = make_float2(v, v);
                v = -(__uint_as_float(__reduce_max_sync(0xffffffffu, uy)) + oy); bhy = make_float2(v, v);
                v = -(__uint_as_float(__reduce_max_sync(0xffffffffu, uz)) + oz); bhz = make_float2(v, v);
            }
            HomePair hm;
#pragma unroll
            for (int j = 0; j < HB; j++) hm.ax[j] = hm.ay[j] = hm.az[j] = make_float2(0.f, 0.f);

            // ---- 2. stage: flattened units, culled by the box, compacted ------------------
            int nt = 0;
            for (int kin = 0; kin < U; kin += 32) {
                const int k = min(kin + lane, last);
                // row of k = number of row ends <= k: binary search over the ends held in
                // lanes 0..15 (4 shuffles), then the row's shift (1 shuffle)
                int row = 0;
#pragma unroll
                for (int step = 8; step > 0; step >>= 1) {
                    const int t = __shfl_sync(0xffffffffu, incl, row + step - 1);
                    row += k >= t ? step : 0;
                }
                const int u = k + __shfl_sync(0xffffffffu, shift, row);
                const float4 a = pair_ld<NC>(&uxy[u]);
                float4 b = pair_ld<NC>(&uzA[u]);
                float4 bb = make_float4(0.f, 0.f, 0.f, 0.f);
                if (!PLAIN) bb = pair_ld<NC>(&uB[u]);
                // distance of both beads of the unit from the box (packed)
                const float2 x2 = make_float2(a.x, a.y), y2 = make_float2(a.z, a.w), z2 = make_float2(b.x, b.y);
                const float2 lx = __fadd2_rn(blx, make_float2(-x2.x, -x2.y)), hx = __fadd2_rn(x2, bhx);
                const float2 ly = __fadd2_rn(bly, make_float2(-y2.x, -y2.y)), hy = __fadd2_rn(y2, bhy);
                const float2 lz = __fadd2_rn(blz, make_float2(-z2.x, -z2.y)), hz = __fadd2_rn(z2, bhz);
                const float2 ex = make_float2(fmaxf(fmaxf(lx.x, hx.x), 0.0f), fmaxf(fmaxf(lx.y, hx.y), 0.0f));
                const float2 ey = make_float2(fmaxf(fmaxf(ly.x, hy.x), 0.0f), fmaxf(fmaxf(ly.y, hy.y), 0.0f));
                const float2 ez = make_float2(fmaxf(fmaxf(lz.x, hz.x), 0.0f), fmaxf(fmaxf(lz.y, hz.y), 0.0f));
                const float2 d2 = __ffma2_rn(ex, ex, __ffma2_rn(ey, ey, __fmul2_rn(ez, ez)));
#ifdef GC_NO_CULL
                const bool keep = kin + lane < U;
#else
                const bool keep = (kin + lane < U) & (fminf(d2.x, d2.y) <= cull2);
#endif
                const unsigned m = __ballot_sync(0xffffffffu, keep);
                if (keep) {
                    PairUnit *dst = tile + (nt + __popc(m & lt_mask));
                    dst->xy = a;
                    b.z = PLAIN ? hA : hA * b.z;               // A' = 1/2 cA a_j
                    b.w = PLAIN ? hA : hA * b.w;
                    dst->zA = b;
                    if (!PLAIN) {
                        bb.x *= hB;                            // B' = 1/2 cB b_j
                        bb.y *= hB;
                        dst->B = bb;
                    }
                }
                nt += __popc(m);
                if (nt > TILE_UNITS - 32 || kin + 32 >= U) {   // tile full or input exhausted
                    __syncwarp();
                    // ---- 3. H2 home pairs x S slices, branch-free ---------------------
                    // (the home beads' constants are loaded here, not before staging: their
                    //  registers would otherwise be live across the staging loop)
                    if (sl < S) {
#pragma unroll
                        for (int j = 0; j < HB; j++) {
                            const int t = h0 + g0 + min(HB * hl + j, ng - 1);
                            const float4 me = pair_ld<NC>(&sp[t]);
                            hm.nx[j] = -me.x; hm.ny[j] = -me.y; hm.nz[j] = -me.z;
                            hm.A[j] = hA; hm.B[j] = 0.0f;
                            if (!PLAIN) {
                                const float2 abi = pair_ld<NC>(&sab[t]);
                                hm.A[j] = hA * abi.x; hm.B[j] = hB * abi.y;
                            }
                        }
                        unit_pass<PLAIN>(tile, nt, S, sl, kA2, kB2, hm);
                    }
                    __syncwarp();                              // tile fully consumed
                    nt = 0;
                }
            }
            // ---- 4. fold the slices, store (F_i = -sum fac e') ----------------------------
#pragma unroll
            for (int j = 0; j < HB; j++) {
                float fx = -(hm.ax[j].x + hm.ax[j].y), fy = -(hm.ay[j].x + hm.ay[j].y);
                float fz = -(hm.az[j].x + hm.az[j].y);
                cell_fold(H2, S, sl, fx, fy, fz);
                if (lane < H2 && HB * lane + j < ng) pforce[h0 + g0 + HB * lane + j] = make_float4(fx, fy, fz, 0.0f);
            }
        }
        }   // cells of this claim
    }
}

template<bool PLAIN, int MINB>
__global__ void __launch_bounds__(CW * 32, MINB) k_pair_cells(const __grid_constant__ Params P, unsigned mask_arg,
                                                        int use_enabled)
{
    GC_PDL_ENTRY();
    __shared__ __align__(16) PairUnit s_tile[CW][TILE_UNITS];

    const unsigned mask = use_enabled ? P.enabled : (mask_arg & P.enabled);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const bool pair_on = P.pair.mode != PAIR_NONE && (mask >> P.pair.term & 1u);

    if (!pair_on) {                                            // block-uniform
        const size_t total = (size_t) P.n_rep * P.n;
        for (size_t t = (size_t) blockIdx.x * (CW * 32) + threadIdx.x; t < total;
             t += (size_t) gridDim.x * (CW * 32)) {
            P.pforce[t] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        return;
    }

    // Blocks are not tied to a replica: every warp works through the queues of all
    // replicas round-robin, starting at a different one.
    for (int rr = 0; rr < P.n_rep; rr++) {
        const int r = (blockIdx.x + rr) % P.n_rep;
#ifdef GC_STATIC_CLAIM
        pair_cells_replica<PLAIN, true, true>(P, P.st[r], r, s_tile[warp], lane, blockIdx.x * CW + warp,
                                              gridDim.x * CW);
#else
        pair_cells_replica<PLAIN, false, true>(P, P.st[r], r, s_tile[warp], lane, 0, 1);
#endif
    }
}

} // namespace gc

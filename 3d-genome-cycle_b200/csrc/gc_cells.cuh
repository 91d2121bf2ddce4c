// gc_cells.cuh -- nonbonded soft-core forces, warp-per-cell (the production
// pair kernel).
//
// Replaces the pair loop of md::neighbor_pairwise_forcefield
// (MD/forcefield/neighbor_pairwise_forcefield.hpp:68-84) together with the
// Verlet list it walks (MD/forcefield/detail/neighbor_list.hpp:162-230): no pair
// list exists here.  Each WARP pops non-empty home cells from a per-replica work
// queue (Params::cell_list, filled by k_gather) and, privately, with no block
// barrier anywhere:
//   1. lanes 0..(2M+1)^2-1 fetch the slot ranges of the contiguous x-rows of the
//      neighbouring cells (9 rows of 3 cells for M = 1; the grid carries an M-cell
//      halo, so there are no bounds checks); a warp scan flattens them into one
//      candidate list;
//   2. the warp stages that list into ITS OWN shared-memory tile, coalesced per
//      row, as (x, y, z, A') + B' with the A/B mixing factors pre-scaled;
//   3. lanes split into H home beads x S candidate slices (H * S <= 32, chosen per
//      cell; cells with more than 16 beads are served in groups of 16, each with its own
//      culled tile) and evaluate EVERY staged candidate branch-free: the clamped soft
//      cores vanish identically beyond their diameter, so there is no distance
//      test, no ballot and no compaction (packed FP32x2 math, 16.5 issued
//      instructions per pair);
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
#ifndef GC_PACK_UNROLL
#define GC_PACK_UNROLL 2
#endif
constexpr int kPackUnroll = GC_PACK_UNROLL;   // candidate pairs in flight per lane
constexpr int CCAP = GC_CCAP;            // candidates per tile chunk
// kRecip[d] = 65536 / d + 1: x / d == (x * kRecip[d]) >> 16 for x * d < 65536 (lane splits,
// slice counts and tile lengths are all far below); index 0 is unused
__constant__ unsigned int kRecip[33] = {0, 65537, 32769, 21846, 16385, 13108, 10923, 9363, 8193, 7282, 6554, 5958, 5462, 5042, 4682, 4370, 4097, 3856, 3641, 3450, 3277, 3121, 2979, 2850, 2731, 2622, 2521, 2428, 2341, 2260, 2185, 2115, 2049};
// Home beads per staged tile.  (A 32-bead tile evaluated in two 16-bead passes was
// measured against this: the second pass' live registers cost more than re-staging the
// few cells with more than 16 beads, and the per-group bounding box culls tighter.)
constexpr int CHMAX = 16;

// One soft-core pair, branchless (see header).  sA = 0.5 cA (a_i + a_j) and
// sB = 0.5 cB (b_i + b_j) arrive pre-scaled:
//   A: 6 eps_A kA2 (1 - u)^2,  B: 24 eps_B kB2 (1 - u^4)^2 u^3   (gc_math.cuh)
__device__ __forceinline__ void pair_accumulate(float xi, float yi, float zi, float Ai, float Bi,
                                                float4 q, float Bj, float kA2, float kB2,
                                                float &fx, float &fy, float &fz)
{
    const float ex = xi - q.x, ey = yi - q.y, ez = zi - q.z;
    const float r2 = fmaf(ex, ex, fmaf(ey, ey, ez * ez));
    const float gA = fmaxf(fmaf(-r2, kA2, 1.0f), 0.0f);
    const float uB = r2 * kB2;
    const float uB2 = uB * uB;
    const float gB = fmaxf(fmaf(-uB2, uB2, 1.0f), 0.0f);
    const float tA = gA * gA;
    const float tB = (gB * gB) * (uB2 * uB);
    const float fac = fmaf(Bi + Bj, tB, (Ai + q.w) * tA);
    fx = fmaf(fac, ex, fx);
    fy = fmaf(fac, ey, fy);
    fz = fmaf(fac, ez, fz);
}

// The tile is stored as PAIRS of candidates per component so that the hot loop can
// use Blackwell's packed FP32x2 instructions (fma.rn.f32x2 & co., SASS FFMA2 / FMUL2 /
// FADD2): the FP32 pipe still retires 128 lanes per clock, but every issued
// instruction now carries two pair evaluations, which takes the kernel off the
// issue-slot limit (profiles/r1_notes.md).
template<int CAP>
struct CellWarpSmemT {
    static_assert(CAP % 2 == 0, "candidates are staged in pairs");
    float2 tx[CAP / 2], ty[CAP / 2], tz[CAP / 2];
    float2 tA[CAP / 2], tB[CAP / 2];     // pre-scaled mixing factors A', B'
    int row_off[33];                     // first flattened index of each row (+ sentinel)
    int row_shift[32];                   // slot = flattened index + row_shift[row]
};
using CellWarpSmem = CellWarpSmemT<CCAP>;
// The pair kernel exists in two shapes: GC_CELL_MINBLOCKS = 8 resident blocks per SM (64
// registers) for launches that own the GPU, and 10 (48 registers) for the replica-group
// branches of the step graph, where several pair kernels share the SMs and the extra
// resident warps pay (profiles/r1_notes.md).  Both use the same 256-candidate tile
// (10 blocks x 21 KB fit an SM), hence the same chunking of long candidate lists and the same
// summation order: a trajectory does not depend on which shape ran a step.
#ifndef GC_PAIR_CAP
#define GC_PAIR_CAP 256
#endif
#ifndef GC_DENSE_MINBLOCKS
#define GC_DENSE_MINBLOCKS 10
#endif
constexpr int PAIR_CAP = GC_PAIR_CAP, MINBLOCKS_DENSE = GC_DENSE_MINBLOCKS;

__device__ __forceinline__ float2 neg2(float2 v) { return make_float2(-v.x, -v.y); }
__device__ __forceinline__ float2 max0(float2 v) { return make_float2(fmaxf(v.x, 0.0f), fmaxf(v.y, 0.0f)); }

// Home-bead pass over the staged tile [0, nt): lanes = H beads x S slices; slice s
// takes candidate pairs s, s + S, s + 2 S, ... (packed math), then the ragged tail
// one candidate at a time.
template<class SM>
__device__ __forceinline__ void cell_pass(const SM &sm, int nt, int S, int s,
                                          float4 me, float Ai, float Bi, float kA2, float kB2,
                                          float &fx, float &fy, float &fz)
{
    // iterations in which every slice has a whole pair: (nt / 2) / S by reciprocal
    // multiplication (exact for nt / 2 <= 2047, S <= 32)
    const int full = (int) (((unsigned) (nt >> 1) * kRecip[S]) >> 16);
    const float2 xi = make_float2(me.x, me.x), yi = make_float2(me.y, me.y), zi = make_float2(me.z, me.z);
    const float2 Ai2 = make_float2(Ai, Ai), Bi2 = make_float2(Bi, Bi);
    const float2 nkA = make_float2(-kA2, -kA2), kB = make_float2(kB2, kB2), one = make_float2(1.0f, 1.0f);
    float2 ax = make_float2(0.f, 0.f), ay = ax, az = ax;
#pragma unroll kPackUnroll
    for (int it = 0; it < full; it++) {
        const int p = s + it * S;
        const float2 ex = __fadd2_rn(xi, neg2(sm.tx[p]));
        const float2 ey = __fadd2_rn(yi, neg2(sm.ty[p]));
        const float2 ez = __fadd2_rn(zi, neg2(sm.tz[p]));
        const float2 r2 = __ffma2_rn(ex, ex, __ffma2_rn(ey, ey, __fmul2_rn(ez, ez)));
        const float2 gA = max0(__ffma2_rn(r2, nkA, one));
        const float2 uB = __fmul2_rn(r2, kB);
        const float2 uB2 = __fmul2_rn(uB, uB);
        const float2 gB = max0(__ffma2_rn(neg2(uB2), uB2, one));
        const float2 tA = __fmul2_rn(gA, gA);
        const float2 tB = __fmul2_rn(__fmul2_rn(gB, gB), __fmul2_rn(uB2, uB));
        const float2 fac = __ffma2_rn(__fadd2_rn(Bi2, sm.tB[p]), tB,
                                      __fmul2_rn(__fadd2_rn(Ai2, sm.tA[p]), tA));
        ax = __ffma2_rn(fac, ex, ax);
        ay = __ffma2_rn(fac, ey, ay);
        az = __ffma2_rn(fac, ez, az);
    }
    fx += ax.x + ax.y;
    fy += ay.x + ay.y;
    fz += az.x + az.y;
    // ragged tail: candidates [2 full S, nt), fewer than 2 S + 1 of them
    const float *sx = reinterpret_cast<const float *>(sm.tx), *sy = reinterpret_cast<const float *>(sm.ty);
    const float *sz = reinterpret_cast<const float *>(sm.tz), *sA = reinterpret_cast<const float *>(sm.tA);
    const float *sB = reinterpret_cast<const float *>(sm.tB);
    for (int k = 2 * full * S + s; k < nt; k += S) {
        pair_accumulate(me.x, me.y, me.z, Ai, Bi, make_float4(sx[k], sy[k], sz[k], sA[k]), sB[k], kA2,
                        kB2, fx, fy, fz);
    }
}

// Sum over the S slices of each home bead, in fixed order; valid in lanes < H.
__device__ __forceinline__ void cell_fold(int H, int S, int lane, float &fx, float &fy, float &fz)
{
    for (int k = 1; k < S; k++) {
        const int src = (lane + k * H) & 31;
        const float gx = __shfl_sync(0xffffffffu, fx, src);
        const float gy = __shfl_sync(0xffffffffu, fy, src);
        const float gz = __shfl_sync(0xffffffffu, fz, src);
        if (lane < H) { fx += gx; fy += gy; fz += gz; }
    }
}

template<bool PLAIN, int MINB>
__global__ void __launch_bounds__(CW * 32, MINB) k_pair_cells(const __grid_constant__ Params P, unsigned mask_arg,
                                                        int use_enabled)
{
    GC_PDL_ENTRY();
    constexpr int CAP = PAIR_CAP;
    using CellSmem = CellWarpSmemT<CAP>;
    __shared__ CellSmem s_warp[CW];

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

    const int M = P.cells_per_cut;
    const int side = 2 * M + 1, nrows = side * side;
    constexpr bool plain = PLAIN;          // one soft core, no A/B mixing (mitotic stages)
    CellSmem &sm = s_warp[warp];
    // tx, ty, tz, tA, tB are consecutive arrays of CAP floats (static_assert below)
    const unsigned tile_addr = (unsigned) __cvta_generic_to_shared(sm.tx);
    static_assert(offsetof(CellSmem, ty) == sizeof(float2) * (CAP / 2) &&
                  offsetof(CellSmem, tz) == 2 * sizeof(float2) * (CAP / 2) &&
                  offsetof(CellSmem, tA) == 3 * sizeof(float2) * (CAP / 2) &&
                  offsetof(CellSmem, tB) == 4 * sizeof(float2) * (CAP / 2), "tile layout");

    // Blocks are not tied to a replica: every warp works through the queues of all
    // replicas round-robin, starting at a different one, so an ensemble is balanced
    // across the whole GPU whatever part of the grid is resident.
    for (int rr = 0; rr < P.n_rep; rr++) {
    const int r = (blockIdx.x + rr) % P.n_rep;
    RepState &st = P.st[r];
    const StepConst &sc = st.sc;
    const size_t ro = (size_t) r * P.n;
    const float4 *__restrict__ sp = P.spos[st.parity] + ro;
    const float2 *__restrict__ sab = P.sab + ro;
    float4 *__restrict__ pforce = P.pforce + ro;
    const int *__restrict__ start = P.cell_start + (size_t) r * (P.ncell_max + 1);
    const int4 *__restrict__ cells = P.cell_list + ro;
    const int nheavy = st.n_used_cells, nused = nheavy + st.n_light_cells;
    const int dx = st.dims[0], dxy = st.dims[0] * st.dims[1];
    const float kA2 = sc.kA2, kB2 = sc.kB2;
    const float hA = 0.5f * sc.cA, hB = plain ? 0.0f : 0.5f * sc.cB;
    // squared cutoff of the pair term (the larger soft core), with a hair of margin
    const float cull2 = 1.0002f / (plain ? kA2 : fminf(kA2, kB2));
    const unsigned lt_mask = (1u << lane) - 1u;
    // row `lane` of the (2M+1)^2 neighbour rows, relative to the home cell: the
    // grid has an M-cell halo (plan_grid), so every row exists
    const int row_rel = lane < nrows ? (lane % side - M) * dx + (lane / side - M) * dxy - M : 0;

    for (;;) {
        int kk = 0;
        if (lane == 0) kk = atomicAdd(&st.cell_cursor, 1);
        kk = __shfl_sync(0xffffffffu, kk, 0);
        if (kk >= nused) break;
        const int4 cell = __ldg(&cells[kk < nheavy ? kk : P.n - 1 - (kk - nheavy)]);
        const int c = cell.x, h0 = cell.y, nhome = cell.z;

        // ---- 1. candidate rows -> flattened list -------------------------------
        int lo = 0, len = 0;
        if (lane < nrows) {
            lo = __ldg(&start[c + row_rel]);
            len = __ldg(&start[c + row_rel + side]) - lo;
        }
        int incl = len;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += v;
        }
        const int total = __shfl_sync(0xffffffffu, incl, 31);
        __syncwarp();                                          // previous cell fully consumed
        sm.row_off[lane] = incl - len;                         // lanes >= nrows hold `total`
        sm.row_shift[lane] = lo - (incl - len);
        if (lane == 0) sm.row_off[32] = total;
        __syncwarp();

        for (int g0 = 0; g0 < nhome; g0 += CHMAX) {            // home beads, <= 16 per tile
            const int ng = min(CHMAX, nhome - g0);
            // lanes = ng home beads x S candidate slices
            const unsigned rh = kRecip[ng];
            const int S = (int) ((32u * rh) >> 16);
            const int sl = (int) (((unsigned) lane * rh) >> 16), hl = lane - sl * ng;
            const float4 me = __ldg(&sp[h0 + g0 + hl]);
            float Ai = hA, Bi = 0.0f;
            if (!plain) {
                const float2 abi = __ldg(&sab[h0 + g0 + hl]);
                Ai = hA * abi.x; Bi = hB * abi.y;
            }
            float fx = 0.f, fy = 0.f, fz = 0.f;

            // ---- 2. stage: flattened, coalesced within each row, culled, compacted --------
            // The loads of the next 32 candidates are issued before the current 32 are
            // tested and stored (and before the home beads' bounding box is reduced).
            // (tail lanes re-read the last candidate and are masked out of the ballot, so the
            // loop body has no predicated blocks)
            int kin = 0, row = 0, end = sm.row_off[1], shift = sm.row_shift[0];
            float4 qn;
            float2 abn = make_float2(0.f, 0.f);
            const int last = total - 1;
            auto fetch = [&](int k0) {
                const int k = min(k0 + lane, last);
                if (k >= end) {
                    do {
                        row++;
                        end = sm.row_off[row + 1];
                    } while (k >= end);
                    shift = sm.row_shift[row];
                }
                qn = __ldg(&sp[k + shift]);
                if (!plain) abn = __ldg(&sab[k + shift]);
            };
            fetch(0);
            // Bounding box of this group's home beads: a candidate further than the cutoff
            // from the box is further than the cutoff from every home bead, where both
            // clamped soft cores are exactly zero -- it is dropped while staging, which
            // removes 35-45 % of the branch-free pair evaluations.
            float blo[3], bhi[3];
            {
                const float4 hp = __ldg(&sp[h0 + g0 + min(lane, ng - 1)]);
                blo[0] = ordered_to_float(__reduce_min_sync(0xffffffffu, float_to_ordered(hp.x)));
                blo[1] = ordered_to_float(__reduce_min_sync(0xffffffffu, float_to_ordered(hp.y)));
                blo[2] = ordered_to_float(__reduce_min_sync(0xffffffffu, float_to_ordered(hp.z)));
                bhi[0] = ordered_to_float(__reduce_max_sync(0xffffffffu, float_to_ordered(hp.x)));
                bhi[1] = ordered_to_float(__reduce_max_sync(0xffffffffu, float_to_ordered(hp.y)));
                bhi[2] = ordered_to_float(__reduce_max_sync(0xffffffffu, float_to_ordered(hp.z)));
            }
            int nt = 0;
            while (kin < total) {
                const float4 q = qn;
                const float2 ab = abn;
                const bool v = kin + lane < total;
                kin += 32;
                if (kin < total) fetch(kin);
                const float ex = fmaxf(fmaxf(blo[0] - q.x, q.x - bhi[0]), 0.0f);
                const float ey = fmaxf(fmaxf(blo[1] - q.y, q.y - bhi[1]), 0.0f);
                const float ez = fmaxf(fmaxf(blo[2] - q.z, q.z - bhi[2]), 0.0f);
#ifdef GC_NO_CULL
                const bool keep = v;
#else
                const bool keep = v && fmaf(ex, ex, fmaf(ey, ey, ez * ez)) <= cull2;
#endif
                const unsigned m = __ballot_sync(0xffffffffu, keep);
                if (keep) {
                    // explicit shared-space stores: one address, five immediate offsets
                    const unsigned a = tile_addr + 4u * (unsigned) (nt + __popc(m & lt_mask));
                    constexpr unsigned kStride = (unsigned) (sizeof(float2) * (CAP / 2));
                    asm volatile("st.shared.f32 [%0], %1;" ::"r"(a), "f"(q.x) : "memory");
                    asm volatile("st.shared.f32 [%0], %1;" ::"r"(a + kStride), "f"(q.y) : "memory");
                    asm volatile("st.shared.f32 [%0], %1;" ::"r"(a + 2 * kStride), "f"(q.z) : "memory");
                    asm volatile("st.shared.f32 [%0], %1;" ::"r"(a + 3 * kStride), "f"(plain ? hA : hA * ab.x) : "memory");
                    asm volatile("st.shared.f32 [%0], %1;" ::"r"(a + 4 * kStride), "f"(plain ? 0.0f : hB * ab.y) : "memory");
                }
                nt += __popc(m);
                if (nt > CAP - 32 || kin >= total) {          // tile full or input exhausted
                    __syncwarp();
                    // ---- 3. H home beads x S slices, branch-free -------------------
                    cell_pass(sm, nt, S, sl, me, Ai, Bi, kA2, kB2, fx, fy, fz);
                    __syncwarp();                              // tile fully consumed
                    nt = 0;
                }
            }
            // ---- 4. fold the slices, store -----------------------------------------
            cell_fold(ng, S, lane, fx, fy, fz);
            if (lane < ng) pforce[h0 + g0 + lane] = make_float4(fx, fy, fz, 0.0f);
        }
    }
    }   // replicas
}

} // namespace gc

// gc_tile.cuh -- cell-tile step kernel (the production hot path).
//
// A thread block owns ONE home cell at a time (grid-stride over the cells of a
// replica).  All beads of a cell share the same 27 (M = 1) neighbouring cells,
// so the block
//   1. builds the flattened candidate list of the cell ONCE (9 contiguous
//      x-rows -> one list) and stages it in shared memory with coalesced loads:
//      positions as float4 and (a, b) factors as float2;
//   2. lets each warp walk that CONTIGUOUS shared-memory tile for one home bead
//      at a time: LDS.128 + distance test + ballot compaction into the warp's
//      ring buffer, dense evaluation of 32 hits per flush;
//   3. runs the single-bead terms, the noise and the BAOAB-limit update for the
//      cell's beads, one thread per bead.
// Compared with the warp-per-bead walk over global memory (gc_pairwalk.cuh,
// still used by the query kernels) the row bookkeeping is paid once per cell
// instead of once per bead and the hot loop has no address arithmetic left.
#pragma once

#include "gc_step.cuh"

namespace gc {

constexpr int TB = 128;                  // threads per block
constexpr int TW = TB / 32;              // warps per block
constexpr int TCAP = 768;                // candidates staged per tile chunk
constexpr int HCAP = 64;                 // home beads per batch

struct TileSmem {
    float4 tile[TCAP];                   // x, y, z, original index bits
    float2 tile_ab[TCAP];
    float4 queue[TW][PW_Q];
    float4 force[HCAP];
    int row_lo[32];
    int row_off[32];
    int total;
};

struct TileAcc {                         // per-thread accumulators over many beads
    float lo[3], hi[3];
    float disp2, dtmin;
    long long w[4];                      // wall statistics, fixed point
    bool bad;
};

__device__ __forceinline__ void tile_acc_init(TileAcc &a)
{
    for (int k = 0; k < 3; k++) { a.lo[k] = INFINITY; a.hi[k] = -INFINITY; }
    a.disp2 = 0.0f;
    a.dtmin = INFINITY;
    for (int k = 0; k < 4; k++) a.w[k] = 0;
    a.bad = false;
}

__device__ __forceinline__ void tile_acc_merge(TileAcc &a, const StepRed &s)
{
    for (int k = 0; k < 3; k++) { a.lo[k] = fminf(a.lo[k], s.lo[k]); a.hi[k] = fmaxf(a.hi[k], s.hi[k]); }
    a.disp2 = s.disp2 != s.disp2 ? INFINITY : fmaxf(a.disp2, s.disp2);
    a.bad |= s.bad || s.disp2 != s.disp2;
    a.dtmin = fminf(a.dtmin, s.dtmin);
    a.w[0] += __double2ll_rn((double) s.w.ax * GC_FX_SCALE);
    a.w[1] += __double2ll_rn((double) s.w.ay * GC_FX_SCALE);
    a.w[2] += __double2ll_rn((double) s.w.az * GC_FX_SCALE);
    a.w[3] += __double2ll_rn((double) s.w.reaction * GC_FX_SCALE);
}

// Block reduction of the accumulators + one set of atomics + ticket.
// Returns true in every thread of the LAST block of this replica.
__device__ __forceinline__ bool tile_commit(RepState &st, TileAcc &a, bool with_wall, bool with_geom,
                                            bool with_dt)
{
    __shared__ float sh_f[TW][8];
    __shared__ long long sh_l[TW][4];
    __shared__ int sh_bad, sh_last;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (threadIdx.x == 0) sh_bad = 0;
    float v[8];
    for (int k = 0; k < 3; k++) { v[k] = warp_min(a.lo[k]); v[3 + k] = warp_max(a.hi[k]); }
    v[6] = warp_max(a.disp2);
    v[7] = warp_min(a.dtmin);
    long long q[4];
    for (int k = 0; k < 4; k++) q[k] = warp_sum(a.w[k]);
    bool bad = __any_sync(0xffffffffu, a.bad);
    __syncthreads();
    if (lane == 0) {
        for (int k = 0; k < 8; k++) sh_f[wid][k] = v[k];
        for (int k = 0; k < 4; k++) sh_l[wid][k] = q[k];
        if (bad) sh_bad = 1;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < TW; w++) {
            for (int k = 0; k < 3; k++) {
                v[k] = fminf(v[k], sh_f[w][k]);
                v[3 + k] = fmaxf(v[3 + k], sh_f[w][3 + k]);
            }
            v[6] = fmaxf(v[6], sh_f[w][6]);
            v[7] = fminf(v[7], sh_f[w][7]);
            for (int k = 0; k < 4; k++) q[k] += sh_l[w][k];
        }
        if (with_geom) {
            for (int k = 0; k < 3; k++) {
                atomicMin(&st.bbox_min[k], float_to_ordered(v[k]));
                atomicMax(&st.bbox_max[k], float_to_ordered(v[3 + k]));
            }
            atomicMax(&st.maxdisp2_bits, __float_as_uint(v[6]));
        }
        if (with_dt) atomicMin(&st.dtmin_bits, __float_as_uint(v[7]));
        if (with_wall) {
            for (int k = 0; k < 3; k++) {
                if (q[k]) atomicAdd((unsigned long long *) &st.axial_fx[k], (unsigned long long) q[k]);
            }
            if (q[3]) atomicAdd((unsigned long long *) &st.reaction_fx, (unsigned long long) q[3]);
        }
        if (sh_bad) atomicExch(&st.error, 1);
        __threadfence();
        unsigned int prev = atomicAdd(&st.ticket, 1u);
        sh_last = prev == gridDim.x - 1;
        if (sh_last) __threadfence();
    }
    __syncthreads();
    return sh_last != 0;
}

// Walk the staged tile [0, ntile) for one home bead: ~22 instructions per 32
// candidates.  visit(dx, dy, dz, k) is called by exactly one lane per hit, k is
// the candidate's position in the tile.
template<typename Visitor>
__device__ __forceinline__ void tile_walk(const float4 *tile, int ntile, float4 self, float rc2,
                                          float4 *queue, int lane, Visitor &&visit)
{
    const unsigned lt_mask = (1u << lane) - 1u;
    int head = 0, qn = 0;
    for (int base = 0; base < ntile; base += 32) {
        const int k = base + lane;
        const float4 q = tile[min(k, ntile - 1)];
        float4 e;
        e.x = self.x - q.x;
        e.y = self.y - q.y;
        e.z = self.z - q.z;
        e.w = __int_as_float(k);
        const float r2 = fmaf(e.x, e.x, fmaf(e.y, e.y, e.z * e.z));
        const bool in = (r2 < rc2) & (k < ntile);
        const unsigned m = __ballot_sync(0xffffffffu, in);
        if (in) queue[(head + qn + __popc(m & lt_mask)) & (PW_Q - 1)] = e;
        qn += __popc(m);
        if (qn >= 32) {
            __syncwarp();
            const float4 h = queue[(head + lane) & (PW_Q - 1)];
            visit(h.x, h.y, h.z, __float_as_int(h.w));
            __syncwarp();
            head = (head + 32) & (PW_Q - 1);
            qn -= 32;
        }
    }
    __syncwarp();
    if (lane < qn) {
        const float4 h = queue[(head + lane) & (PW_Q - 1)];
        visit(h.x, h.y, h.z, __float_as_int(h.w));
    }
    __syncwarp();
}

// Nonbonded soft-core force of every bead -> P.pforce (sorted slot order).
__global__ void __launch_bounds__(TB, 8) k_pair_tile(Params *Pp, unsigned mask_arg, int use_enabled)
{
    __shared__ TileSmem sm;

    Params &P = *Pp;
    const int r = blockIdx.y;
    const RepState &st = P.st[r];
    const StepConst &sc = st.sc;
    const unsigned mask = use_enabled ? P.enabled : (mask_arg & P.enabled);
    const size_t ro = (size_t) r * P.n;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;

    const float4 *__restrict__ sp = P.spos[st.parity] + ro;
    const float2 *__restrict__ sab = P.sab + ro;
    float4 *__restrict__ pforce = P.pforce + ro;
    const int *__restrict__ start = P.cell_start + (size_t) r * (P.ncell_max + 1);
    const int dx = st.dims[0], dy = st.dims[1], dz = st.dims[2];
    const int ncell = st.ncell;
    const int M = P.cells_per_cut;
    const int side = 2 * M + 1, nrows = side * side;
    const bool pair_on = P.pair.mode != PAIR_NONE && (mask >> P.pair.term & 1u);
    const bool plain = P.pair.mode == PAIR_PLAIN;
    const PairCoef pc = {sc.kA2, sc.kB2, sc.cA, sc.cB, sc.eA, sc.eB, sc.rc2};

    if (!pair_on) {                                            // block-uniform
        for (int t = blockIdx.x * TB + threadIdx.x; t < P.n; t += gridDim.x * TB) {
            pforce[t] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        return;
    }

    for (int c = blockIdx.x; c < ncell; c += gridDim.x) {
        const int h0 = start[c], nhome = start[c + 1] - h0;
        if (nhome == 0) continue;                              // block-uniform

        // ---- 1. the cell's candidate rows -> flattened list -----------------
        __syncthreads();                                       // previous cell fully consumed
        if (warp == 0) {
            int lo = 0, len = 0;
            if (lane < nrows) {
                const int cx = c % dx, cy = (c / dx) % dy, cz = c / (dx * dy);
                const int y = cy + lane % side - M, z = cz + lane / side - M;
                if (y >= 0 && y < dy && z >= 0 && z < dz) {
                    const int row = dx * (y + dy * z);
                    lo = start[row + max(cx - M, 0)];
                    len = start[row + min(cx + M, dx - 1) + 1] - lo;
                }
            }
            int incl = len;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                int v = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += v;
            }
            sm.row_lo[lane] = lo;
            sm.row_off[lane] = incl - len;                     // lanes >= nrows: total (sentinel)
            if (lane == 31) sm.total = incl;
        }
        __syncthreads();
        const int total = sm.total;

        for (int c0 = 0; c0 < total; c0 += TCAP) {             // tile chunks (usually one)
            const int nt = min(TCAP, total - c0);
            if (c0 > 0) __syncthreads();                       // previous chunk fully consumed
            // ---- 2. stage the chunk (coalesced within each row) -------------
            {
                int row = 0;
                for (int k = c0 + threadIdx.x; k < c0 + nt; k += TB) {
                    while (k >= sm.row_off[row + 1]) row++;
                    const int slot = sm.row_lo[row] + (k - sm.row_off[row]);
                    sm.tile[k - c0] = __ldg(&sp[slot]);
                    if (!plain) sm.tile_ab[k - c0] = __ldg(&sab[slot]);
                }
            }
            __syncthreads();
            // ---- 3. every warp walks the tile for its home beads ------------
            for (int b = warp; b < nhome; b += TW) {
                const float4 self4 = sp[h0 + b];
                const float2 abi = plain ? make_float2(0.f, 0.f) : sab[h0 + b];
                float fx = 0.0f, fy = 0.0f, fz = 0.0f;
                tile_walk(sm.tile, nt, self4, pc.rc2, sm.queue[warp], lane,
                          [&](float ex, float ey, float ez, int k) {
                    // the bead itself has d = 0 and contributes exactly 0
                    float a = 1.0f, bb = 0.0f;
                    if (!plain) {
                        const float2 abj = sm.tile_ab[k];
                        a = 0.5f * (abi.x + abj.x);
                        bb = 0.5f * (abi.y + abj.y);
                    }
                    const float r2 = fmaf(ex, ex, fmaf(ey, ey, ez * ez));
                    const float fac = pair_force_factor(pc, a, bb, r2);
                    fx = fmaf(fac, ex, fx);
                    fy = fmaf(fac, ey, fy);
                    fz = fmaf(fac, ez, fz);
                });
                fx = warp_sum_f(fx);
                fy = warp_sum_f(fy);
                fz = warp_sum_f(fz);
                if (lane == 0) {
                    float4 f = make_float4(fx, fy, fz, 0.0f);
                    if (c0 > 0) {
                        const float4 prev = pforce[h0 + b];
                        f.x += prev.x; f.y += prev.y; f.z += prev.z;
                    }
                    pforce[h0 + b] = f;
                }
            }
        }
    }
}

// Single-bead terms + update (or force output), one thread per sorted slot, full
// occupancy; consumes the nonbonded force of k_pair_tile.
template<int MODE>
__global__ void __launch_bounds__(BS) k_finish(Params *Pp, unsigned mask_arg, int use_enabled)
{
    Params &P = *Pp;
    const int r = blockIdx.y;
    RepState &st = P.st[r];
    const unsigned mask = use_enabled ? P.enabled : (mask_arg & P.enabled);
    const int t = blockIdx.x * BS + threadIdx.x;
    StepRed red;
    red_init(red);

    if (t < P.n) {
        const size_t ro = (size_t) r * P.n;
        const float4 self4 = P.spos[st.parity][ro + t];
        const int idx = __float_as_int(self4.w);
        const float2 abi = P.sab[ro + t];
        f3 F = bead_force<false>(P, st, r, t, idx, self4, abi, mask, red.w);
        const float4 fp = P.pforce[ro + t];
        F.x += fp.x;
        F.y += fp.y;
        F.z += fp.z;
        if (MODE == STEP_FORCE) {
            P.force[ro + idx] = make_float4(F.x, F.y, F.z, 0.0f);
            if (P.spacestep != 0.0f) {
                red.dtmin = solve_timestep(P.spacestep, P.mob[idx], dot(F, F), P.kT, P.dt);
            }
            red.bad = !(fabsf(F.x) < INFINITY && fabsf(F.y) < INFINITY && fabsf(F.z) < INFINITY);
        } else {
            integrate_bead(P, st, r, t, idx, self4, F, P.dt, P.dt, red);
        }
    }

    const bool last = block_reduce_commit<BS>(st, red, true, MODE == STEP_FUSED, MODE == STEP_FORCE);
    if (last && threadIdx.x == 0) {
        finalize_stats(st);
        if (MODE == STEP_FUSED) {
            finalize_step(P, st, P.dt_cfg);
        } else {
            float dtc = P.dt;
            if (P.spacestep != 0.0f) {
                dtc = fminf(dtc, __uint_as_float(*(volatile unsigned int *) &st.dtmin_bits));
            }
            st.sc.dt_cur = dtc;
            st.dtmin_bits = 0x7f800000u;
            st.ticket = 0u;
        }
    }
}

} // namespace gc

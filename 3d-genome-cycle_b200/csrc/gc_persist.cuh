// gc_persist.cuh -- small systems (the ~600 / ~1,200-bead mitotic stages,
// SIM/stage_anatelophase/simulation_driver.cpp:240-310, stage_prometaphase/...:213-237), fixed
// dt: MANY steps per launch.
//
// The three kernels of a small-system step (k_pair_cells, k_finish, k_finalize_rebuild) are
// dependency chains of L2 round trips -- 10-14 us each for ~1,000 beads however few beads there
// are.  Here ONE 1,024-thread block per replica keeps everything a step rewrites in shared
// memory -- both position buffers, the sorted copy, the pair-interleaved copy, the pair forces,
// the work queue, RepState -- and runs pair forces -> single-bead terms + update -> callback
// body -> cell list with block barriers in between, for as many steps as the caller asked.
// Only the cell histogram / starts / member lists (sized by the grid, not by the beads) and the
// read-only topology stay in global memory.
//
// It is the SAME code: the block builds a copy of Params in shared memory whose pointers lead
// to the shared arrays (generic addressing) and calls pair_cells_replica, bead_force,
// integrate_bead, finalize_step and block_rebuild on it, so every sum is taken in the order of
// the launch-per-step path and the trajectory is bit-identical to it (GC_NO_PERSIST=1;
// tests/test_gpu_parity.py::test_persistent_small_system_kernel_is_bit_identical).
#pragma once

#include "gc_cells.cuh"
#include "gc_step.cuh"

namespace gc {

struct PersistLayout {       // byte offsets into the dynamic shared memory of k_persist_small
    unsigned q, st, sh, pos0, pos1, spos, pforce, sab, punit, cells, tiles, total;
    int pair_warps;
};

__host__ __device__ inline unsigned persist_align(unsigned x) { return (x + 15u) & ~15u; }

// pair_warps < 0: as many as fit into smem_max
__host__ __device__ inline PersistLayout persist_layout(int n, int n_units, int pair_warps, unsigned smem_max)
{
    PersistLayout L;
    unsigned o = 0;
    L.q = o;      o = persist_align(o + (unsigned) sizeof(Params));
    L.st = o;     o = persist_align(o + (unsigned) sizeof(RepState));
    L.sh = o;     o = persist_align(o + 32u * (unsigned) sizeof(int));
    L.pos0 = o;   o = persist_align(o + (unsigned) n * 16u);
    L.pos1 = o;   o = persist_align(o + (unsigned) n * 16u);
    L.spos = o;   o = persist_align(o + (unsigned) n * 16u);
    L.pforce = o; o = persist_align(o + (unsigned) n * 16u);
    L.sab = o;    o = persist_align(o + (unsigned) n * 8u);
    L.punit = o;  o = persist_align(o + 3u * (unsigned) n_units * 16u);
    L.cells = o;  o = persist_align(o + (unsigned) n * 16u);
    L.tiles = o;
    const unsigned per_warp = (unsigned) (TILE_UNITS * sizeof(PairUnit));
    int fit = smem_max > o ? (int) ((smem_max - o) / per_warp) : 0;
    if (fit > SMALL_BS / 32) fit = SMALL_BS / 32;
    L.pair_warps = pair_warps < 0 ? fit : (pair_warps < fit ? pair_warps : fit);
    L.total = o + (unsigned) (L.pair_warps > 0 ? L.pair_warps : 0) * per_warp;
    return L;
}

template<bool PLAIN>
__global__ void __launch_bounds__(SMALL_BS, 1) k_persist_small(const __grid_constant__ Params P, long long steps,
                                                               int pair_warps)
{
    GC_PDL_ENTRY();
    extern __shared__ __align__(16) unsigned char smem[];
    const int r = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const PersistLayout L = persist_layout(P.n, P.n_units, pair_warps, 0xffffffffu);
    Params &Q = *reinterpret_cast<Params *>(smem + L.q);
    RepState &S = *reinterpret_cast<RepState *>(smem + L.st);
    int *sh = reinterpret_cast<int *>(smem + L.sh);
    float4 *pos_s[2] = {reinterpret_cast<float4 *>(smem + L.pos0), reinterpret_cast<float4 *>(smem + L.pos1)};

    // ---- a Params whose hot arrays live in this block's shared memory, windowed to replica r
    {
        const int *src = reinterpret_cast<const int *>(&P);
        int *dst = reinterpret_cast<int *>(&Q);
        for (int k = tid; k < (int) (sizeof(Params) / sizeof(int)); k += SMALL_BS) dst[k] = src[k];
    }
    if (wid == 0) repstate_load(S, P.st[r], lane);
    __syncthreads();
    if (tid == 0) {
        const size_t o = (size_t) r * (size_t) P.n;
        Q.n_rep = 1;
        Q.st = &S;
        Q.pos[0] = pos_s[0];
        Q.pos[1] = pos_s[1];
        Q.spos[0] = Q.spos[1] = reinterpret_cast<float4 *>(smem + L.spos);   // no skin: only the current one is read
        Q.pforce = reinterpret_cast<float4 *>(smem + L.pforce);
        Q.sab = reinterpret_cast<float2 *>(smem + L.sab);
        Q.punit = reinterpret_cast<float4 *>(smem + L.punit);
        Q.unit_plane = (size_t) P.n_units;
        Q.cell_list = reinterpret_cast<int4 *>(smem + L.cells);
        Q.sref += o;                       // the rest stays in global memory, advanced to replica r
        Q.scell += o;
        Q.members += o;
        Q.cell_of += o;
        Q.force += o;
        Q.cell_count += (size_t) r * P.ncell_max;
        Q.cell_start += (size_t) r * (P.ncell_max + 1);
        Q.scan_status += (size_t) r * P.scan_blocks;
    }
    {
        const float4 *src = P.pos[S.parity] + (size_t) r * P.n;
        float4 *dst = pos_s[S.parity];
        for (int i = tid; i < P.n; i += SMALL_BS) dst[i] = src[i];
    }
    __syncthreads();
    const unsigned mask = Q.enabled;
    const bool pair_on = Q.pair.mode != PAIR_NONE && (mask >> Q.pair.term & 1u);
    if (!pair_on) {
        for (int i = tid; i < P.n; i += SMALL_BS) Q.pforce[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    // the sorted copies of the state the block starts from
    block_rebuild(Q, 0, sh);
    __syncthreads();

    for (long long step = 0; step < steps; step++) {
        // ---- 1. nonbonded forces: the first pair_warps warps share the replica's home cells
        if (pair_on && wid < pair_warps) {
            PairUnit *tile = reinterpret_cast<PairUnit *>(smem + L.tiles) + (size_t) wid * TILE_UNITS;
            pair_cells_replica<PLAIN, true>(Q, S, 0, tile, lane, wid, pair_warps);
        }
        __syncthreads();

        // ---- 2. single-bead terms, noise, update (k_finish<STEP_FUSED>) -----------------
        {
            StepRed red;
            red_init(red);
            for (int t = tid; t < P.n; t += SMALL_BS) {
                const float4 self4 = Q.spos[S.parity][t];
                const int idx = __float_as_int(self4.w);
                const float2 abi = Q.sab[t];
                StepRed one;
                red_init(one);
                f3 F = bead_force(Q, S, 0, t, idx, self4, abi, mask, one.w);
                const float4 fp = Q.pforce[t];
                F.x += fp.x;
                F.y += fp.y;
                F.z += fp.z;
                integrate_bead(Q, S, 0, t, idx, self4, F, Q.dt, one);
                for (int k = 0; k < 3; k++) {
                    red.lo[k] = fminf(red.lo[k], one.lo[k]);
                    red.hi[k] = fmaxf(red.hi[k], one.hi[k]);
                }
                red.bad |= one.bad;
            }
            // (no wall in the stages this kernel serves: the wall sums are not accumulated)
            block_reduce_commit<SMALL_BS>(S, red, false, true, false);
        }
        __syncthreads();

        // ---- 3. the per-step callback body (k_finalize_step), on the shared RepState ----
        if (wid == 0) {
            if (lane == 0) {
                finalize_stats(S);
                finalize_step(Q, S, Q.dt_cfg);
            }
            __syncwarp();
            refresh_consts_item(Q, S, lane);
        }
        __syncthreads();

        // ---- 4. cell list of the next step ----------------------------------------------
        block_rebuild(Q, 0, sh);
        __syncthreads();
    }

    // ---- back to global memory: positions, state (the caller re-bins the global copies)
    {
        float4 *dst = P.pos[S.parity] + (size_t) r * P.n;
        const float4 *src = pos_s[S.parity];
        for (int i = tid; i < P.n; i += SMALL_BS) dst[i] = src[i];
    }
    if (wid == 0) repstate_store(P.st[r], S, lane);
}

} // namespace gc

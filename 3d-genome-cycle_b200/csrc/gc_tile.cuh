// gc_tile.cuh -- k_finish: single-bead terms + update after the pair kernel.
//
// One thread per sorted slot, full occupancy: bonded pairs and triples, tethers,
// nucleolar droplet, point sources, wall (+ its reaction statistics), Philox
// noise and the BAOAB-limit update; consumes the nonbonded force k_pair_cells
// left in P.pforce.  The per-step callback body runs in k_finalize_step, the
// launch behind this one.
#pragma once

#include "gc_step.cuh"

namespace gc {

// Single-bead terms + update (or force output), one thread per sorted slot, full
// occupancy; consumes the nonbonded force of k_pair_cells.
#ifndef GC_FINISH_MINB
#define GC_FINISH_MINB 1
#endif
template<int MODE>
__global__ void __launch_bounds__(BS, GC_FINISH_MINB) k_finish(const __grid_constant__ Params P, unsigned mask_arg, int use_enabled)
{
    GC_PDL_ENTRY();
    const int r = blockIdx.y;
    RepState &st = P.st[r];
    const unsigned mask = use_enabled ? P.enabled : (mask_arg & P.enabled);
    // a decomposed run moves the sorted slots of the layers it owns (contiguous: z is the
    // slowest cell axis)
    int t = blockIdx.x * BS + threadIdx.x;
    int t_lo = 0, t_hi = P.n;
    if (P.decomp) {
        // the grid covers Params::down_bound slots from the first owned one (k_decomp_header
        // reports a rank that owns more)
        const int *start = P.cell_start + (size_t) r * (P.ncell_max + 1);
        const int dxy = st.dims[0] * st.dims[1];
        t_lo = start[st.zown_lo * dxy];
        t_hi = start[st.zown_hi * dxy];
        t += t_lo;
        if ((int) blockIdx.x * BS >= t_hi - t_lo) return;   // block-uniform
    }
    StepRed red;
    red_init(red);

    if (t >= t_lo && t < t_hi) {
        const size_t ro = (size_t) r * P.n;
        const float4 self4 = P.spos[st.parity][ro + t];
        const int idx = __float_as_int(self4.w);
        const float2 abi = P.sab[ro + t];
        f3 F = bead_force(P, st, r, t, idx, self4, abi, mask, red.w);
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
            integrate_bead(P, st, r, t, idx, self4, F, P.dt, red);
        }
    }

    block_reduce_commit<BS>(st, red, true, MODE == STEP_FUSED, MODE == STEP_FORCE);
}

} // namespace gc

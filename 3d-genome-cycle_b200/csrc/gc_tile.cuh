// gc_tile.cuh -- k_finish: single-bead terms + update after the pair kernel.
//
// One thread per sorted slot, full occupancy: bonded pairs and triples, tethers,
// nucleolar droplet, point sources, wall (+ its reaction statistics), Philox
// noise and the BAOAB-limit update; consumes the nonbonded force k_pair_cells
// left in P.pforce.  The last block of a replica runs the per-step callback
// body (finalize_step).
#pragma once

#include "gc_step.cuh"

namespace gc {

// Single-bead terms + update (or force output), one thread per sorted slot, full
// occupancy; consumes the nonbonded force of k_pair_tile.
template<int MODE>
__global__ void __launch_bounds__(BS) k_finish(const __grid_constant__ Params P, unsigned mask_arg, int use_enabled)
{
    GC_PDL_ENTRY();
    const int r = blockIdx.y;
    RepState &st = P.st[r];
    const unsigned mask = use_enabled ? P.enabled : (mask_arg & P.enabled);
    // a decomposed run launches only the blocks of the owned slot range
    const int t = (P.decomp ? P.own_lo : 0) + blockIdx.x * BS + threadIdx.x;
    StepRed red;
    red_init(red);

    if (t < (P.decomp ? P.own_hi : P.n)) {
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

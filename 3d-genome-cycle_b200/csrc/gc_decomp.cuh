// gc_decomp.cuh -- one large replica spread over several GPUs (BASELINE configs[4],
// SURVEY.md 8e row 2; not in the reference, which is single-process).
//
// Decomposition by SORTED SLOT RANGE.  Every rank holds all N positions and builds the
// same cell list (x-fastest, z-slowest, so a contiguous slot range is a z-slab of the
// nucleus with ragged edges).  Rank k evaluates the nonbonded forces of the cells that
// hold slots [k C, (k + 1) C), C = ceil(N / ranks), and moves exactly those beads
// (k_gather filters the work queue, k_finish the slots).  The new positions leave in
// slot order as (x, y, z, bead) in Params::dsend, followed by a 32-byte header with this
// rank's share of the wall reaction sums; ONE all-gather (NCCL over NVLink) puts every
// rank's buffer into Params::drecv, and k_decomp_scatter below writes all N beads back
// by bead index, rebuilds the bounding box and sums the headers in rank order.  Every
// bead's force is a gather over its full neighbourhood and the noise is keyed by (seed,
// bead, step), so the trajectory is bit-identical to the single-GPU one whatever the
// number of ranks (tests/test_decomp.py).
#pragma once

#include "gc_step.cuh"

namespace gc {

// after k_finish: this rank's partial wall statistics -> header of the send buffer
__global__ void k_decomp_header(const __grid_constant__ Params P)
{
    GC_PDL_ENTRY();
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    RepState &st = P.st[0];
    long long *h = reinterpret_cast<long long *>(P.dsend + P.dchunk);
    h[0] = st.axial_fx[0];
    h[1] = st.axial_fx[1];
    h[2] = st.axial_fx[2];
    h[3] = st.reaction_fx;
}

// after the all-gather: every bead's new position by bead index, bounding box of all
// beads (the next grid plan), wall statistics summed over ranks in rank order
__global__ void __launch_bounds__(256) k_decomp_scatter(const __grid_constant__ Params P)
{
    GC_PDL_ENTRY();
    RepState &st = P.st[0];
    float lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
    bool bad = false;
    float4 *__restrict__ out = P.pos[st.parity ^ 1];
    for (int t = blockIdx.x * 256 + threadIdx.x; t < P.n; t += gridDim.x * 256) {      // slot
        const int k = t / P.dchunk;
        const float4 v = P.drecv[(size_t) k * (P.dchunk + 2) + (t - k * P.dchunk)];
        out[__float_as_int(v.w)] = make_float4(v.x, v.y, v.z, 0.0f);
        lo[0] = fminf(lo[0], v.x); hi[0] = fmaxf(hi[0], v.x);
        lo[1] = fminf(lo[1], v.y); hi[1] = fmaxf(hi[1], v.y);
        lo[2] = fminf(lo[2], v.z); hi[2] = fmaxf(hi[2], v.z);
        bad |= !(fabsf(v.x) < INFINITY && fabsf(v.y) < INFINITY && fabsf(v.z) < INFINITY);
    }
    // block reduction, then 6 atomics per block
    __shared__ float sh[8][6];
    __shared__ int sh_bad;
    if (threadIdx.x == 0) sh_bad = 0;
    for (int k = 0; k < 3; k++) {
        for (int o = 16; o > 0; o >>= 1) {
            lo[k] = fminf(lo[k], __shfl_xor_sync(0xffffffffu, lo[k], o));
            hi[k] = fmaxf(hi[k], __shfl_xor_sync(0xffffffffu, hi[k], o));
        }
    }
    __syncthreads();
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (lane == 0) {
        for (int k = 0; k < 3; k++) { sh[wid][k] = lo[k]; sh[wid][3 + k] = hi[k]; }
    }
    if (bad) sh_bad = 1;
    __syncthreads();
    if (threadIdx.x < 3) {
        float l = sh[0][threadIdx.x], h = sh[0][3 + threadIdx.x];
        for (int w = 1; w < 8; w++) { l = fminf(l, sh[w][threadIdx.x]); h = fmaxf(h, sh[w][3 + threadIdx.x]); }
        if (l <= h) {
            atomicMin(&st.bbox_min[threadIdx.x], float_to_ordered(l));
            atomicMax(&st.bbox_max[threadIdx.x], float_to_ordered(h));
        }
    }
    if (threadIdx.x == 0 && sh_bad) atomicExch(&st.error, 1);
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        long long sum[4] = {0, 0, 0, 0};
        for (int k = 0; k < P.dworld; k++) {
            const long long *h = reinterpret_cast<const long long *>(P.drecv + (size_t) k * (P.dchunk + 2) + P.dchunk);
            for (int c = 0; c < 4; c++) sum[c] += h[c];
        }
        st.axial_fx[0] = sum[0];
        st.axial_fx[1] = sum[1];
        st.axial_fx[2] = sum[2];
        st.reaction_fx = sum[3];
    }
}

} // namespace gc

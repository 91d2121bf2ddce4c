// gc_decomp.cuh -- one large replica spread over several GPUs (BASELINE configs[4],
// SURVEY.md 8e row 2; not in the reference, which is single-process).
//
// SLAB + HALO decomposition.  The cell grid is x-fastest / z-slowest, so a range of z
// layers is a contiguous range of cells AND of sorted slots.  Rank k owns the layers
// [slab_start(k), slab_start(k + 1)) of the global grid (equal-volume slabs of the
// bounding ellipsoid) and bins only the beads whose position is CURRENT on it: its own
// beads plus the halo copies its two neighbours sent after the previous step.  Every bead
// carries the step at which this rank last wrote it (Params::dstamp); a bead that left
// the halo simply stops being current and drops out of the cell list.
//
// One step on rank k (send / receive scheme):
//   gc_decomp_step_compute   cell list over the window [z0 - 1, z1 + 1) built from the work
//                            list of current beads (k_count takes the halo copies received
//                            after the previous step in on its way), pair forces of the owned
//                            layers, k_finish moves the owned beads (new position in pos[] and,
//                            in slot order, spos[]; stamp = step + 1), k_decomp_header writes
//                            the rank's share of the per-step reductions (bounding box, wall
//                            sums, error, bead count) to a 128-byte header
//   all-gather of the headers (world x 128 bytes; NCCL)
//   gc_decomp_step_pack      k_finalize_step folds the headers (decomp_reduce: reductions,
//                            load balance), its per-step callback body plans the NEXT grid
//                            (identical on every rank); k_decomp_pack appends each moved bead
//                            whose new layer lies within dhalo layers of the neighbour's new
//                            slab -- or inside it: ownership follows the position -- to the
//                            send buffer of that neighbour, as (x, y, z, bead)
//   send/recv with rank - 1 and rank + 1 (NCCL point-to-point over NVLink)
// Peer mode (gc_decomp_peer_*): k_decomp_pack runs inside gc_decomp_step_compute, under the
// grid of the step just taken plus GC_DECOMP_PEER_MARGIN layers, and stores straight into the
// neighbours' double-buffered receive buffers over NVLink; the header all-gather is then the
// step's only collective, and the next k_count verifies the shipped set against the new grid.
//
// The exchange is O(surface): 2 neighbours x (beads in dhalo + 1 layers) x 16 bytes instead
// of round 1's all-gather of all N beads to every rank.  Every bead's force is still a
// gather over its complete neighbourhood, evaluated in the order of the same cell list
// arithmetic, and the noise is keyed by (seed, bead, step, epoch): the trajectory is
// bit-identical to the single-GPU one whatever the number of ranks (tests/test_decomp.py).
#pragma once

#include "gc_step.cuh"

namespace gc {

// after k_finish: this rank's share of the per-step reductions
__global__ void k_decomp_header(const __grid_constant__ Params P)
{
    GC_PDL_ENTRY();
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    RepState &st = P.st[0];
    double *h = P.dhdr;
    long long *hl = reinterpret_cast<long long *>(h);
    int *hi = reinterpret_cast<int *>(h + 4);
    hl[0] = st.axial_fx[0];
    hl[1] = st.axial_fx[1];
    hl[2] = st.axial_fx[2];
    hl[3] = st.reaction_fx;
    for (int k = 0; k < 3; k++) { hi[k] = st.bbox_min[k]; hi[3 + k] = st.bbox_max[k]; }
    hi[6] = st.error;
    // beads moved by this rank (diagnostics: the counts of all ranks must add up to N)
    const int *start = P.cell_start;
    const int dxy = st.dims[0] * st.dims[1];
    hi[7] = start[st.zown_hi * dxy] - start[st.zown_lo * dxy];
    if (hi[7] > P.down_bound && !hi[6]) hi[6] = st.error = 4;      // k_finish's grid did not reach them all
    // the work list of the next cell-list build: what this rank just moved (+ the halos)
    st.dprev_lo = start[st.zown_lo * dxy];
    st.dprev_hi = start[st.zown_hi * dxy];
    st.dfull = 0;
    // ... and the grid it was moved under (peer mode ships under this grid)
    st.dold_oz = st.origin[2];
    st.dold_inv = st.inv_cell;
    st.dold_gdimz = st.gdimz;
    st.dold_z0 = st.zoff + st.zown_lo;
    st.dold_z1 = st.zoff + st.zown_hi;
    // the send buffers restart empty
    reinterpret_cast<int *>(P.dsend_lo)[0] = 0;
    reinterpret_cast<int *>(P.dsend_hi)[0] = 0;
}

// Ships the moved beads the neighbours need.
//   NCCL mode: after the callback body planned the next grid; beads within dhalo layers of a
//     neighbour's NEW slab go to the send buffers, which the caller exchanges.
//   Peer mode: right after k_finish, under the grid of the step just taken with
//     GC_DECOMP_PEER_MARGIN layers more (the next k_count checks the set against the new grid), and straight into the
//     neighbour's receive buffer of this step's parity over NVLink; k_decomp_publish then
//     writes the counts.  The header all-gather that follows is the step's only collective
//     and orders these writes before the neighbour reads them.
__global__ void __launch_bounds__(256) k_decomp_pack(const __grid_constant__ Params P)
{
    GC_PDL_ENTRY();
    RepState &st = P.st[0];
    __shared__ int zs[4];       // slab starts of ranks drank - 1 .. drank + 2
    if (threadIdx.x < 4) zs[threadIdx.x] = slab_start(P, st, P.drank - 1 + threadIdx.x);
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const int peer = P.dpeer;
    // peer mode runs before the callback flipped the parity and advanced the step
    const float4 *__restrict__ moved = P.spos[peer ? st.parity ^ 1 : st.parity] + st.dprev_lo;
    const int reach = P.dhalo + (peer ? GC_DECOMP_PEER_MARGIN : 0);
    const int z0 = peer ? st.dold_z0 : zs[1], z1 = peer ? st.dold_z1 : zs[2];
    const int wb = (int) (st.step & 1);
    float4 *out_lo = peer ? P.dpeer_lo[wb] : P.dsend_lo;
    float4 *out_hi = peer ? P.dpeer_hi[wb] : P.dsend_hi;
    // the beads this rank moved in the step just taken, left in slot order by k_finish
    const int nw = st.dprev_hi - st.dprev_lo;
    const int nw_pad = (nw + 31) & ~31;             // whole warps: the appends below are warp-aggregated
    for (int w = blockIdx.x * 256 + threadIdx.x; w < nw_pad; w += gridDim.x * 256) {
        const bool live = w < nw;
        float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
        if (live) x = moved[w];                     // new position, bead index in .w
        const int layer = cell_layer(st, x.z, P.cells_per_cut);
        // a bead that jumped past a neighbour's whole slab would be lost (peer mode: k_count)
        if (!peer && live && (layer < zs[0] || layer >= zs[3])) atomicExch(&st.error, 32);
        const bool to_lo = live && P.drank > 0 && layer < z0 + reach;
        const bool to_hi = live && P.drank + 1 < P.dworld && layer >= z1 - reach;
#pragma unroll
        for (int side = 0; side < 2; side++) {
            const unsigned m = __ballot_sync(0xffffffffu, side ? to_hi : to_lo);
            if (!m) continue;
            // the count lives in the LOCAL send buffer in both modes
            int *counter = reinterpret_cast<int *>(side ? P.dsend_hi : P.dsend_lo);
            float4 *buf = side ? out_hi : out_lo;
            int base = 0;
            if (lane == __ffs(m) - 1) base = atomicAdd(counter, __popc(m));
            base = __shfl_sync(0xffffffffu, base, __ffs(m) - 1);
            if (side ? to_hi : to_lo) {
                const int k = base + __popc(m & ((1u << lane) - 1u));
                if (k < P.dcap) buf[1 + k] = x; else atomicExch(&st.error, 4);
            }
        }
    }
    // peer stores are posted writes over NVLink: every thread makes its own visible system-wide
    // before the grid -- and with it the stream order the all-gather hangs on -- completes
    if (peer) __threadfence_system();
}

// peer mode, after k_decomp_pack: the bead counts into the neighbours' buffers
__global__ void k_decomp_publish(const __grid_constant__ Params P)
{
    GC_PDL_ENTRY();
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    const RepState &st = P.st[0];
    const int wb = (int) (st.step & 1);
    if (P.drank > 0) reinterpret_cast<int *>(P.dpeer_lo[wb])[0] = min(reinterpret_cast<const int *>(P.dsend_lo)[0], P.dcap);
    if (P.drank + 1 < P.dworld) reinterpret_cast<int *>(P.dpeer_hi[wb])[0] = min(reinterpret_cast<const int *>(P.dsend_hi)[0], P.dcap);
    __threadfence_system();
}

// Received beads by bead index, stamped current.  The next k_count does this on its way;
// this kernel serves the queries between two steps (gc_decomp_owned).
__global__ void __launch_bounds__(256) k_decomp_unpack(const __grid_constant__ Params P)
{
    GC_PDL_ENTRY();
    RepState &st = P.st[0];
    float4 *__restrict__ pos = P.pos[st.parity];
    const int side = blockIdx.y;
    if (st.dfull) return;
    if (side == 0 && P.drank == 0) return;
    if (side == 1 && P.drank + 1 >= P.dworld) return;
    const int rb = decomp_recv_index(P, st);
    const float4 *buf = side == 0 ? P.drecv_lo[rb] : P.drecv_hi[rb];
    const int count = min(reinterpret_cast<const int *>(buf)[0], P.dcap);
    for (int k = blockIdx.x * 256 + threadIdx.x; k < count; k += gridDim.x * 256) {
        const float4 v = buf[1 + k];
        const int idx = __float_as_int(v.w);
        pos[idx] = make_float4(v.x, v.y, v.z, 0.0f);
        P.dstamp[idx] = (int) st.step;
    }
}

// gc_run(steps = 0, restart = 1): every rank was handed all N positions
__global__ void __launch_bounds__(256) k_decomp_stamp_all(const __grid_constant__ Params P)
{
    GC_PDL_ENTRY();
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i == 0) {
        P.st[0].dfull = 1;
        for (int k = 0; k <= P.dworld; k++) P.st[0].dfrac[k] = (double) k / (double) P.dworld;   // equal volumes
    }
    if (i < P.n) P.dstamp[i] = (int) P.st[0].step;
}

// mask of the beads this rank owns now (position current and inside its slab)
__global__ void __launch_bounds__(256) k_decomp_owned(const __grid_constant__ Params P, unsigned char *mask)
{
    RepState &st = P.st[0];
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= P.n) return;
    unsigned char own = 0;
    if (P.dstamp[i] == (int) st.step) {
        const int layer = cell_layer(st, P.pos[st.parity][i].z, P.cells_per_cut);
        own = layer >= slab_start(P, st, P.drank) && layer < slab_start(P, st, P.drank + 1);
    }
    mask[i] = own;
}

} // namespace gc

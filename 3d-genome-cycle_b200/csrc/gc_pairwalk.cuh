// gc_pairwalk.cuh -- the neighbour walk shared by the force, contact and
// pair-list kernels.
//
// ONE WARP walks the neighbourhood of ONE bead:
//   1. lanes 0 .. (2M+1)^2-1 each fetch the [lo, hi) slot range of one
//      contiguous x-row of neighbouring cells; a warp scan turns the row
//      lengths into offsets of ONE flattened candidate list (so short rows do
//      not waste lanes);
//   2. the warp streams that list 32 candidates at a time (coalesced float4
//      loads within a row), tests the distance, and COMPACTS the candidates
//      inside the cutoff (typically 15-25 %) into a ring buffer in shared
//      memory with a warp ballot;
//   3. whenever 32 hits have accumulated they are evaluated densely, one per
//      lane, through the visitor -- the potential math runs on full warps
//      instead of being predicated off most of the time.
//
// Replaces the pair loop of md::neighbor_pairwise_forcefield
// (MD/forcefield/neighbor_pairwise_forcefield.hpp:68-84) and the bucket x bucket
// loop of md::neighbor_searcher::search (MD/misc/neighbor_searcher.hpp:75-100).
#pragma once

#include "gc_kernels.cuh"

namespace gc {

constexpr int PW_Q = 64;                // ring-buffer entries per warp
constexpr int PW_MAXROWS = 25;          // (2M+1)^2 with M <= 2
constexpr int PW_ROWINTS = 2 * 32;      // per-warp ints: row lo[32] + row offset[32]

struct PairHit {                        // what the visitor receives
    float dx, dy, dz;                   // self - other
    int j;                              // slot (or original index, BY_INDEX) of the other bead
};

struct PairWalkSmem {                   // per warp
    float4 queue[PW_Q];
    int row_lo[32];
    int row_off[32];
};

__device__ __forceinline__ float warp_sum_f(float v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// All 32 lanes call this with the same (t, self).  visit(hit) is invoked by
// exactly one lane per in-range candidate.
// EXACT = true keeps candidates with r2 <= rc2 * (1 + 1e-4) so that the visitor
// can redo a borderline test in FP64; otherwise r2 < rc2.
// BY_INDEX = true reports ORIGINAL bead indices in hit.j and keeps only
// candidates whose original index is above the walking bead's (each unordered
// pair once, as md::neighbor_searcher does).
template<bool EXACT, bool BY_INDEX, typename Visitor>
__device__ __forceinline__ void pair_walk(const Params &P, const RepState &st, int r, int t,
                                          float4 self, float rc2, PairWalkSmem &sm, int lane,
                                          Visitor &&visit)
{
    const size_t ro = (size_t) r * P.n;
    const float4 *__restrict__ sp = P.spos[st.parity] + ro;
    const int *__restrict__ start = P.cell_start + (size_t) r * (P.ncell_max + 1);
    const int dx = st.dims[0], dy = st.dims[1], dz = st.dims[2];
    const int M = P.cells_per_cut;
    const int side = 2 * M + 1;
    const int nrows = side * side;
    const float lim = EXACT ? rc2 * 1.0001f : rc2;
    const int self_idx = __float_as_int(self.w);

    // ---- 1. row ranges -> flattened candidate list ---------------------------
    int c = P.scell[ro + t];
    int cx = c % dx, cy = (c / dx) % dy, cz = c / (dx * dy);
    int lo = 0, len = 0;
    if (lane < nrows) {
        int y = cy + lane % side - M, z = cz + lane / side - M;
        if (y >= 0 && y < dy && z >= 0 && z < dz) {
            int row = dx * (y + dy * z);
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
    const int total = __shfl_sync(0xffffffffu, incl, 31);
    __syncwarp();                       // previous bead's readers are done
    sm.row_lo[lane] = lo;
    sm.row_off[lane] = incl - len;      // lanes >= nrows hold `total`: the sentinel
    __syncwarp();

    // ---- 2. stream, test, compact ----------------------------------------------
    // Hot loop: ~30 instructions per 32 candidates.  The row of a lane's k-th
    // candidate only moves forward, so (end of current row, slot - k) live in
    // registers and shared memory is touched once per row change.
    const float4 *spo = sp;
    float4 *queue = sm.queue;
    asm volatile("" : "+l"(spo), "+l"(queue));     // keep both in registers (no rematerialisation)
    const unsigned lt_mask = (1u << lane) - 1u;
    int head = 0, qn = 0, row = -1, cur_end = 0, shift = 0;
    const int last = total - 1;
    for (int base = 0; base < total; base += 32) {
        const int k = base + lane;
        const int kk = min(k, last);                 // tail lanes re-test the last candidate
        if (kk >= cur_end) {
            do {
                row++;
                cur_end = sm.row_off[row + 1];
            } while (kk >= cur_end);
            shift = sm.row_lo[row] - sm.row_off[row];
        }
        const float4 q = __ldg(spo + (unsigned) (kk + shift));
        float4 e;
        e.x = self.x - q.x;
        e.y = self.y - q.y;
        e.z = self.z - q.z;
        e.w = BY_INDEX ? q.w : __int_as_float(kk + shift);
        const float r2 = fmaf(e.x, e.x, fmaf(e.y, e.y, e.z * e.z));
        bool in = (EXACT ? (r2 <= lim) : (r2 < lim)) & (k <= last);
        if (BY_INDEX) in = in & (__float_as_int(q.w) > self_idx);
        const unsigned m = __ballot_sync(0xffffffffu, in);
        if (in) queue[(head + qn + __popc(m & lt_mask)) & (PW_Q - 1)] = e;
        qn += __popc(m);
        if (qn >= 32) {
            // ---- 3. evaluate 32 hits densely -----------------------------------
            __syncwarp();
            float4 h = queue[(head + lane) & (PW_Q - 1)];
            visit(PairHit{h.x, h.y, h.z, __float_as_int(h.w)});
            __syncwarp();
            head = (head + 32) & (PW_Q - 1);
            qn -= 32;
        }
    }
    __syncwarp();
    if (lane < qn) {
        float4 h = queue[(head + lane) & (PW_Q - 1)];
        visit(PairHit{h.x, h.y, h.z, __float_as_int(h.w)});
    }
}

} // namespace gc

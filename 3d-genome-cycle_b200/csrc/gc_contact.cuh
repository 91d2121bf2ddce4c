// gc_contact.cuh -- contact-map update on the cell tiles (contact_map::update,
// SIM/stage_interphase/contact_map.cpp:33-63).
//
// Same front end as k_pair_cells (gc_cells.cuh): a warp pops a non-empty home cell,
// flattens its neighbour rows and stages (x, y, z, bead) of every candidate ONCE in its
// private shared-memory tile; lanes then split into H home beads x S candidate slices and
// test r^2 against the contact distance.  Hits (each unordered pair once: bead_j > bead_i)
// are ballot-compacted into a 64-entry ring and inserted 32 at a time, so the table
// atomics of a whole warp are in flight together.  The predicate is the one of k_pairs
// (gc_query.cuh): FP32 r^2 decides except inside a 1e-4 relative band around rc^2, where
// the test is redone in FP64 on the exactly representable FP32 coordinates -- the set
// equals the FP64 reference's (tests/test_gpu_parity.py::test_contact_update_exact).
// The previous version walked the cells once per BEAD (k_pairs<PAIRS_CONTACT>, candidates
// re-read from L2 by every bead).
#pragma once

#include "gc_cells.cuh"
#include "gc_query.cuh"

namespace gc {

__global__ void __launch_bounds__(CW * 32, GC_CELL_MINBLOCKS) k_contact_cells(const __grid_constant__ Params P, int radius_sel)
{
    GC_PDL_ENTRY();
    __shared__ CellWarpSmem s_warp[CW];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int M = P.cells_per_cut;
    const int side = 2 * M + 1, nrows = side * side;
    CellWarpSmem &sm = s_warp[warp];
    float *tx = reinterpret_cast<float *>(sm.tx), *ty = reinterpret_cast<float *>(sm.ty);
    float *tz = reinterpret_cast<float *>(sm.tz);
    int *tj = reinterpret_cast<int *>(sm.tA);                 // bead index of the candidate
    uint2 *q_ij = reinterpret_cast<uint2 *>(sm.tB);           // ring: (bead_i, bead_j)
    float *q_r2 = reinterpret_cast<float *>(sm.tB) + 128;     //       r^2
    static_assert(CCAP >= 192, "the hit ring lives in the unused B' tile");
    const unsigned lt_mask = (1u << lane) - 1u;

    for (int rr = 0; rr < P.n_rep; rr++) {
    const int r = (blockIdx.x + rr) % P.n_rep;
    RepState &st = P.st[r];
    const double rc = radius_sel == RADIUS_CONTACT ? st.contact_distance : st.contact_distance_prev;
    if (!(rc > 0.0)) continue;
    const double rc2d = rc * rc;
    const float rc2 = (float) rc2d;
    const float band = 1.0e-4f * rc2, lim = rc2 * 1.0001f;
    const size_t ro = (size_t) r * P.n;
    const float4 *__restrict__ sp = P.spos[st.parity] + ro;
    const float4 *__restrict__ pos = P.pos[st.parity] + ro;
    const int *__restrict__ start = P.cell_start + (size_t) r * (P.ncell_max + 1);
    const int4 *__restrict__ cells = P.cell_list + ro;
    const int nheavy = st.n_used_cells, nused = nheavy + st.n_light_cells;
    const int dx = st.dims[0], dxy = st.dims[0] * st.dims[1];
    const int row_rel = lane < nrows ? (lane % side - M) * dx + (lane / side - M) * dxy - M : 0;
    const float ox = st.origin[0], oy = st.origin[1], oz = st.origin[2];
    int head = 0, qn = 0;

    // one lane per queued hit: exact predicate, then the table insert
    auto flush = [&](int count) {
        __syncwarp();
        if (lane < count) {
            const uint2 ij = q_ij[(head + lane) & 63];
            const float r2 = q_r2[(head + lane) & 63];
            bool in = r2 <= rc2;
            if (fabsf(r2 - rc2) <= band) {
                const float4 a = pos[ij.x], b = pos[ij.y];
                const double fx = (double) a.x - (double) b.x, fy = (double) a.y - (double) b.y;
                const double fz = (double) a.z - (double) b.z;
                in = fx * fx + fy * fy + fz * fz <= rc2d;
            }
            if (in) contact_insert(P.contact, P.n, r, st, ij.x, ij.y);
        }
        __syncwarp();
    };

    for (;;) {
        int kk = 0;
        if (lane == 0) kk = atomicAdd(&st.contact_cursor, 1);
        kk = __shfl_sync(0xffffffffu, kk, 0);
        if (kk >= nused) break;
        const int4 cell = __ldg(&cells[kk < nheavy ? kk : P.n - 1 - (kk - nheavy)]);
        const int c = cell.x, h0 = cell.y, nhome = cell.z;

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
        __syncwarp();
        sm.row_off[lane] = incl - len;
        sm.row_shift[lane] = lo - (incl - len);
        if (lane == 0) sm.row_off[32] = total;
        __syncwarp();

        // Bounding box of the cell's home beads: a candidate further than the contact distance
        // from the box is further than that from every home bead.  x - origin > 0 for every
        // bead (the origin lies a cell below the bounding box), so the raw bits of the shifted
        // coordinates order like the floats: one REDUX per bound.
        float blo[3], bhi[3];
        {
            unsigned ulo[3] = {0xffffffffu, 0xffffffffu, 0xffffffffu}, uhi[3] = {0u, 0u, 0u};
            for (int g0 = 0; g0 < nhome; g0 += 32) {
                const float4 hp = __ldg(&sp[h0 + min(g0 + lane, nhome - 1)]);
                const unsigned u[3] = {__float_as_uint(hp.x - ox), __float_as_uint(hp.y - oy),
                                       __float_as_uint(hp.z - oz)};
                for (int k = 0; k < 3; k++) {
                    ulo[k] = min(ulo[k], __reduce_min_sync(0xffffffffu, u[k]));
                    uhi[k] = max(uhi[k], __reduce_max_sync(0xffffffffu, u[k]));
                }
            }
            blo[0] = __uint_as_float(ulo[0]) + ox; bhi[0] = __uint_as_float(uhi[0]) + ox;
            blo[1] = __uint_as_float(ulo[1]) + oy; bhi[1] = __uint_as_float(uhi[1]) + oy;
            blo[2] = __uint_as_float(ulo[2]) + oz; bhi[2] = __uint_as_float(uhi[2]) + oz;
        }

        for (int c0 = 0; c0 < total; c0 += CCAP) {             // tile chunks (usually one)
            const int nin = min(CCAP, total - c0);
            int nt = 0;                                        // candidates kept in the tile
            if (c0 > 0) __syncwarp();
            {
                int row = 0, end = sm.row_off[1], shift = sm.row_shift[0];
                for (int kk = c0; kk < c0 + nin; kk += 32) {
                    const int k = kk + lane;
                    bool keep = false;
                    float4 q = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (k < c0 + nin) {
                        if (k >= end) {
                            do {
                                row++;
                                end = sm.row_off[row + 1];
                            } while (k >= end);
                            shift = sm.row_shift[row];
                        }
                        q = __ldg(&sp[k + shift]);
                        const float ex = fmaxf(fmaxf(blo[0] - q.x, q.x - bhi[0]), 0.0f);
                        const float ey = fmaxf(fmaxf(blo[1] - q.y, q.y - bhi[1]), 0.0f);
                        const float ez = fmaxf(fmaxf(blo[2] - q.z, q.z - bhi[2]), 0.0f);
                        // (a hair of slack on top of the band limit: the box bounds were
                        //  rebuilt from shifted coordinates)
                        keep = fmaf(ex, ex, fmaf(ey, ey, ez * ez)) <= lim * 1.0002f;
                    }
                    const unsigned m = __ballot_sync(0xffffffffu, keep);
                    if (keep) {
                        const int t = nt + __popc(m & lt_mask);
                        tx[t] = q.x;
                        ty[t] = q.y;
                        tz[t] = q.z;
                        tj[t] = __float_as_int(q.w);
                    }
                    nt += __popc(m);
                }
            }
            __syncwarp();
            for (int g0 = 0; g0 < nhome; g0 += 32) {           // home beads: H x S lanes
                const int H = min(32, nhome - g0), S = 32 / H;
                const int s = (lane * (65536 / H + 1)) >> 16, h = lane - s * H;
                const bool live = s < S;
                const float4 me = __ldg(&sp[h0 + g0 + h]);
                const int bead_i = __float_as_int(me.w);
                const int iters = (nt + S - 1) / S;
                for (int it = 0; it < iters; it++) {
                    const int k = s + it * S;
                    bool in = false;
                    float r2 = 0.0f;
                    int bead_j = 0;
                    if (live && k < nt) {
                        const float ex = me.x - tx[k], ey = me.y - ty[k], ez = me.z - tz[k];
                        r2 = fmaf(ex, ex, fmaf(ey, ey, ez * ez));
                        bead_j = tj[k];
                        in = r2 <= lim && bead_j > bead_i;
                    }
                    const unsigned m = __ballot_sync(0xffffffffu, in);
                    if (in) {
                        const int slot = (head + qn + __popc(m & lt_mask)) & 63;
                        q_ij[slot] = make_uint2((unsigned) bead_i, (unsigned) bead_j);
                        q_r2[slot] = r2;
                    }
                    qn += __popc(m);
                    if (qn >= 32) {
                        flush(32);
                        head = (head + 32) & 63;
                        qn -= 32;
                    }
                }
            }
        }
    }
    flush(qn);
    }   // replicas
}

} // namespace gc

// gc_query.cuh -- pair enumeration (neighbor search output, contact map),
// potential energy, observables.  Same cell walk as the force kernel.
#pragma once

#include "gc_pairwalk.cuh"
#include "gc_step.cuh"

namespace gc {

#define GC_EMPTY_KEY 0xffffffffffffffffull
#define GC_MAX_PROBES 512

__device__ __forceinline__ unsigned long long hash64(unsigned long long k)
{
    // splitmix64 finaliser
    k ^= k >> 30; k *= 0xbf58476d1ce4e5b9ull;
    k ^= k >> 27; k *= 0x94d049bb133111ebull;
    k ^= k >> 31;
    return k;
}

// contact_map::update's `contacts[{i, j}] += 1`
// (SIM/stage_interphase/contact_map.cpp:52-58), i < j.
__device__ __forceinline__ void contact_insert(const ContactParams &C, int n_beads, int r,
                                               RepState &st, unsigned int i, unsigned int j)
{
    // ---- bead i's private bucket --------------------------------------------------
    const unsigned int S = (unsigned int) C.bucket_slots;
    unsigned long long *b = C.bucket + ((size_t) r * n_beads + i) * S;
    const unsigned int h = (j * 0x9E3779B1u) >> 16;
    for (unsigned int p = 0; p < S; p++) {
        unsigned long long *w = b + ((h + p) & (S - 1));
        unsigned long long cur = *(volatile unsigned long long *) w;
        if (cur == GC_EMPTY_KEY) {
            cur = atomicCAS(w, GC_EMPTY_KEY, ((unsigned long long) j << 32) | 1ull);
            if (cur == GC_EMPTY_KEY) {
                atomicAdd(&C.used[r], 1u);
                return;
            }
        }
        if ((unsigned int) (cur >> 32) == j) {
            atomicAdd(w, 1ull);          // the count lives in the low word
            return;
        }
    }
    // ---- bucket full: spill table, bounded linear probing ------------------------------
    if (st.error == 2) return;
    unsigned long long key = ((unsigned long long) i << 32) | j;
    unsigned long long mask = C.capacity - 1;
    unsigned long long *keys = C.keys + (size_t) r * C.capacity;
    unsigned int *vals = C.vals + (size_t) r * C.capacity;
    unsigned long long hh = hash64(key) & mask;
    for (int probe = 0; probe < GC_MAX_PROBES; probe++) {
        unsigned long long cur = keys[hh];
        if (cur == GC_EMPTY_KEY) {
            cur = atomicCAS(&keys[hh], GC_EMPTY_KEY, key);
            if (cur == GC_EMPTY_KEY) {
                atomicAdd(&C.used[r], 1u);
                cur = key;
            }
        }
        if (cur == key) {
            atomicAdd(&vals[hh], 1u);
            return;
        }
        hh = (hh + 1) & mask;
    }
    atomicExch(&st.error, 2);   // both levels full -> GC_ERR_CAPACITY
}

constexpr int SBS = 256;                 // threads per block of k_pairs
constexpr int SWARPS = SBS / 32;
constexpr int SBEADS = 32;               // beads per block
constexpr int SPERWARP = SBEADS / SWARPS;

enum PairsMode : int { PAIRS_LIST = 0, PAIRS_CONTACT = 1 };
enum RadiusSel : int { RADIUS_ARG = 0, RADIUS_CONTACT = 1, RADIUS_CONTACT_PREV = 2 };

// Emits every unordered pair (i < j, original indices) with |xi - xj|^2 <= rc^2
// exactly once -- the contract of md::neighbor_searcher::search
// (MD/misc/neighbor_searcher.hpp:75-100, inclusive test at line 90).  The FP32
// distance decides except inside a 1e-4 relative band around rc^2, where the
// test is redone in FP64 on the (exactly representable) FP32 coordinates, so
// the set equals the FP64 reference's on the same inputs.
template<int MODE>
__global__ void __launch_bounds__(SBS) k_pairs(const __grid_constant__ Params P, double rc_arg, int radius_sel,
                                               int replica, uint2 *out,
                                               unsigned long long out_cap,
                                               unsigned long long *out_count)
{
    GC_PDL_ENTRY();
    __shared__ PairWalkSmem s_walk[SWARPS];
    const int r = blockIdx.y;
    RepState &st = P.st[r];
    if (replica >= 0 && r != replica) return;
    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    double rc = radius_sel == RADIUS_ARG ? rc_arg
              : radius_sel == RADIUS_CONTACT ? st.contact_distance : st.contact_distance_prev;
    if (!(rc > 0.0)) return;
    const double rc2d = rc * rc;
    const float rc2 = (float) rc2d;
    const float band = 1.0e-4f * rc2;

    for (int b = 0; b < SPERWARP; b++) {
    const int t = blockIdx.x * SBEADS + warp * SPERWARP + b;
    if (t >= P.n) return;                       // warp-uniform
    const float4 self = P.spos[st.parity][(size_t) r * P.n + t];
    const int idx = __float_as_int(self.w);

    pair_walk<true, true>(P, st, r, t, self, rc2, s_walk[warp], lane, [&](PairHit h) {
        float r2 = fmaf(h.dx, h.dx, fmaf(h.dy, h.dy, h.dz * h.dz));
        bool in = r2 <= rc2;
        if (fabsf(r2 - rc2) <= band) {
            // FP32 differences of FP32 inputs within a factor 2 are exact; redo
            // the squares and the comparison in FP64 from the original coordinates
            float4 q = P.pos[st.parity][(size_t) r * P.n + h.j];
            double fx = (double) self.x - (double) q.x;
            double fy = (double) self.y - (double) q.y;
            double fz = (double) self.z - (double) q.z;
            in = fx * fx + fy * fy + fz * fz <= rc2d;
        }
        if (!in) return;
        if (MODE == PAIRS_CONTACT) {
            contact_insert(P.contact, P.n, r, st, (unsigned) idx, (unsigned) h.j);
        } else {
            unsigned long long k = atomicAdd(out_count, 1ull);
            if (k < out_cap) out[k] = make_uint2((unsigned) idx, (unsigned) h.j);
        }
    });
    }
}

// Compacts the stored pairs of one replica (replica >= 0) or of all replicas
// (replica < 0) into (key = i << 32 | j, count) arrays: buckets first, then spill.
__global__ void k_contact_compact(const __grid_constant__ Params P, int replica,
                                  unsigned long long *out_keys, unsigned int *out_vals,
                                  unsigned long long *out_count)
{
    GC_PDL_ENTRY();
    const ContactParams &C = P.contact;
    const unsigned long long nrep = replica < 0 ? (unsigned long long) P.n_rep : 1ull;
    const unsigned long long r0 = replica < 0 ? 0ull : (unsigned long long) replica;
    const unsigned long long per_rep = (unsigned long long) P.n * C.bucket_slots;
    const unsigned long long stride = (unsigned long long) gridDim.x * blockDim.x;
    const unsigned long long tid = blockIdx.x * (unsigned long long) blockDim.x + threadIdx.x;
    for (unsigned long long s = tid; s < nrep * per_rep; s += stride) {
        const unsigned long long w = C.bucket[r0 * per_rep + s];
        if (w == GC_EMPTY_KEY) continue;
        const unsigned long long i = (s % per_rep) / C.bucket_slots;
        const unsigned long long o = atomicAdd(out_count, 1ull);
        out_keys[o] = (i << 32) | (w >> 32);
        out_vals[o] = (unsigned int) w;
    }
    for (unsigned long long s = tid; s < nrep * C.capacity; s += stride) {
        const unsigned long long k = C.keys[r0 * C.capacity + s];
        if (k == GC_EMPTY_KEY) continue;
        const unsigned long long o = atomicAdd(out_count, 1ull);
        out_keys[o] = k;
        out_vals[o] = C.vals[r0 * C.capacity + s];
    }
}

// ---------------------------------------------------------------------------
// potential energy (md::system::compute_potential_energy, MD/system.hpp:249-252)
// every interaction is counted by exactly one thread; per-block partial sums
// in FP64 are added on the host in block order (deterministic).
// ---------------------------------------------------------------------------
__global__ void __launch_bounds__(BS) k_energy(const __grid_constant__ Params P, int r, unsigned mask_arg,
                                               double *partials)
{
    GC_PDL_ENTRY();
    RepState &st = P.st[r];
    const StepConst &sc = st.sc;
    const unsigned mask = mask_arg & P.enabled;
    const int t = blockIdx.x * BS + threadIdx.x;
    double E = 0.0;

    if (t < P.n) {
        const size_t ro = (size_t) r * P.n;
        const int par = st.parity;
        const float4 *__restrict__ sp = P.spos[par] + ro;
        const float4 *__restrict__ pos = P.pos[par] + ro;
        const float4 self4 = sp[t];
        const f3 self = make_f3(self4);
        const int idx = __float_as_int(self4.w);
        const float2 abi = P.sab[ro + t];
        const int flags = P.flags[idx];

        if (P.pair.mode != PAIR_NONE && (mask >> P.pair.term & 1u)) {
            PairCoef pc = {sc.kA2, sc.kB2, sc.cA, sc.cB, sc.eA, sc.eB, sc.rc2};
            const bool plain = P.pair.mode == PAIR_PLAIN;
            const float2 *__restrict__ sab = P.sab + ro;
            const int *__restrict__ start = P.cell_start + (size_t) r * (P.ncell_max + 1);
            const int dx = st.dims[0], dy = st.dims[1], dz = st.dims[2];
            const int M = P.cells_per_cut;
            int c = P.scell[ro + t];
            int cx = c % dx, cy = (c / dx) % dy, cz = c / (dx * dy);
            int x0 = max(cx - M, 0), x1 = min(cx + M, dx - 1);
            for (int z = max(cz - M, 0); z <= min(cz + M, dz - 1); z++) {
                for (int y = max(cy - M, 0); y <= min(cy + M, dy - 1); y++) {
                    int row = dx * (y + dy * z);
                    int lo = max(start[row + x0], t + 1), hi = start[row + x1 + 1];
                    for (int j = lo; j < hi; j++) {   // slots above t only: each pair once
                        float4 q = __ldg(&sp[j]);
                        f3 d = {self.x - q.x, self.y - q.y, self.z - q.z};
                        float r2 = dot(d, d);
                        if (r2 < pc.rc2) {
                            float a = 1.0f, b = 0.0f;
                            if (!plain) {
                                float2 abj = __ldg(&sab[j]);
                                a = 0.5f * (abi.x + abj.x);
                                b = 0.5f * (abi.y + abj.y);
                            }
                            E += (double) pair_energy(pc, a, b, r2);
                        }
                    }
                }
            }
        }
        for (int e = P.bond_off[idx]; e < P.bond_off[idx + 1]; e++) {
            BondEntry be = P.bond_ent[e];
            const BondTerm &bt = P.bonds[be.bterm];
            if (!(mask >> bt.term & 1u) || be.partner < idx) continue;
            if (be.partner == idx) continue;
            f3 d = self - make_f3(pos[be.partner]);
            float K = sc.bK_a[be.bterm], l = sc.bl_a[be.bterm];
            if (bt.mix_ab) {
                float2 abj = P.ab[be.partner];
                float a = 0.5f * (abi.x + abj.x), b = 0.5f * (abi.y + abj.y);
                K = a * sc.bK_a[be.bterm] + b * sc.bK_b[be.bterm];
                l = a * sc.bl_a[be.bterm] + b * sc.bl_b[be.bterm];
            }
            E += (double) bond_energy(bt.kind, K, l, dot(d, d));
        }
        if (flags & 1) {
            for (int e = P.triple_off[idx]; e < P.triple_off[idx + 1]; e++) {
                TripleEntry te = P.triple_ent[e];
                int tt = te.role_term >> 2, role = te.role_term & 3;
                if (role != 1 || !(mask >> P.triples[tt].term & 1u)) continue;
                f3 pi = make_f3(pos[te.i]), pj = make_f3(pos[te.j]), pk = make_f3(pos[te.k]);
                E += (double) bending_energy(sc.tE[tt], pi - pj, pj - pk);
            }
        }
        if (flags & 2) {
            for (int e = P.tether_off[idx]; e < P.tether_off[idx + 1]; e++) {
                TetherEntry te = P.tether_ent[e];
                if (!(mask >> te.term & 1u)) continue;
                f3 d = {self.x - sc.pole[te.pole][0], self.y - sc.pole[te.pole][1],
                        self.z - sc.pole[te.pole][2]};
                E += (double) bond_energy(2, te.spring, te.length, dot(d, d));
            }
        }
        if ((flags & 4) && P.well.term >= 0 && (mask >> P.well.term & 1u)) {
            for (int k = 0; k < P.well.n_targets; k++) {
                int j = P.well.targets[k];
                if (j <= idx) continue;
                f3 d = self - make_f3(pos[j]);
                E += (double) well_energy(sc.well_eps, sc.well_k2, sc.well_cut2, dot(d, d));
            }
        }
        for (int k = 0; k < P.n_sources; k++) {
            if (!(mask >> P.sources[k].term & 1u)) continue;
            f3 d = {self.x - sc.src_pt[k][0], self.y - sc.src_pt[k][1], self.z - sc.src_pt[k][2]};
            float r2 = dot(d, d);
            E += (double) (P.sources[k].kind == 0 ? bond_energy(0, sc.src_p0[k], sc.src_p1[k], r2)
                                                  : flux_energy(sc.src_p0[k], sc.src_p1[k], r2));
        }
        if (P.wall.term >= 0 && (mask >> P.wall.term & 1u)) {
            double wx = self.x, wy = self.y, wz = self.z;
            double ux = sc.qd[0] * wx, uy = sc.qd[1] * wy, uz = sc.qd[2] * wz;
            double ddd = ux * ux + uy * uy + uz * uz;
            if (ddd != 0.0) {
                double phid = ux * wx + uy * wy + uz * wz - 1.0;
                double kd = phid / (2.0 * ddd);
                float phi = (float) phid;
                f3 delta = {(float) (kd * ux), (float) (kd * uy), (float) (kd * uz)};
                float d2 = dot(delta, delta);
                if (phi < 0.0f) {
                    PairCoef wc = {sc.w_kA2, sc.w_kB2, sc.w_cA, sc.w_cB, sc.w_eA, sc.w_eB, 0.0f};
                    float a = 0.5f * (abi.x + sc.wall_a), b = 0.5f * (abi.y + sc.wall_b);
                    E += (double) pair_energy(wc, a, b, d2);
                } else {
                    E += (double) (0.5f * sc.wall_K * d2);
                }
            }
        }
    }

    __shared__ double sh[BS / 32];
    for (int o = 16; o > 0; o >>= 1) E += __shfl_xor_sync(0xffffffffu, E, o);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = E;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0.0;
        for (int w = 0; w < BS / 32; w++) s += sh[w];
        partials[blockIdx.x] = s;
    }
}

__global__ void k_debug_noise(const __grid_constant__ Params P, int r, long long step, float *out)
{
    GC_PDL_ENTRY();
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P.n) return;
    f3 z = normal3(P.st[r].seed, (uint32_t) i, step, (uint32_t) P.st[r].epoch);
    out[3 * i] = z.x;
    out[3 * i + 1] = z.y;
    out[3 * i + 2] = z.z;
}

// Squared radius of gyration of bead ranges [start, end): one block per range.
__global__ void k_rg2(const __grid_constant__ Params P, int r, int n_ranges, const int *ranges, double *out)
{
    GC_PDL_ENTRY();
    int g = blockIdx.x;
    if (g >= n_ranges) return;
    const float4 *pos = P.pos[P.st[r].parity] + (size_t) r * P.n;
    int s = ranges[2 * g], e = ranges[2 * g + 1];
    double sx = 0, sy = 0, sz = 0, s2 = 0;
    for (int i = s + threadIdx.x; i < e; i += blockDim.x) {
        float4 p = pos[i];
        sx += p.x; sy += p.y; sz += p.z;
        s2 += (double) p.x * p.x + (double) p.y * p.y + (double) p.z * p.z;
    }
    __shared__ double sh[4][32];
    double v[4] = {sx, sy, sz, s2};
    for (int k = 0; k < 4; k++) {
        for (int o = 16; o > 0; o >>= 1) v[k] += __shfl_xor_sync(0xffffffffu, v[k], o);
        if ((threadIdx.x & 31) == 0) sh[k][threadIdx.x >> 5] = v[k];
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        double a[4] = {0, 0, 0, 0};
        for (int k = 0; k < 4; k++)
            for (int w = 0; w < (int) (blockDim.x + 31) / 32; w++) a[k] += sh[k][w];
        double m = (double) (e - s);
        out[g] = m > 0 ? a[3] / m - (a[0] * a[0] + a[1] * a[1] + a[2] * a[2]) / (m * m) : 0.0;
    }
}

// Trajectory frame as the driver stores it (simulation_store::save_positions,
// SIM/common/simulation_store.cpp:22-32,192-216): every coordinate rounded to
// `bits` significant binary digits (frexp / nearbyint / ldexp) and packed as f32 x 3.
// Scaling by powers of two is exact and the scaled mantissa is < 2^(bits + 1), so the
// FP32 arithmetic below returns the reference's value for FP32 inputs.
__device__ __forceinline__ float quantize_coord(float v, int bits)
{
    int e;
    const float m = frexpf(v, &e);
    return ldexpf(rintf(ldexpf(m, bits)), e - bits);
}

__global__ void __launch_bounds__(256) k_frame(const __grid_constant__ Params P, int r0, int nrep, int bits,
                                               float *__restrict__ out)
{
    GC_PDL_ENTRY();
    const size_t total = (size_t) nrep * P.n;
    for (size_t t = (size_t) blockIdx.x * 256 + threadIdx.x; t < total; t += (size_t) gridDim.x * 256) {
        const int r = r0 + (int) (t / P.n);
        const float4 p = P.pos[P.st[r].parity][(size_t) r * P.n + (t % P.n)];
        out[3 * t] = quantize_coord(p.x, bits);
        out[3 * t + 1] = quantize_coord(p.y, bits);
        out[3 * t + 2] = quantize_coord(p.z, bits);
    }
}

// Host transfers of whole ensembles (gc_set/get_positions_all_f32): the packed
// [R][n][3] f32 image of view_positions() is repacked to / from the float4 device
// layout here, so that each direction is ONE asynchronous copy.
__global__ void __launch_bounds__(256) k_unpack_positions(const __grid_constant__ Params P,
                                                          const float *__restrict__ in)
{
    GC_PDL_ENTRY();
    const size_t total = (size_t) P.n_rep * P.n;
    for (size_t t = (size_t) blockIdx.x * 256 + threadIdx.x; t < total; t += (size_t) gridDim.x * 256) {
        const int r = (int) (t / P.n);
        P.pos[P.st[r].parity][t] = make_float4(in[3 * t], in[3 * t + 1], in[3 * t + 2], 0.0f);
    }
}

__global__ void __launch_bounds__(256) k_pack_positions(const __grid_constant__ Params P,
                                                        float *__restrict__ out)
{
    GC_PDL_ENTRY();
    const size_t total = (size_t) P.n_rep * P.n;
    for (size_t t = (size_t) blockIdx.x * 256 + threadIdx.x; t < total; t += (size_t) gridDim.x * 256) {
        const int r = (int) (t / P.n);
        const float4 p = P.pos[P.st[r].parity][t];
        out[3 * t] = p.x;
        out[3 * t + 1] = p.y;
        out[3 * t + 2] = p.z;
    }
}

} // namespace gc

// gc_step.cuh -- single-bead force terms, BAOAB-limit Brownian update, per-step
// reductions and the on-device body of the driver's per-step callback.
//
// One thread owns one bead (in cell-sorted order) and GATHERS every term that
// acts on it (instead of the reference's half-list scatter,
// MD/forcefield/bonded_pairwise_forcefield.hpp:78-93), so there are no force
// atomics and the result is bit-reproducible run to run.
#pragma once

#include "gc_kernels.cuh"

namespace gc {

enum StepMode : int {
    STEP_FUSED = 0,   // force + update + per-step callback body (fixed dt)
    STEP_FORCE = 1    // force only -> P.force (adaptive stepper, gc_compute_force)
};

// ---------------------------------------------------------------------------
// FP64 context -> FP32 step constants (one thread, once per step per replica)
// ---------------------------------------------------------------------------
// The constants are independent of one another, so the finalize kernels spread
// the items over the lanes of a warp (FP64 divisions are slow in one thread).
__device__ inline void refresh_consts_item(const Params &P, RepState &st, int item)
{
    StepConst &c = st.sc;
    if (item == 0) {
        double s = P.pair.scale_by_core ? st.core_scale : 1.0;
        PairCoef pc = make_pair_coef(P.pair.eps_a, P.pair.sigma_a, P.pair.eps_b, P.pair.sigma_b, s,
                                     P.pair.mode == PAIR_NONE ? PAIR_PLAIN : P.pair.mode);
        c.kA2 = pc.kA2; c.kB2 = pc.kB2; c.cA = pc.cA; c.cB = pc.cB;
        c.eA = pc.eA; c.eB = pc.eB; c.rc2 = pc.rc2;
    } else if (item == 1) {
        if (P.wall.term >= 0) {
            double s = P.wall.scale_by_core ? st.core_scale : 1.0;
            // inward soft cores use HALF diameters
            // (SIM/stage_interphase/simulation_driver_forcefield.cpp:214-221)
            PairCoef wc = make_pair_coef(P.wall.eps_a, P.wall.sigma_a * 0.5, P.wall.eps_b,
                                         P.wall.sigma_b * 0.5, s, PAIR_AB);
            c.w_kA2 = wc.kA2; c.w_kB2 = wc.kB2; c.w_cA = wc.cA; c.w_cB = wc.cB;
            c.w_eA = wc.eA; c.w_eB = wc.eB;
            c.wall_a = (float) P.wall.wall_a;
            c.wall_b = (float) P.wall.wall_b;
            c.wall_K = (float) P.wall.outward_spring;
        }
    } else if (item < 5) {
        if (P.wall.term >= 0) {
            const int k = item - 2;
            c.qd[k] = 1.0 / (st.semiaxes[k] * st.semiaxes[k]);
            c.q[k] = (float) c.qd[k];
        }
    } else if (item == 5) {
        if (P.well.term >= 0) {
            c.well_eps = (float) P.well.eps;
            c.well_k2 = (float) (1.0 / (P.well.decay * P.well.decay));
            c.well_cut2 = (float) (P.well.cutoff * P.well.cutoff);
        }
    } else if (item >= 8 && item < 8 + GC_MAX_BOND_TERMS) {
        const int b = item - 8;
        if (b < P.n_bond_terms) {
            const BondTerm &bt = P.bonds[b];
            double s = bt.scale ? st.bond_scale : 1.0;
            double is2 = 1.0 / (s * s);
            c.bK_a[b] = (float) (bt.K_a * is2);
            c.bK_b[b] = (float) (bt.K_b * is2);
            c.bl_a[b] = (float) (bt.l_a * s);
            c.bl_b[b] = (float) (bt.l_b * s);
        }
    } else if (item >= 16 && item < 16 + GC_MAX_TRIPLE_TERMS) {
        const int t = item - 16;
        if (t < P.n_triple_terms) c.tE[t] = (float) P.triples[t].energy;
    } else if (item >= 20 && item < 20 + GC_MAX_SOURCES) {
        const int k = item - 20;
        if (k < P.n_sources) {
            for (int d = 0; d < 3; d++) c.src_pt[k][d] = (float) P.sources[k].point[d];
            c.src_p0[k] = (float) P.sources[k].p0;
            c.src_p1[k] = (float) P.sources[k].p1;
        }
    } else if (item >= 24 && item < 24 + GC_MAX_POLES) {
        const int k = item - 24;
        if (k < P.n_poles) {
            for (int d = 0; d < 3; d++) c.pole[k][d] = (float) P.poles[k][d];
        }
    }
}
static_assert(8 + GC_MAX_BOND_TERMS <= 16 && 16 + GC_MAX_TRIPLE_TERMS <= 20 &&
              20 + GC_MAX_SOURCES <= 24 && 24 + GC_MAX_POLES <= 32, "item layout of refresh_consts");

__device__ inline void refresh_consts(const Params &P, RepState &st)
{
    for (int item = 0; item < 32; item++) refresh_consts_item(P, st, item);
}

// ---------------------------------------------------------------------------
// per-bead force: all enabled terms acting on the bead in sorted slot t
// ---------------------------------------------------------------------------
struct WallAcc {
    float ax, ay, az, reaction;
    int stale;            // decomposed run: a bonded partner was not current on this rank
};

// Single-bead terms only: the nonbonded soft cores come from k_pair_cells.
__device__ __forceinline__ f3 bead_force(const Params &P, const RepState &st, int r, int t,
                                         int idx, float4 self4, float2 abi, unsigned mask,
                                         WallAcc &wacc)
{
    const StepConst &sc = st.sc;
    const size_t ro = (size_t) r * P.n;
    const int par = st.parity;
    const f3 self = make_f3(self4);
    f3 F = {0.0f, 0.0f, 0.0f};

    const float4 *__restrict__ pos = P.pos[par] + ro;   // original order
    const int flags = P.flags[idx];

    // ---- bonded pairs (MD/forcefield/bonded_pairwise_forcefield.hpp:78-93) -
    {
        int e0 = P.bond_off[idx], e1 = P.bond_off[idx + 1];
        for (int e = e0; e < e1; e++) {
            BondEntry be = P.bond_ent[e];
            const BondTerm &bt = P.bonds[be.bterm];
            if (!(mask >> bt.term & 1u)) continue;
            float4 q = pos[be.partner];
            // decomposed run: the partner's position must be current here, i.e. lie within
            // the position halo (Params::dhalo layers) of this rank's slab.  (Stamps of beads
            // this kernel has already moved read step + 1: current is >= step.)
            if (P.decomp && P.dstamp[be.partner] < (int) st.step) wacc.stale = 1;
            f3 d = {self.x - q.x, self.y - q.y, self.z - q.z};
            float K = sc.bK_a[be.bterm], l = sc.bl_a[be.bterm];
            if (bt.mix_ab) {
                float2 abj = P.ab[be.partner];
                float a = 0.5f * (abi.x + abj.x), b = 0.5f * (abi.y + abj.y);
                K = a * sc.bK_a[be.bterm] + b * sc.bK_b[be.bterm];
                l = a * sc.bl_a[be.bterm] + b * sc.bl_b[be.bterm];
            }
            float fac = bond_force_factor(bt.kind, K, l, dot(d, d));
            F.x = fmaf(fac, d.x, F.x);
            F.y = fmaf(fac, d.y, F.y);
            F.z = fmaf(fac, d.z, F.z);
        }
    }

    // ---- bending triples (MD/forcefield/bonded_triplewise_forcefield.hpp) --
    if (flags & 1) {
        int e0 = P.triple_off[idx], e1 = P.triple_off[idx + 1];
        for (int e = e0; e < e1; e++) {
            TripleEntry te = P.triple_ent[e];
            int tt = te.role_term >> 2, role = te.role_term & 3;
            if (!(mask >> P.triples[tt].term & 1u)) continue;
            if (P.decomp && (P.dstamp[te.i] < (int) st.step || P.dstamp[te.j] < (int) st.step ||
                             P.dstamp[te.k] < (int) st.step)) wacc.stale = 1;
            f3 pi = make_f3(pos[te.i]), pj = make_f3(pos[te.j]), pk = make_f3(pos[te.k]);
            f3 f = bending_force(sc.tE[tt], pi - pj, pj - pk, role);
            F = F + f;
        }
    }

    // ---- kinetochore fibres (SIM/common/forcefield/kinetochore_fiber_forcefield.cpp:70-84)
    if (flags & 2) {
        int e0 = P.tether_off[idx], e1 = P.tether_off[idx + 1];
        for (int e = e0; e < e1; e++) {
            TetherEntry te = P.tether_ent[e];
            if (!(mask >> te.term & 1u)) continue;
            f3 d = {self.x - sc.pole[te.pole][0], self.y - sc.pole[te.pole][1],
                    self.z - sc.pole[te.pole][2]};
            float fac = bond_force_factor(2 /* GC_BOND_SPRING */, te.spring, te.length, dot(d, d));
            F = F + fac * d;
        }
    }

    // ---- nucleolar droplet: brute force over the (few) target beads --------
    if ((flags & 4) && P.well.term >= 0 && (mask >> P.well.term & 1u)) {
        for (int k = 0; k < P.well.n_targets; k++) {
            int j = P.well.targets[k];
            if (j == idx) continue;
            if (P.decomp && P.dstamp[j] < (int) st.step) {
                // not current here: further than the position halo from this rank's slab,
                // where the well must have ended (its term is exactly 0 beyond the cutoff)
                const float reach = (float) P.dhalo * st.cell_size;
                if (sc.well_cut2 > reach * reach) wacc.stale = 1;
                continue;
            }
            float4 q = pos[j];
            f3 d = {self.x - q.x, self.y - q.y, self.z - q.z};
            float fac = well_force_factor(sc.well_eps, sc.well_k2, sc.well_cut2, dot(d, d));
            F = F + fac * d;
        }
    }

    // ---- point sources (MD/forcefield/point_source_forcefield.hpp:88-110) --
    for (int k = 0; k < P.n_sources; k++) {
        if (!(mask >> P.sources[k].term & 1u)) continue;
        f3 d = {self.x - sc.src_pt[k][0], self.y - sc.src_pt[k][1], self.z - sc.src_pt[k][2]};
        float r2 = dot(d, d);
        float fac = P.sources[k].kind == 0
                        ? bond_force_factor(0, sc.src_p0[k], sc.src_p1[k], r2)
                        : flux_force_factor(sc.src_p0[k], sc.src_p1[k], r2);
        F = F + fac * d;
    }

    // ---- ellipsoidal wall, inward + outward fused --------------------------
    // MD/forcefield/ellipsoid_surface_forcefield.hpp:55-211.  The potential is
    // central in delta, so iso = f and aniso = 0 and the reference's
    // iso + (aniso - iso) o strain reduces to f o (1 - strain).
    if (P.wall.term >= 0 && (mask >> P.wall.term & 1u)) {
        // The implicit function phi = sum q x^2 - 1 cancels to ~1e-2 near the wall
        // and the steep softcore<8,3> amplifies its error, so the point
        // evaluation (evaluate_point, lines 55-90) is done in FP64: ~20 DFMA per
        // bead-step against several hundred FP32 pair candidates.
        double wx = self.x, wy = self.y, wz = self.z;
        double ux = sc.qd[0] * wx, uy = sc.qd[1] * wy, uz = sc.qd[2] * wz;
        double ddd = ux * ux + uy * uy + uz * uz;
        if (ddd != 0.0) {
            double phid = ux * wx + uy * wy + uz * wz - 1.0;
            double kd = phid / (2.0 * ddd);
            float phi = (float) phid;
            float kappa = (float) kd;
            f3 delta = {(float) (kd * ux), (float) (kd * uy), (float) (kd * uz)};
            float d2 = dot(delta, delta);
            float fac;
            if (phi < 0.0f) {
                PairCoef wc = {sc.w_kA2, sc.w_kB2, sc.w_cA, sc.w_cB, sc.w_eA, sc.w_eB, 0.0f};
                float a = 0.5f * (abi.x + sc.wall_a), b = 0.5f * (abi.y + sc.wall_b);
                fac = pair_force_factor(wc, a, b, d2);
            } else {
                fac = -sc.wall_K;
            }
            f3 f = {fac * delta.x * (1.0f - kappa * sc.q[0]),
                    fac * delta.y * (1.0f - kappa * sc.q[1]),
                    fac * delta.z * (1.0f - kappa * sc.q[2])};
            F = F + f;
            // statistics (ellipsoid_surface_forcefield.hpp:199-209): the outward
            // normal has the sign of the bead coordinate on every axis.
            wacc.ax -= self.x > 0.0f ? f.x : -f.x;
            wacc.ay -= self.y > 0.0f ? f.y : -f.y;
            wacc.az -= self.z > 0.0f ? f.z : -f.z;
            if (d2 > 0.0f) {
                float sgn = phi < 0.0f ? -1.0f : 1.0f;
                wacc.reaction -= sgn * dot(f, delta) * rsqrtf(d2);
            }
        }
    }
    return F;
}

// ---------------------------------------------------------------------------
// block reductions
// ---------------------------------------------------------------------------
__device__ __forceinline__ float warp_max(float v)
{
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ float warp_min(float v)
{
    for (int o = 16; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ long long warp_sum(long long v)
{
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

struct StepRed {          // per-thread contributions to the per-step reductions
    float lo[3], hi[3];   // bounding box of the new positions
    float disp2;          // squared displacement from the rebuild reference
    float dtmin;          // adaptive time-step candidate
    WallAcc w;
    bool bad;
};

__device__ __forceinline__ void red_init(StepRed &s)
{
    for (int k = 0; k < 3; k++) { s.lo[k] = INFINITY; s.hi[k] = -INFINITY; }
    s.disp2 = 0.0f;
    s.dtmin = INFINITY;
    s.w = {0.0f, 0.0f, 0.0f, 0.0f, 0};
    s.bad = false;
}

// Reduce over the block and push one set of atomics per block.  The per-step
// scalars are consumed by k_finalize_*, a separate one-warp launch: the kernel
// boundary orders it after every block, so no fence/ticket is needed here.
template<int NT>
__device__ __forceinline__ void block_reduce_commit(RepState &st, StepRed &s, bool with_wall,
                                                    bool with_geom, bool with_dt)
{
    constexpr int NW = NT / 32;
    __shared__ int sh_i[NW][8];
    __shared__ long long sh_l[NW][4];
    __shared__ int sh_bad, sh_wall;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (threadIdx.x == 0) { sh_bad = 0; sh_wall = 0; }

    // warp level: one REDUX per value on order-preserving integer images of the floats
    // (disp2 >= 0 and dt > 0 are ordered as unsigned bit patterns already)
    int v[8] = {0, 0, 0, 0, 0, 0, 0, 0x7f800000};
    if (with_geom) {
        for (int k = 0; k < 3; k++) {
            v[k] = __reduce_min_sync(0xffffffffu, float_to_ordered(s.lo[k]));
            v[3 + k] = __reduce_max_sync(0xffffffffu, float_to_ordered(s.hi[k]));
        }
        v[6] = (int) __reduce_max_sync(0xffffffffu, __float_as_uint(s.disp2 != s.disp2 ? NAN : s.disp2));
    }
    if (with_dt) v[7] = (int) __reduce_min_sync(0xffffffffu, __float_as_uint(s.dtmin));
    // the wall touches few beads: warps without a contribution skip the 64-bit sums
    long long q[4] = {0, 0, 0, 0};
    const bool wall_any = with_wall && __any_sync(0xffffffffu, s.w.ax != 0.0f || s.w.ay != 0.0f ||
                                                                  s.w.az != 0.0f || s.w.reaction != 0.0f);
    if (wall_any) {
        q[0] = warp_sum(__double2ll_rn((double) s.w.ax * GC_FX_SCALE));
        q[1] = warp_sum(__double2ll_rn((double) s.w.ay * GC_FX_SCALE));
        q[2] = warp_sum(__double2ll_rn((double) s.w.az * GC_FX_SCALE));
        q[3] = warp_sum(__double2ll_rn((double) s.w.reaction * GC_FX_SCALE));
    }
    bool bad = __any_sync(0xffffffffu, s.bad);
    if (__any_sync(0xffffffffu, s.w.stale != 0) && lane == 0) atomicExch(&st.error, 34);
    __syncthreads();
    if (lane == 0) {
        for (int k = 0; k < 8; k++) sh_i[wid][k] = v[k];
        for (int k = 0; k < 4; k++) sh_l[wid][k] = q[k];
        if (bad) sh_bad = 1;
        if (wall_any) sh_wall = 1;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < NW; w++) {
            for (int k = 0; k < 3; k++) {
                v[k] = min(v[k], sh_i[w][k]);
                v[3 + k] = max(v[3 + k], sh_i[w][3 + k]);
            }
            v[6] = (int) max((unsigned) v[6], (unsigned) sh_i[w][6]);
            v[7] = (int) min((unsigned) v[7], (unsigned) sh_i[w][7]);
            for (int k = 0; k < 4; k++) q[k] += sh_l[w][k];
        }
        if (with_geom) {
            for (int k = 0; k < 3; k++) {
                if (v[k] <= v[3 + k]) {                     // a block of idle threads holds +-inf
                    atomicMin(&st.bbox_min[k], v[k]);
                    atomicMax(&st.bbox_max[k], v[3 + k]);
                }
            }
            // NaN displacement: bits compare above every finite value
            atomicMax(&st.maxdisp2_bits, (unsigned) v[6]);
        }
        if (with_dt) atomicMin(&st.dtmin_bits, (unsigned) v[7]);
        if (with_wall && sh_wall) {
            for (int k = 0; k < 3; k++) {
                if (q[k]) atomicAdd((unsigned long long *) &st.axial_fx[k], (unsigned long long) q[k]);
            }
            if (q[3]) atomicAdd((unsigned long long *) &st.reaction_fx, (unsigned long long) q[3]);
        }
        if (sh_bad) atomicExch(&st.error, 1);
    }
}

// Wall statistics of the force pass just finished: fixed point -> double.
__device__ inline void finalize_stats(RepState &st)
{
    volatile long long *fx = st.axial_fx;
    for (int k = 0; k < 3; k++) {
        st.axial[k] = (double) fx[k] / GC_FX_SCALE;
        st.axial_fx[k] = 0;
    }
    st.reaction = (double) *(volatile long long *) &st.reaction_fx / GC_FX_SCALE;
    st.reaction_fx = 0;
}

// The body of the driver's per-step callback plus list maintenance, executed by
// one thread of the last block after every bead has been updated:
//   SIM/stage_interphase/simulation_driver_interphase.cpp:16-18 (time),
//   67-76 (core/bond scale, contact distance), 79-90 (wall ODE)
//   MD/forcefield/detail/neighbor_list.hpp:101-159 (consistency check)
__device__ inline void finalize_step(const Params &P, RepState &st, double dt_used)
{
    st.dt_prev = st.dt_last;
    st.dt_last = dt_used;
    st.step += 1;
    st.time = (double) st.step * P.dt_cfg;

    st.contact_distance_prev = st.contact_distance;
    if (P.sched.enabled) {
        st.core_scale = 1.0 - (1.0 - P.sched.core_init) * exp(-st.time / P.sched.core_tau);
        st.bond_scale = 1.0 - (1.0 - P.sched.bond_init) * exp(-st.time / P.sched.bond_tau);
    }
    if (P.contact.enabled && P.contact.scale_by_core) {
        st.contact_distance = P.contact.base_distance * st.core_scale;
    }
    if (P.wall.term >= 0 && P.wall.ode_enabled) {
        for (int k = 0; k < 3; k++) {
            double net = st.axial[k] - P.wall.ode_spring[k] * st.semiaxes[k];
            st.semiaxes[k] += P.dt_cfg * P.wall.ode_mobility * net;
        }
    }
    st.parity ^= 1;

    // list maintenance for the NEXT force evaluation
    float maxdisp2 = __uint_as_float(*(volatile unsigned int *) &st.maxdisp2_bits);
    if (maxdisp2 != maxdisp2) st.error = 1;
    double rnow = interaction_radius(P, st);
    bool need = P.skin <= 0.0f || !(rnow + 2.0 * sqrt((double) maxdisp2) <= (double) st.rc_valid);
    if (need) {
        // bbox was accumulated from the positions just written
        int bmin[3], bmax[3];
        for (int k = 0; k < 3; k++) {
            bmin[k] = *(volatile int *) &st.bbox_min[k];
            bmax[k] = *(volatile int *) &st.bbox_max[k];
        }
        for (int k = 0; k < 3; k++) { st.bbox_min[k] = bmin[k]; st.bbox_max[k] = bmax[k]; }
        plan_grid(P, st, rnow * (1.0 + (double) P.skin));
        st.need_rebuild = 1;
        st.maxdisp2_bits = 0u;
        st.rebuilds += 1;
    } else {
        st.need_rebuild = 0;
    }
    for (int k = 0; k < 3; k++) {
        st.bbox_min[k] = 0x7fffffff;
        st.bbox_max[k] = (int) 0x80000000;
    }
    st.dtmin_bits = 0x7f800000u;
    st.cell_cursor = 0;
    st.contact_cursor = 0;
}

// ---------------------------------------------------------------------------
// BAOAB-limit update of one bead
//   MD/simulation/detail/brownian_simulator.hpp:61-75:
//     x += mu dt F;  x += (W_n + W_{n-1}) / 2;  W = sqrt(2 kT mu dt) N(0,1)^3
//   The two increments are summed first so x is rounded once in FP32.
// ---------------------------------------------------------------------------
__device__ __forceinline__ f3 bead_noise(const Params &P, const RepState &st, int r, int idx,
                                         long long n)
{
    if (P.noise) {
        const float *z = P.noise + (((size_t) r * P.noise_steps + (size_t) (n - 1 - P.noise_step0)) *
                                        P.n + idx) * 3;
        return {z[0], z[1], z[2]};
    }
    return normal3(st.seed, (uint32_t) idx, n, (uint32_t) st.epoch);
}

__device__ __forceinline__ void integrate_bead(const Params &P, const RepState &st, int r, int t,
                                               int idx, float4 self4, f3 F, float dt, StepRed &red)
{
    const size_t ro = (size_t) r * P.n;
    const int par = st.parity;
    const long long n = st.step + 1;
    float mu = P.mob[idx];
    float mudt = dt * mu;
    float sigma = sqrtf(2.0f * P.kT * mudt);
    f3 W = sigma * bead_noise(P, st, r, idx, n);
    // W_{n-1} is REGENERATED from the counter-based stream instead of being kept
    // in memory as the reference does (brownian_simulator.hpp:72-74, 83): a stored
    // copy costs a scattered 16-byte read-modify-write per bead, measured at 2.6x
    // the whole kernel's time (profiles/r1_notes.md)
    f3 Wp = {0.0f, 0.0f, 0.0f};
    if (n > 1) {
        float sigma_p = sqrtf(2.0f * P.kT * (float) st.dt_last * mu);
        Wp = sigma_p * bead_noise(P, st, r, idx, n - 1);
    }
    float4 x;
    x.x = self4.x + fmaf(mudt, F.x, 0.5f * (W.x + Wp.x));
    x.y = self4.y + fmaf(mudt, F.y, 0.5f * (W.y + Wp.y));
    x.z = self4.z + fmaf(mudt, F.z, 0.5f * (W.z + Wp.z));
    x.w = 0.0f;
    P.pos[par ^ 1][ro + idx] = x;
    if (P.decomp) {
        P.dstamp[idx] = (int) n;                    // current on this rank for step n
        // ... and in slot order with the bead index, for k_decomp_pack and the next cell-list
        // build (decomp_current): neither then gathers from pos[]
        P.spos[par ^ 1][t] = make_float4(x.x, x.y, x.z, self4.w);
    }
    if (P.skin > 0.0f) {
        // cells may be reused: keep the sorted copy current and track the largest
        // displacement since the rebuild (neighbor_list.hpp:129-158)
        x.w = self4.w;
        P.spos[par ^ 1][ro + t] = x;
        {   // ... and the pair-interleaved copy the pair kernel stages from
            const int dxy = st.dims[0] * st.dims[1];
            const int L = P.scell[ro + t] / dxy;
            const int us = unit_slot(t, L, P.cell_start[(size_t) r * (P.ncell_max + 1) + L * dxy]);
            float *u0 = unit_half(P, r, us, 0);
            u0[0] = x.x;
            u0[2] = x.y;
            unit_half(P, r, us, 1)[0] = x.z;
        }
        float4 ref = P.sref[ro + t];
        float ex = x.x - ref.x, ey = x.y - ref.y, ez = x.z - ref.z;
        red.disp2 = fmaf(ex, ex, fmaf(ey, ey, ez * ez));
    }   // skin == 0: every step re-bins from pos, nothing else reads these
    red.lo[0] = red.hi[0] = x.x;
    red.lo[1] = red.hi[1] = x.y;
    red.lo[2] = red.hi[2] = x.z;
    red.bad = !(fabsf(x.x) < INFINITY && fabsf(x.y) < INFINITY && fabsf(x.z) < INFINITY);
}

// MD/simulation/detail/brownian_timestepper.hpp:72-91 (root in the
// cancellation-free form 2c / (b + sqrt(b^2 + 4ac)))
__device__ __forceinline__ float solve_timestep(float spacestep, float mu, float f2, float kT,
                                                float dt_max)
{
    if (mu == 0.0f && f2 == 0.0f) return dt_max;
    float a = mu * mu * f2;
    float b = 2.55f * mu * kT;
    float c = spacestep * spacestep;
    if (a * c < 1.0e-6f * b * b) return c / b;
    return 2.0f * c / (b + sqrtf(fmaf(b, b, 4.0f * a * c)));
}

// Second half of an adaptive step: update with the globally reduced dt.
__global__ void __launch_bounds__(BS) k_integrate(const __grid_constant__ Params P)
{
    GC_PDL_ENTRY();
    const int r = blockIdx.y;
    RepState &st = P.st[r];
    const int t = blockIdx.x * BS + threadIdx.x;
    StepRed red;
    red_init(red);
    const float dt = st.sc.dt_cur;

    if (t < P.n) {
        const size_t ro = (size_t) r * P.n;
        float4 self4 = P.spos[st.parity][ro + t];
        int idx = __float_as_int(self4.w);
        float4 F4 = P.force[ro + idx];
        f3 F = {F4.x, F4.y, F4.z};
        integrate_bead(P, st, r, t, idx, self4, F, dt, red);
    }
    block_reduce_commit<BS>(st, red, false, true, false);
}

// ---------------------------------------------------------------------------
// per-step epilogues: one thread per replica, launched right after the kernel
// whose reductions they consume
// ---------------------------------------------------------------------------
// One warp per replica: the RepState is staged in shared memory (coalesced copy
// in, serial FP64 bookkeeping by lane 0, coalesced copy out) instead of ~100
// dependent global round trips.
__device__ __forceinline__ void repstate_load(RepState &dst, const RepState &src, int lane)
{
    // all loads first (one round trip), then the stores
    constexpr int W = (int) (sizeof(RepState) / sizeof(int)), J = (W + 31) / 32;
    const int *s = reinterpret_cast<const int *>(&src);
    int *d = reinterpret_cast<int *>(&dst);
    int tmp[J];
#pragma unroll
    for (int j = 0; j < J; j++) tmp[j] = lane + 32 * j < W ? s[lane + 32 * j] : 0;
#pragma unroll
    for (int j = 0; j < J; j++) {
        if (lane + 32 * j < W) d[lane + 32 * j] = tmp[j];
    }
    __syncwarp();
}
__device__ __forceinline__ void repstate_store(RepState &dst, const RepState &src, int lane)
{
    __syncwarp();
    const int *s = reinterpret_cast<const int *>(&src);
    int *d = reinterpret_cast<int *>(&dst);
    for (int k = lane; k < (int) (sizeof(RepState) / sizeof(int)); k += 32) d[k] = s[k];
}
static_assert(sizeof(RepState) % sizeof(int) == 0, "RepState is copied word by word");

constexpr int DHDR = 16;     // doubles per rank header of a decomposed run (gc_decomp.cuh)

// Decomposed run, after the header all-gather: the reductions of the whole system, folded in
// rank order (integer sums and min/max: order-independent anyway)
__device__ inline void decomp_reduce(const Params &P, RepState &st)
{
    long long sum[4] = {0, 0, 0, 0};
    int bmin[3] = {0x7fffffff, 0x7fffffff, 0x7fffffff};
    int bmax[3] = {(int) 0x80000000, (int) 0x80000000, (int) 0x80000000};
    int err = 0, moved = 0;
    for (int k = 0; k < P.dworld; k++) {
        const double *h = P.dhdr_all + (size_t) k * DHDR;
        const long long *hl = reinterpret_cast<const long long *>(h);
        const int *hi = reinterpret_cast<const int *>(h + 4);
        for (int c = 0; c < 4; c++) sum[c] += hl[c];
        for (int c = 0; c < 3; c++) { bmin[c] = min(bmin[c], hi[c]); bmax[c] = max(bmax[c], hi[3 + c]); }
        if (!err) err = hi[6];
        moved += hi[7];
    }
    for (int c = 0; c < 3; c++) { st.axial_fx[c] = sum[c]; st.bbox_min[c] = bmin[c]; st.bbox_max[c] = bmax[c]; }
    st.reaction_fx = sum[3];
    if (!err && moved != P.n) err = 4;           // a bead was lost or moved twice
    if (err) st.error = err;
    // Load balance: nudge every slab boundary towards equal bead counts (the equal-volume cut
    // assumes a uniformly filled ellipsoid).  Integral control on the cumulative counts, a
    // tenth of the error per step -- far less than a layer, so a boundary moves by at most one
    // layer between two plans, which the halo logic relies on.  Same inputs on every rank.
    if (!err) {
        long long below = 0;
        for (int k = 1; k < P.dworld; k++) {
            const int *hi = reinterpret_cast<const int *>(P.dhdr_all + (size_t) (k - 1) * DHDR + 4);
            below += hi[7];
            const double want = (double) k / (double) P.dworld;
            double f = st.dfrac[k] + 0.1 * (want - (double) below / (double) P.n);
            const double lo = st.dfrac[k - 1] + 1.0e-4;
            f = f < lo ? lo : f;
            st.dfrac[k] = f > 1.0 - 1.0e-4 ? 1.0 - 1.0e-4 : f;
        }
    }
}

// after k_finish<STEP_FUSED> (fixed dt) or k_integrate (adaptive dt)
__global__ void __launch_bounds__(32) k_finalize_step(const __grid_constant__ Params P, int adaptive)
{
    GC_PDL_ENTRY();
    __shared__ RepState s;
    const int r = blockIdx.x, lane = threadIdx.x;
    repstate_load(s, P.st[r], lane);
    if (lane == 0) {
        if (P.decomp) decomp_reduce(P, s);
        if (adaptive) {
            finalize_step(P, s, (double) s.sc.dt_cur);
        } else {
            finalize_stats(s);
            finalize_step(P, s, P.dt_cfg);
        }
    }
    __syncwarp();
    refresh_consts_item(P, s, lane);       // constants of the next step, one item per lane
    repstate_store(P.st[r], s, lane);
}

// Small systems, fixed dt: the epilogue of step n and the cell list of step n + 1 in one
// launch (one SMALL_BS block per replica; warp 0 is k_finalize_step, then the whole block
// re-bins the positions just written).  A mitotic-stage step is then 3 launches
// (k_pair_cells, k_finish, this) instead of 7.
__global__ void __launch_bounds__(SMALL_BS) k_finalize_rebuild(const __grid_constant__ Params P)
{
    GC_PDL_ENTRY();
    __shared__ RepState s;
    __shared__ int sh[32];
    const int r = blockIdx.x;
    if (threadIdx.x < 32) {
        const int lane = threadIdx.x;
        repstate_load(s, P.st[r], lane);
        if (lane == 0) {
            finalize_stats(s);
            finalize_step(P, s, P.dt_cfg);
        }
        __syncwarp();
        refresh_consts_item(P, s, lane);
        repstate_store(P.st[r], s, lane);
    }
    __syncthreads();                       // the new RepState is visible to the whole block
    block_rebuild(P, r, sh);
}

// after k_finish<STEP_FORCE>: wall statistics and the adaptive time step
// (MD/simulation/detail/brownian_timestepper.hpp:109-136: global minimum)
__global__ void __launch_bounds__(32) k_finalize_force(const __grid_constant__ Params P)
{
    GC_PDL_ENTRY();
    __shared__ RepState s;
    const int r = blockIdx.x, lane = threadIdx.x;
    repstate_load(s, P.st[r], lane);
    if (lane == 0) {
        finalize_stats(s);
        float dtc = P.dt;
        if (P.spacestep != 0.0f) dtc = fminf(dtc, __uint_as_float(s.dtmin_bits));
        s.sc.dt_cur = dtc;
        s.dtmin_bits = 0x7f800000u;
        s.cell_cursor = 0;
        s.contact_cursor = 0;
    }
    repstate_store(P.st[r], s, lane);
}

// radius_override > 0 forces the list radius (gc_neighbor_pairs).
__global__ void k_plan(const __grid_constant__ Params P, double radius_override)
{
    GC_PDL_ENTRY();
    int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= P.n_rep) return;
    RepState &st = P.st[r];
    double rl = radius_override > 0.0 ? radius_override
                                      : interaction_radius(P, st) * (1.0 + (double) P.skin);
    plan_grid(P, st, rl);
    st.need_rebuild = 1;
    st.maxdisp2_bits = 0u;
    st.rebuilds += 1;
    st.cell_cursor = 0;
    st.contact_cursor = 0;
    // the next step accumulates its own bounding box from scratch
    for (int k = 0; k < 3; k++) {
        st.bbox_min[k] = 0x7fffffff;
        st.bbox_max[k] = (int) 0x80000000;
    }
    refresh_consts(P, st);
}

} // namespace gc

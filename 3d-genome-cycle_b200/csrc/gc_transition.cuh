// gc_transition.cuh -- inter-stage transitions on the device, so that a
// multi-generation lineage never leaves the GPU.  The reference runs them as a separate
// `transition` program between stage binaries, through the trajectory file:
//   telophase -> relaxation      SIM/stage_transition/transition_interphase.cpp:13-84
//   interphase -> prometaphase   SIM/stage_transition/transition_prometaphase.cpp:24-105
//   prometaphase -> anaphase     SIM/stage_transition/transition_cycle.cpp:25-76
// Positions are read from one context's pos[] (float4, original bead order) and written to
// another's; all arithmetic is FP64, as in the reference.  `bits` > 0 reproduces the file
// hand-off: the input is rounded as the stored last frame would be and the output as
// save_positions rounds the new stage's frame 0 (SIM/common/simulation_store.cpp:22-32).
#pragma once

#include "gc_types.cuh"

namespace gc {

__device__ __forceinline__ double quantize_d(double v, int bits)
{
    if (bits <= 0) return v;
    int e;
    const double m = frexp(v, &e);
    return (double) (float) ldexp(rint(ldexp(m, bits)), e - bits);   // f32 store of the frame
}

// Not-a-knot cubic spline through n uniform knots t_k = (k + 1/2) / n, as
// submodules/github.com/snsinfu/cxx-spline/include/spline.hpp builds it (second derivatives
// from a tridiagonal system eliminated with row pivoting).  One thread per coordinate solves
// its system in shared memory; W = L | D | U | Y, each n doubles; values in Y on entry,
// second derivatives in Y on exit, node values kept in V.
__device__ inline void spline_second_derivatives(int n, double *L, double *D, double *U, double *Y,
                                                 const double *V)
{
    const int ns = n - 1;                        // segments
    const double h = 1.0 / (double) n;           // uniform interval (the knots are (k + 1/2) / n)
    for (int i = 0; i <= ns; i++) { L[i] = 0.0; D[i] = 0.0; U[i] = 0.0; Y[i] = 0.0; }
    auto slope = [&](int i) { return (V[i + 1] - V[i]) / (((0.5 + (double) (i + 1)) / (double) n) -
                                                          ((0.5 + (double) i) / (double) n)); };
    auto interval = [&](int i) { return ((0.5 + (double) (i + 1)) / (double) n) - ((0.5 + (double) i) / (double) n); };
    (void) h;
    if (ns == 1) {
        D[0] = 1.0; D[ns] = 1.0;                 // natural: a straight line
    } else {
        {
            const double h0 = interval(0), h1 = interval(1), s0 = slope(0), s1 = slope(1);
            D[0] = h0 - h1; U[0] = 2 * h0 + h1; Y[0] = -6 * h0 / (h0 + h1) * (s0 - s1);
        }
        {
            const double h0 = interval(ns - 1), h1 = interval(ns - 2), s0 = slope(ns - 1), s1 = slope(ns - 2);
            D[ns] = h0 - h1; L[ns] = 2 * h0 + h1; Y[ns] = 6 * h0 / (h0 + h1) * (s0 - s1);
        }
    }
    for (int i = 1; i < ns; i++) {
        L[i] = interval(i - 1);
        D[i] = 2 * (interval(i - 1) + interval(i));
        U[i] = interval(i);
        Y[i] = 6 * (slope(i) - slope(i - 1));
    }
    const int dim = ns + 1;
    for (int i = 0; i + 1 < dim; i++) {
        if (fabs(D[i]) >= fabs(L[i + 1])) {
            const double w = L[i + 1] / D[i];
            D[i + 1] -= w * U[i];
            Y[i + 1] -= w * Y[i];
            L[i + 1] = 0.0;
        } else {                                 // interchange rows i and i + 1, then eliminate
            const double w = D[i] / L[i + 1], u = D[i + 1];
            D[i] = L[i + 1];
            D[i + 1] = U[i] - w * u;
            L[i + 1] = U[i + 1];
            U[i + 1] *= -w;
            U[i] = u;
            const double r = Y[i];
            Y[i] = Y[i + 1];
            Y[i + 1] = r - w * Y[i + 1];
        }
    }
    Y[dim - 1] /= D[dim - 1];
    for (int rev = 2; rev <= dim; rev++) {
        const int i = dim - rev;
        if (i == dim - 2) Y[i] -= U[i] * Y[i + 1];
        else Y[i] -= U[i] * Y[i + 1] + L[i + 1] * Y[i + 2];
        Y[i] /= D[i];
    }
}

// One block per chain: spline through the source chain, resampled along the destination
// chain (resample_chain, transition_interphase.cpp:13-41).  Dynamic shared memory:
// 3 coordinates x (V | L | D | U | Y) x max_knots doubles.
__global__ void __launch_bounds__(256) k_transition_spline(const float4 *__restrict__ src, float4 *__restrict__ dst,
                                                           const int *__restrict__ src_ranges,
                                                           const int *__restrict__ dst_ranges, int max_knots, int bits)
{
    extern __shared__ double sh_spline[];
    const int c = blockIdx.x;
    const int s0 = src_ranges[2 * c], n = src_ranges[2 * c + 1] - s0;
    const int d0 = dst_ranges[2 * c], m = dst_ranges[2 * c + 1] - d0;
    if (n <= 0 || m <= 0) return;
    double *V[3], *L[3], *D[3], *U[3], *Y[3];
    for (int k = 0; k < 3; k++) {
        double *base = sh_spline + (size_t) k * 5 * max_knots;
        V[k] = base; L[k] = base + max_knots; D[k] = base + 2 * max_knots; U[k] = base + 3 * max_knots;
        Y[k] = base + 4 * max_knots;
    }
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const float4 p = src[s0 + i];
        V[0][i] = quantize_d((double) p.x, bits);
        V[1][i] = quantize_d((double) p.y, bits);
        V[2][i] = quantize_d((double) p.z, bits);
    }
    __syncthreads();
    // n >= 4: the reference's not-a-knot system; n == 2: a line (its "natural" fallback);
    // n == 3: that system is singular in the reference (NaN); the parabola through the three
    // points is used instead; n == 1: the point itself.
    if (threadIdx.x < 3 && (n >= 4 || n == 2)) {
        spline_second_derivatives(n, L[threadIdx.x], D[threadIdx.x], U[threadIdx.x], Y[threadIdx.x], V[threadIdx.x]);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < m; i += blockDim.x) {
        const double t = (0.5 + (double) i) / (double) m;
        double out[3];
        if (n == 1) {
            for (int k = 0; k < 3; k++) out[k] = V[k][0];
        } else if (n == 3) {
            const double t0 = 0.5 / 3.0, t1 = 1.5 / 3.0, t2 = 2.5 / 3.0;
            const double l0 = (t - t1) * (t - t2) / ((t0 - t1) * (t0 - t2));
            const double l1 = (t - t0) * (t - t2) / ((t1 - t0) * (t1 - t2));
            const double l2 = (t - t0) * (t - t1) / ((t2 - t0) * (t2 - t1));
            for (int k = 0; k < 3; k++) out[k] = V[k][0] * l0 + V[k][1] * l1 + V[k][2] * l2;
        } else {
            const int ns = n - 1;
            auto knot = [&](int k) { return (0.5 + (double) k) / (double) n; };
            int seg;
            if (t <= knot(0)) seg = 0;                     // find_spline: boundary pieces extend outwards
            else if (t >= knot(ns)) seg = ns - 1;
            else {
                seg = min(max((int) floor(t * (double) n - 0.5), 0), ns - 1);
                while (seg + 1 < ns && t >= knot(seg + 1)) seg++;
                while (seg > 0 && t < knot(seg)) seg--;
            }
            const double hk = knot(seg + 1) - knot(seg), dt = t - knot(seg);
            for (int k = 0; k < 3; k++) {
                const double M0 = Y[k][seg], M1 = Y[k][seg + 1];
                const double sl = (V[k][seg + 1] - V[k][seg]) / hk;
                const double c0 = V[k][seg], c1 = sl - (M1 + 2 * M0) * hk / 6, c2 = M0 / 2, c3 = (M1 - M0) / (6 * hk);
                double v = c3;
                v = v * dt + c2;
                v = v * dt + c1;
                v = v * dt + c0;
                out[k] = v;
            }
        }
        dst[d0 + i] = make_float4((float) quantize_d(out[0], bits), (float) quantize_d(out[1], bits),
                                  (float) quantize_d(out[2], bits), 0.0f);
    }
}

// dst[pairs[k].y] = dst[pairs[k].x]: nucleolar beads start on top of their NOR bead
// (transition_interphase.cpp:76-78)
__global__ void k_transition_copy(float4 *__restrict__ dst, const int *__restrict__ pairs, int n_pairs)
{
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < n_pairs) dst[pairs[2 * k + 1]] = dst[pairs[2 * k]];
}

// One thread per coarse bead: centroid of `cg` consecutive source beads, written to the
// target chromatid and, displaced, to the sister (transition_prometaphase.cpp:81-98; the
// window is the reference's, clipped only at the end of the source array).
__global__ void __launch_bounds__(128) k_transition_coarse(const float4 *__restrict__ src, int n_src,
                                                           float4 *__restrict__ dst, int n_chains,
                                                           const int *__restrict__ src_ranges,
                                                           const int *__restrict__ target_ranges,
                                                           const int *__restrict__ sister_ranges, int cg,
                                                           double dx, double dy, double dz, int bits)
{
    const int c = blockIdx.y;
    if (c >= n_chains) return;
    const int off = blockIdx.x * blockDim.x + threadIdx.x;
    const int ts = target_ranges[2 * c], coarse_len = target_ranges[2 * c + 1] - ts;
    if (off >= coarse_len) return;
    const int s0 = src_ranges[2 * c], slen = src_ranges[2 * c + 1] - s0;
    const int a = min(s0 + cg * off, n_src - 1);
    const int b = min(min(a + cg, a + slen), n_src);
    double sx = 0.0, sy = 0.0, sz = 0.0, count = 0.0;
    for (int k = a; k < b; k++) {
        const float4 p = src[k];
        sx += quantize_d((double) p.x, bits);
        sy += quantize_d((double) p.y, bits);
        sz += quantize_d((double) p.z, bits);
        count += 1.0;
    }
    sx /= count; sy /= count; sz /= count;
    dst[ts + off] = make_float4((float) quantize_d(sx, bits), (float) quantize_d(sy, bits),
                                (float) quantize_d(sz, bits), 0.0f);
    if (sister_ranges) {
        const int ss = sister_ranges[2 * c];
        dst[ss + off] = make_float4((float) quantize_d(sx + dx, bits), (float) quantize_d(sy + dy, bits),
                                    (float) quantize_d(sz + dz, bits), 0.0f);
    }
}

// dst chain = src chain + displacement (transition_cycle.cpp:55-70)
__global__ void __launch_bounds__(128) k_transition_shift(const float4 *__restrict__ src, float4 *__restrict__ dst,
                                                          int n_chains, const int *__restrict__ src_ranges,
                                                          const int *__restrict__ dst_ranges, double dx, double dy,
                                                          double dz, int bits)
{
    const int c = blockIdx.y;
    if (c >= n_chains) return;
    const int off = blockIdx.x * blockDim.x + threadIdx.x;
    const int s0 = src_ranges[2 * c], d0 = dst_ranges[2 * c];
    const int len = min(src_ranges[2 * c + 1] - s0, dst_ranges[2 * c + 1] - d0);
    if (off >= len) return;
    const float4 p = src[s0 + off];
    dst[d0 + off] = make_float4((float) quantize_d(quantize_d((double) p.x, bits) + dx, bits),
                                (float) quantize_d(quantize_d((double) p.y, bits) + dy, bits),
                                (float) quantize_d(quantize_d((double) p.z, bits) + dz, bits), 0.0f);
}

} // namespace gc

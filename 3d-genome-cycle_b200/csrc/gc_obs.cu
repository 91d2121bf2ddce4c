// gc_obs.cu -- contact-map observables on the device.
//
// The reference computes these offline in Python from the trajectory file: contact triples ->
// cooler matrix (src/cool.py:80-118), expected contacts / P(s) = cis decay profile
// (src/pc1.py:156-183), O/E transform + column standardisation + first principal component
// by power iteration (src/pc1.py:36-101, src/pkg/pc1/math.py:13-37), NCI
// (src/pkg/nci/nci.py:9-35).  Here the ensemble contact map is already resident on the GPU as
// sorted (key = i << 32 | j, count) runs (gc_contacts_reduce_device / _merge_device), so the
// binned dense matrix and everything derived from it are built there and only the small
// results (profiles, tracks, scalars) cross to the host.
//
// All kernels are HBM-streaming passes over the n_bins x n_bins float32 matrix (the power
// iteration reads it twice per step); FP64 accumulators, float32 storage as in the reference.
#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <string>
#include <vector>

#include <cuda_runtime.h>

#include "../../include/gc_b200.h"

struct gc_obs {
    int device = 0;
    cudaStream_t stream = nullptr;
    int n_bins = 0, n_chroms = 0, max_len = 0, n_beads = 0;
    std::string error;
    float *d_M = nullptr, *d_D = nullptr;       // raw counts / working copy (O/E, standardised)
    int *d_chrom_of_bin = nullptr, *d_chrom_ranges = nullptr;
    int *d_bead_bin = nullptr, *d_bead_chain = nullptr;
    double *d_intra = nullptr;                  // [2][max_len]: sums, counts
    double *d_scal = nullptr;                   // inter sum, inter count, delta, norm, ...
    float *d_intra_f = nullptr;                 // intra as float32
    double *d_colstat = nullptr;                // [3][n]: nansum, count, aux
    float *d_mean = nullptr, *d_std = nullptr;
    int *d_sel = nullptr;
    float *d_v = nullptr, *d_w = nullptr, *d_part = nullptr;
    int row_chunks = 0;
    std::vector<void *> allocs;
};

namespace {

constexpr int ROWS_PER_CHUNK = 64;

int ofail(gc_obs *o, int code, const char *fmt, ...)
{
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (o) o->error = buf;
    return code;
}

#define OBS_CUDA(o, expr)                                                                       \
    do {                                                                                        \
        cudaError_t e_ = (expr);                                                                \
        if (e_ != cudaSuccess) {                                                                \
            return ofail(o, GC_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e_), \
                         __FILE__, __LINE__);                                                   \
        }                                                                                       \
    } while (0)

template<typename T>
int oalloc(gc_obs *o, T **p, size_t count)
{
    void *q = nullptr;
    OBS_CUDA(o, cudaMalloc(&q, std::max<size_t>(count, 1) * sizeof(T)));
    OBS_CUDA(o, cudaMemsetAsync(q, 0, std::max<size_t>(count, 1) * sizeof(T), o->stream));
    o->allocs.push_back(q);
    *p = (T *) q;
    return GC_OK;
}

// ---- accumulation -----------------------------------------------------------------------
// cooler's matrix() of upper-triangle pixels: symmetric, the diagonal counted once
__global__ void k_obs_add_runs(const unsigned long long *__restrict__ keys, const unsigned int *__restrict__ counts,
                               size_t n, const int *__restrict__ bead_bin, int n_beads, float *M, int nb)
{
    for (size_t t = (size_t) blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (size_t) gridDim.x * blockDim.x) {
        const unsigned long long k = keys[t];
        const unsigned i = (unsigned) (k >> 32), j = (unsigned) k;
        if (i >= (unsigned) n_beads || j >= (unsigned) n_beads) continue;
        const int bi = bead_bin[i], bj = bead_bin[j];
        if (bi < 0 || bj < 0 || bi >= nb || bj >= nb) continue;
        const float c = (float) counts[t];
        atomicAdd(&M[(size_t) bi * nb + bj], c);
        if (bi != bj) atomicAdd(&M[(size_t) bj * nb + bi], c);
    }
}

__global__ void k_obs_add_triples(const unsigned int *__restrict__ ijv, size_t n, const int *__restrict__ bead_bin,
                                  int n_beads, float *M, int nb)
{
    for (size_t t = (size_t) blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (size_t) gridDim.x * blockDim.x) {
        const unsigned i = ijv[3 * t], j = ijv[3 * t + 1];
        if (i >= (unsigned) n_beads || j >= (unsigned) n_beads) continue;
        const int bi = bead_bin[i], bj = bead_bin[j];
        if (bi < 0 || bj < 0 || bi >= nb || bj >= nb) continue;
        const float c = (float) ijv[3 * t + 2];
        atomicAdd(&M[(size_t) bi * nb + bj], c);
        if (bi != bj) atomicAdd(&M[(size_t) bj * nb + bi], c);
    }
}

// P(s) at BEAD resolution straight from the sorted runs (the numpy pass of round 1's
// ensemble.contact_probability): hist[s] += count for cis pairs with j - i = s <= max_s
__global__ void k_obs_ps_runs(const unsigned long long *__restrict__ keys, const unsigned int *__restrict__ counts,
                              size_t n, const int *__restrict__ bead_chain, int n_beads, int max_s,
                              unsigned long long *hist)
{
    for (size_t t = (size_t) blockIdx.x * blockDim.x + threadIdx.x; t < n; t += (size_t) gridDim.x * blockDim.x) {
        const unsigned long long k = keys[t];
        const unsigned i = (unsigned) (k >> 32), j = (unsigned) k;
        if (i >= (unsigned) n_beads || j >= (unsigned) n_beads || j <= i) continue;
        const int ci = bead_chain[i];
        if (ci < 0 || ci != bead_chain[j]) continue;
        const unsigned s = j - i;
        if (s <= (unsigned) max_s) atomicAdd(&hist[s], (unsigned long long) counts[t]);
    }
}

// ---- expected contacts (src/pc1.py:156-183) ------------------------------------------------
// one block per row i: finite cis entries on the upper diagonals -> intra[j - i], every finite
// trans entry -> inter
__global__ void __launch_bounds__(256) k_obs_expected(const float *__restrict__ M, int nb,
                                                      const int *__restrict__ chrom_of_bin, double *intra_sum,
                                                      double *intra_cnt, double *inter)
{
    const int i = blockIdx.x, ci = chrom_of_bin[i];
    double ts = 0.0, tc = 0.0;
    for (int j = threadIdx.x; j < nb; j += blockDim.x) {
        const float v = M[(size_t) i * nb + j];
        if (!isfinite(v)) continue;
        if (chrom_of_bin[j] == ci) {
            if (j >= i) {
                atomicAdd(&intra_sum[j - i], (double) v);
                atomicAdd(&intra_cnt[j - i], 1.0);
            }
        } else {
            ts += (double) v;
            tc += 1.0;
        }
    }
    __shared__ double sh[2][8];
    for (int o = 16; o > 0; o >>= 1) {
        ts += __shfl_xor_sync(0xffffffffu, ts, o);
        tc += __shfl_xor_sync(0xffffffffu, tc, o);
    }
    if ((threadIdx.x & 31) == 0) { sh[0][threadIdx.x >> 5] = ts; sh[1][threadIdx.x >> 5] = tc; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double a = 0, b = 0;
        for (int w = 0; w < 8; w++) { a += sh[0][w]; b += sh[1][w]; }
        if (b > 0) { atomicAdd(&inter[0], a); atomicAdd(&inter[1], b); }
    }
}

__global__ void k_obs_intra_finish(const double *intra_sum, const double *intra_cnt, float *intra_f, int n)
{
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s < n) intra_f[s] = (float) (intra_sum[s] / intra_cnt[s]);        // 0 / 0 -> NaN as numpy
}

// ---- O/E (src/pc1.py:44-58) -----------------------------------------------------------------
__global__ void k_obs_oe(const float *__restrict__ M, float *__restrict__ D, int nb,
                         const int *__restrict__ chrom_of_bin, const float *__restrict__ intra_f, float inter,
                         int mask_intra)
{
    const size_t total = (size_t) nb * nb;
    for (size_t t = (size_t) blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (size_t) gridDim.x * blockDim.x) {
        const int i = (int) (t / nb), j = (int) (t % nb);
        const float v = M[t];
        float r;
        if (chrom_of_bin[i] != chrom_of_bin[j]) r = v / inter;
        else if (mask_intra) r = isfinite(v) ? 1.0f : v;
        else r = v / intra_f[abs(i - j)];
        D[t] = r;
    }
}

// ---- column statistics (np.nansum / nanmean / nanstd over axis 0) ---------------------------
// thread per column, rows in chunks: partial (sum, count, sum of squared deviations)
__global__ void k_obs_col_pass1(const float *__restrict__ D, int nb, double *colstat)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= nb) return;
    const int r0 = blockIdx.y * ROWS_PER_CHUNK, r1 = min(r0 + ROWS_PER_CHUNK, nb);
    double s = 0.0, c = 0.0;
    for (int i = r0; i < r1; i++) {
        const float v = D[(size_t) i * nb + j];
        if (v == v) { s += (double) v; c += 1.0; }
    }
    atomicAdd(&colstat[j], s);
    atomicAdd(&colstat[nb + j], c);
}

__global__ void k_obs_col_pass2(const float *__restrict__ D, int nb, const double *colstat, double *var_acc)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= nb) return;
    const int r0 = blockIdx.y * ROWS_PER_CHUNK, r1 = min(r0 + ROWS_PER_CHUNK, nb);
    const float mean = (float) (colstat[j] / colstat[nb + j]);
    double q = 0.0;
    for (int i = r0; i < r1; i++) {
        const float v = D[(size_t) i * nb + j];
        if (v == v) { const double d = (double) (v - mean); q += d * d; }
    }
    atomicAdd(&var_acc[j], q);
}

__global__ void k_obs_col_finish(int nb, const double *colstat, const double *var_acc, float *mean, float *std_,
                                 int *sel)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= nb) return;
    const double s = colstat[j], c = colstat[nb + j];
    sel[j] = s > 0.0 ? 1 : 0;                                        // coverages > 0
    mean[j] = (float) (s / c);
    std_[j] = (float) sqrt(var_acc[j] / c);
}

__global__ void k_obs_standardise(float *__restrict__ D, int nb, const float *__restrict__ mean,
                                  const float *__restrict__ std_, const int *__restrict__ sel, int use_cov)
{
    const size_t total = (size_t) nb * nb;
    for (size_t t = (size_t) blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (size_t) gridDim.x * blockDim.x) {
        const int j = (int) (t % nb);
        float v = 0.0f;                                              // dropped column: takes no part
        if (sel[j]) {
            v = D[t] - mean[j];
            if (!use_cov) v = v / std_[j];
        }
        D[t] = v;
    }
}

// ---- power iteration (src/pkg/pc1/math.py:13-37) ----------------------------------------------
__global__ void __launch_bounds__(256) k_obs_rowdot(const float *__restrict__ D, int nb, const float *__restrict__ v,
                                                    float *__restrict__ w)
{
    const int i = blockIdx.x;
    float s = 0.0f;
    for (int j = threadIdx.x; j < nb; j += blockDim.x) s = fmaf(D[(size_t) i * nb + j], v[j], s);   // NaN propagates
    __shared__ float sh[8];
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        float a = 0.0f;
        for (int k = 0; k < 8; k++) a += sh[k];
        w[i] = a;
    }
}

// partial[chunk][j] = nansum over the chunk's rows of w_i D_ij
__global__ void k_obs_colacc(const float *__restrict__ D, int nb, const float *__restrict__ w, float *__restrict__ part)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= nb) return;
    const int r0 = blockIdx.y * ROWS_PER_CHUNK, r1 = min(r0 + ROWS_PER_CHUNK, nb);
    float s = 0.0f;
    for (int i = r0; i < r1; i++) {
        const float t = w[i] * D[(size_t) i * nb + j];
        if (t == t) s += t;
    }
    part[(size_t) blockIdx.y * nb + j] = s;
}

// one block: u = sum of the partials (fixed order), vec = u / |u|, delta = max |vec - prev|
__global__ void __launch_bounds__(1024) k_obs_iter_finish(int nb, int chunks, float *part,
                                                          float *v, const int *__restrict__ sel, double *scal)
{
    __shared__ double sh[32];
    __shared__ float sh_norm;
    double q = 0.0;
    for (int j = threadIdx.x; j < nb; j += blockDim.x) {
        float u = 0.0f;
        if (sel[j]) {
            for (int c = 0; c < chunks; c++) u += part[(size_t) c * nb + j];
        }
        part[j] = u;                                   // chunk 0's row now holds u
        q += (double) u * (double) u;
    }
    for (int o = 16; o > 0; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = q;
    __syncthreads();
    if (threadIdx.x == 0) {
        double a = 0;
        for (int k = 0; k < (int) (blockDim.x + 31) / 32; k++) a += sh[k];
        sh_norm = (float) sqrt(a);
    }
    __syncthreads();
    const float inv = 1.0f / sh_norm;
    float dmax = 0.0f;
    bool bad = false;
    for (int j = threadIdx.x; j < nb; j += blockDim.x) {
        if (!sel[j]) continue;
        const float nv = part[j] * inv;
        const float d = fabsf(nv - v[j]);
        if (d != d) bad = true;
        dmax = fmaxf(dmax, d);
        v[j] = nv;
    }
    double dm = bad ? (double) NAN : (double) dmax;
    __syncthreads();
    for (int o = 16; o > 0; o >>= 1) {
        const double x = __shfl_xor_sync(0xffffffffu, dm, o);
        dm = (dm != dm || x != x) ? (double) NAN : fmax(dm, x);
    }
    if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = dm;
    __syncthreads();
    if (threadIdx.x == 0) {
        double a = 0;
        for (int k = 0; k < (int) (blockDim.x + 31) / 32; k++) a = (a != a || sh[k] != sh[k]) ? (double) NAN : fmax(a, sh[k]);
        scal[2] = a;
    }
}

__global__ void k_obs_init_vec(int nb, const int *__restrict__ sel, float *v, float value)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j < nb) v[j] = sel[j] ? value : 0.0f;
}

// ---- NCI (src/pkg/nci/nci.py:9-35) ---------------------------------------------------------------
__global__ void k_obs_nci(const float *__restrict__ M, int nb, const int *__restrict__ chrom_of_bin, float *out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nb) return;
    float r = NAN;
    if (i + 1 < nb && chrom_of_bin[i + 1] == chrom_of_bin[i]) {
        const float x = M[(size_t) i * nb + i], y = M[(size_t) (i + 1) * nb + i + 1];
        r = M[(size_t) i * nb + i + 1] / sqrtf(y * x);
    }
    out[i] = r;
}

// ---- contact enrichment by bin type (compartment strength) -----------------------------------------
// sums[a][b] += value, counts[a][b] += 1 over finite entries with i < j and (trans or j - i >= min_sep)
__global__ void __launch_bounds__(256) k_obs_type_enrich(const float *__restrict__ X, int nb,
                                                         const int *__restrict__ chrom_of_bin,
                                                         const int *__restrict__ bin_type, int n_types, int min_sep,
                                                         double *sums, double *counts)
{
    __shared__ double sh_s[64], sh_c[64];
    for (int k = threadIdx.x; k < 64; k += blockDim.x) { sh_s[k] = 0.0; sh_c[k] = 0.0; }
    __syncthreads();
    const int i = blockIdx.x, ti = bin_type[i], ci = chrom_of_bin[i];
    if (ti >= 0 && ti < n_types) {
        for (int j = i + 1 + threadIdx.x; j < nb; j += blockDim.x) {
            const int tj = bin_type[j];
            if (tj < 0 || tj >= n_types) continue;
            if (chrom_of_bin[j] == ci && j - i < min_sep) continue;
            const float v = X[(size_t) i * nb + j];
            if (!isfinite(v)) continue;
            const int a = min(ti, tj), b = max(ti, tj);
            atomicAdd(&sh_s[a * n_types + b], (double) v);
            atomicAdd(&sh_c[a * n_types + b], 1.0);
        }
    }
    __syncthreads();
    for (int k = threadIdx.x; k < n_types * n_types; k += blockDim.x) {
        if (sh_c[k] > 0) { atomicAdd(&sums[k], sh_s[k]); atomicAdd(&counts[k], sh_c[k]); }
    }
}

int grid_for(size_t n, int bs = 256) { return (int) std::min<size_t>((n + bs - 1) / bs, 148 * 16); }

int expected_device(gc_obs *o)
{
    const int nb = o->n_bins;
    OBS_CUDA(o, cudaMemsetAsync(o->d_intra, 0, 2 * (size_t) o->max_len * sizeof(double), o->stream));
    OBS_CUDA(o, cudaMemsetAsync(o->d_scal, 0, 8 * sizeof(double), o->stream));
    k_obs_expected<<<nb, 256, 0, o->stream>>>(o->d_M, nb, o->d_chrom_of_bin, o->d_intra, o->d_intra + o->max_len, o->d_scal);
    k_obs_intra_finish<<<(o->max_len + 127) / 128, 128, 0, o->stream>>>(o->d_intra, o->d_intra + o->max_len,
                                                                     o->d_intra_f, o->max_len);
    return GC_OK;
}

} // namespace

extern "C" {

const char *gc_obs_last_error(const gc_obs *o) { return o ? o->error.c_str() : "null handle"; }

int gc_obs_create(int32_t device, int32_t n_bins, int32_t n_chroms, const int32_t *chrom_ranges, gc_obs **out)
{
    if (!out || n_bins < 1 || n_chroms < 1 || !chrom_ranges) return GC_ERR_INVALID;
    *out = nullptr;
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || device < 0 || device >= count) {
        fprintf(stderr, "gc_obs: no such CUDA device %d (this library has no CPU path)\n", device);
        return GC_ERR_CUDA;
    }
    gc_obs *o = new gc_obs;
    *out = o;
    o->device = device;
    o->n_bins = n_bins;
    o->n_chroms = n_chroms;
    OBS_CUDA(o, cudaSetDevice(device));
    OBS_CUDA(o, cudaStreamCreateWithFlags(&o->stream, cudaStreamNonBlocking));
    std::vector<int> chrom_of((size_t) n_bins, -1);
    int prev_end = 0;
    for (int c = 0; c < n_chroms; c++) {
        const int s = chrom_ranges[2 * c], e = chrom_ranges[2 * c + 1];
        if (s < prev_end || e < s || e > n_bins) return ofail(o, GC_ERR_INVALID, "chromosome range %d = [%d, %d) invalid", c, s, e);
        for (int b = s; b < e; b++) chrom_of[b] = c;
        o->max_len = std::max(o->max_len, e - s);
        prev_end = e;
    }
    // bins outside every chromosome form singleton groups (ids below -1 are distinct)
    for (int b = 0; b < n_bins; b++) if (chrom_of[b] < 0) chrom_of[b] = -2 - b;
    o->max_len = std::max(o->max_len, 1);
    const size_t nn = (size_t) n_bins * n_bins;
    o->row_chunks = (n_bins + ROWS_PER_CHUNK - 1) / ROWS_PER_CHUNK;
    int rc = 0;
    if (!rc) rc = oalloc(o, &o->d_M, nn);
    if (!rc) rc = oalloc(o, &o->d_D, nn);
    if (!rc) rc = oalloc(o, &o->d_chrom_of_bin, (size_t) n_bins);
    if (!rc) rc = oalloc(o, &o->d_chrom_ranges, 2 * (size_t) n_chroms);
    if (!rc) rc = oalloc(o, &o->d_intra, 2 * (size_t) o->max_len);
    if (!rc) rc = oalloc(o, &o->d_intra_f, (size_t) o->max_len);
    if (!rc) rc = oalloc(o, &o->d_scal, 8);
    if (!rc) rc = oalloc(o, &o->d_colstat, 3 * (size_t) n_bins);
    if (!rc) rc = oalloc(o, &o->d_mean, (size_t) n_bins);
    if (!rc) rc = oalloc(o, &o->d_std, (size_t) n_bins);
    if (!rc) rc = oalloc(o, &o->d_sel, (size_t) n_bins);
    if (!rc) rc = oalloc(o, &o->d_v, (size_t) n_bins);
    if (!rc) rc = oalloc(o, &o->d_w, (size_t) n_bins);
    if (!rc) rc = oalloc(o, &o->d_part, (size_t) o->row_chunks * n_bins);
    if (rc) return rc;
    OBS_CUDA(o, cudaMemcpyAsync(o->d_chrom_of_bin, chrom_of.data(), chrom_of.size() * sizeof(int), cudaMemcpyHostToDevice, o->stream));
    OBS_CUDA(o, cudaMemcpyAsync(o->d_chrom_ranges, chrom_ranges, 2 * (size_t) n_chroms * sizeof(int), cudaMemcpyHostToDevice, o->stream));
    OBS_CUDA(o, cudaStreamSynchronize(o->stream));
    return GC_OK;
}

int gc_obs_destroy(gc_obs *o)
{
    if (!o) return GC_ERR_INVALID;
    cudaSetDevice(o->device);
    if (o->stream) cudaStreamSynchronize(o->stream);
    for (void *p : o->allocs) cudaFree(p);
    if (o->stream) cudaStreamDestroy(o->stream);
    delete o;
    return GC_OK;
}

int gc_obs_set_bead_map(gc_obs *o, int32_t n_beads, const int32_t *bead_bin, const int32_t *bead_chain)
{
    if (!o || n_beads < 0 || !bead_bin) return o ? ofail(o, GC_ERR_INVALID, "bad args") : GC_ERR_INVALID;
    OBS_CUDA(o, cudaSetDevice(o->device));
    int rc = oalloc(o, &o->d_bead_bin, (size_t) n_beads);
    if (!rc) rc = oalloc(o, &o->d_bead_chain, (size_t) n_beads);
    if (rc) return rc;
    OBS_CUDA(o, cudaMemcpyAsync(o->d_bead_bin, bead_bin, (size_t) n_beads * sizeof(int), cudaMemcpyHostToDevice, o->stream));
    std::vector<int> none((size_t) n_beads, -1);
    OBS_CUDA(o, cudaMemcpyAsync(o->d_bead_chain, bead_chain ? bead_chain : none.data(), (size_t) n_beads * sizeof(int),
                                cudaMemcpyHostToDevice, o->stream));
    OBS_CUDA(o, cudaStreamSynchronize(o->stream));
    o->n_beads = n_beads;
    return GC_OK;
}

int gc_obs_clear(gc_obs *o)
{
    if (!o) return GC_ERR_INVALID;
    OBS_CUDA(o, cudaSetDevice(o->device));
    OBS_CUDA(o, cudaMemsetAsync(o->d_M, 0, (size_t) o->n_bins * o->n_bins * sizeof(float), o->stream));
    return GC_OK;
}

int gc_obs_add_runs_device(gc_obs *o, const uint64_t *keys, const uint32_t *counts, size_t n)
{
    if (!o) return GC_ERR_INVALID;
    if (!o->d_bead_bin) return ofail(o, GC_ERR_STATE, "gc_obs_set_bead_map has not been called");
    if (n == 0) return GC_OK;
    if (!keys || !counts) return ofail(o, GC_ERR_INVALID, "bad args");
    OBS_CUDA(o, cudaSetDevice(o->device));
    OBS_CUDA(o, cudaDeviceSynchronize());                // the producer's stream is not ours
    k_obs_add_runs<<<grid_for(n), 256, 0, o->stream>>>((const unsigned long long *) keys, counts, n, o->d_bead_bin,
                                                       o->n_beads, o->d_M, o->n_bins);
    OBS_CUDA(o, cudaStreamSynchronize(o->stream));
    OBS_CUDA(o, cudaGetLastError());
    return GC_OK;
}

int gc_obs_add_triples(gc_obs *o, const uint32_t *ijv, size_t n)
{
    if (!o) return GC_ERR_INVALID;
    if (!o->d_bead_bin) return ofail(o, GC_ERR_STATE, "gc_obs_set_bead_map has not been called");
    if (n == 0) return GC_OK;
    if (!ijv) return ofail(o, GC_ERR_INVALID, "bad args");
    OBS_CUDA(o, cudaSetDevice(o->device));
    unsigned int *d = nullptr;
    OBS_CUDA(o, cudaMalloc(&d, 3 * n * sizeof(unsigned int)));
    cudaError_t e = cudaMemcpyAsync(d, ijv, 3 * n * sizeof(unsigned int), cudaMemcpyHostToDevice, o->stream);
    if (e == cudaSuccess) {
        k_obs_add_triples<<<grid_for(n), 256, 0, o->stream>>>(d, n, o->d_bead_bin, o->n_beads, o->d_M, o->n_bins);
        e = cudaStreamSynchronize(o->stream);
    }
    cudaFree(d);
    OBS_CUDA(o, e);
    return GC_OK;
}

int gc_obs_ps_runs_device(gc_obs *o, const uint64_t *keys, const uint32_t *counts, size_t n, int32_t max_s,
                          uint64_t *hist)
{
    if (!o || max_s < 0 || !hist) return o ? ofail(o, GC_ERR_INVALID, "bad args") : GC_ERR_INVALID;
    if (!o->d_bead_chain) return ofail(o, GC_ERR_STATE, "gc_obs_set_bead_map has not been called");
    OBS_CUDA(o, cudaSetDevice(o->device));
    unsigned long long *d = nullptr;
    OBS_CUDA(o, cudaMalloc(&d, ((size_t) max_s + 1) * sizeof(unsigned long long)));
    cudaError_t e = cudaMemsetAsync(d, 0, ((size_t) max_s + 1) * sizeof(unsigned long long), o->stream);
    if (e == cudaSuccess && n > 0) {
        cudaDeviceSynchronize();
        k_obs_ps_runs<<<grid_for(n), 256, 0, o->stream>>>((const unsigned long long *) keys, counts, n, o->d_bead_chain,
                                                          o->n_beads, max_s, d);
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(hist, d, ((size_t) max_s + 1) * sizeof(unsigned long long), cudaMemcpyDeviceToHost, o->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(o->stream);
    cudaFree(d);
    OBS_CUDA(o, e);
    return GC_OK;
}

int gc_obs_get_matrix(gc_obs *o, float *out)
{
    if (!o || !out) return GC_ERR_INVALID;
    OBS_CUDA(o, cudaSetDevice(o->device));
    OBS_CUDA(o, cudaMemcpyAsync(out, o->d_M, (size_t) o->n_bins * o->n_bins * sizeof(float), cudaMemcpyDeviceToHost, o->stream));
    OBS_CUDA(o, cudaStreamSynchronize(o->stream));
    return GC_OK;
}

int gc_obs_set_matrix(gc_obs *o, const float *in)
{
    if (!o || !in) return GC_ERR_INVALID;
    OBS_CUDA(o, cudaSetDevice(o->device));
    OBS_CUDA(o, cudaMemcpyAsync(o->d_M, in, (size_t) o->n_bins * o->n_bins * sizeof(float), cudaMemcpyHostToDevice, o->stream));
    OBS_CUDA(o, cudaStreamSynchronize(o->stream));
    return GC_OK;
}

int gc_obs_expected(gc_obs *o, double *intra, int32_t capacity, double *inter)
{
    if (!o) return GC_ERR_INVALID;
    OBS_CUDA(o, cudaSetDevice(o->device));
    int rc = expected_device(o);
    if (rc) return rc;
    std::vector<double> h(2 * (size_t) o->max_len);
    double sc[2];
    OBS_CUDA(o, cudaMemcpyAsync(h.data(), o->d_intra, h.size() * sizeof(double), cudaMemcpyDeviceToHost, o->stream));
    OBS_CUDA(o, cudaMemcpyAsync(sc, o->d_scal, sizeof sc, cudaMemcpyDeviceToHost, o->stream));
    OBS_CUDA(o, cudaStreamSynchronize(o->stream));
    if (intra) {
        for (int s = 0; s < std::min<int>(capacity, o->max_len); s++) intra[s] = h[s] / h[(size_t) o->max_len + s];
    }
    if (inter) *inter = sc[0] / sc[1];
    return GC_OK;
}

int gc_obs_max_separation(gc_obs *o, int32_t *out)
{
    if (!o || !out) return GC_ERR_INVALID;
    *out = o->max_len;
    return GC_OK;
}

int gc_obs_pc1(gc_obs *o, int32_t mask_intra, int32_t use_covariance, double tolerance, int32_t max_steps,
               float *ev1, float *pc1, double *explained_variance_ratio, int32_t *steps_out)
{
    if (!o || !(tolerance > 0) || max_steps < 1) return o ? ofail(o, GC_ERR_INVALID, "bad args") : GC_ERR_INVALID;
    OBS_CUDA(o, cudaSetDevice(o->device));
    const int nb = o->n_bins;
    const size_t nn = (size_t) nb * nb;
    int rc = expected_device(o);
    if (rc) return rc;
    double sc[2];
    OBS_CUDA(o, cudaMemcpyAsync(sc, o->d_scal, sizeof sc, cudaMemcpyDeviceToHost, o->stream));
    OBS_CUDA(o, cudaStreamSynchronize(o->stream));
    const float inter = (float) (sc[0] / sc[1]);
    k_obs_oe<<<grid_for(nn), 256, 0, o->stream>>>(o->d_M, o->d_D, nb, o->d_chrom_of_bin, o->d_intra_f, inter, mask_intra);
    // column statistics of the O/E matrix: coverage > 0 selects, nanmean / nanstd standardise
    const dim3 cg((nb + 127) / 128, o->row_chunks);
    OBS_CUDA(o, cudaMemsetAsync(o->d_colstat, 0, 3 * (size_t) nb * sizeof(double), o->stream));
    k_obs_col_pass1<<<cg, 128, 0, o->stream>>>(o->d_D, nb, o->d_colstat);
    k_obs_col_pass2<<<cg, 128, 0, o->stream>>>(o->d_D, nb, o->d_colstat, o->d_colstat + 2 * (size_t) nb);
    k_obs_col_finish<<<(nb + 127) / 128, 128, 0, o->stream>>>(nb, o->d_colstat, o->d_colstat + 2 * (size_t) nb, o->d_mean,
                                                            o->d_std, o->d_sel);
    k_obs_standardise<<<grid_for(nn), 256, 0, o->stream>>>(o->d_D, nb, o->d_mean, o->d_std, o->d_sel, use_covariance);
    std::vector<int> sel((size_t) nb);
    OBS_CUDA(o, cudaMemcpyAsync(sel.data(), o->d_sel, sel.size() * sizeof(int), cudaMemcpyDeviceToHost, o->stream));
    OBS_CUDA(o, cudaStreamSynchronize(o->stream));
    int dim = 0;
    for (int s : sel) dim += s;
    if (dim == 0) return ofail(o, GC_ERR_STATE, "no bin with positive coverage");
    k_obs_init_vec<<<(nb + 127) / 128, 128, 0, o->stream>>>(nb, o->d_sel, o->d_v, 1.0f / sqrtf((float) dim));
    int steps = 0;
    double delta = 0.0;
    for (;;) {
        k_obs_rowdot<<<nb, 256, 0, o->stream>>>(o->d_D, nb, o->d_v, o->d_w);
        k_obs_colacc<<<cg, 128, 0, o->stream>>>(o->d_D, nb, o->d_w, o->d_part);
        k_obs_iter_finish<<<1, 1024, 0, o->stream>>>(nb, o->row_chunks, o->d_part, o->d_v, o->d_sel, o->d_scal);
        OBS_CUDA(o, cudaMemcpyAsync(&delta, o->d_scal + 2, sizeof(double), cudaMemcpyDeviceToHost, o->stream));
        OBS_CUDA(o, cudaStreamSynchronize(o->stream));
        steps++;
        if (delta < tolerance || steps >= max_steps || delta != delta) break;
    }
    // pc1 = data @ vec, rescaled so that its variance is the explained variance ratio
    k_obs_rowdot<<<nb, 256, 0, o->stream>>>(o->d_D, nb, o->d_v, o->d_w);
    OBS_CUDA(o, cudaMemsetAsync(o->d_colstat, 0, 3 * (size_t) nb * sizeof(double), o->stream));
    k_obs_col_pass1<<<cg, 128, 0, o->stream>>>(o->d_D, nb, o->d_colstat);
    k_obs_col_pass2<<<cg, 128, 0, o->stream>>>(o->d_D, nb, o->d_colstat, o->d_colstat + 2 * (size_t) nb);
    std::vector<float> v((size_t) nb), w((size_t) nb);
    std::vector<double> cs(3 * (size_t) nb);
    OBS_CUDA(o, cudaMemcpyAsync(v.data(), o->d_v, v.size() * sizeof(float), cudaMemcpyDeviceToHost, o->stream));
    OBS_CUDA(o, cudaMemcpyAsync(w.data(), o->d_w, w.size() * sizeof(float), cudaMemcpyDeviceToHost, o->stream));
    OBS_CUDA(o, cudaMemcpyAsync(cs.data(), o->d_colstat, cs.size() * sizeof(double), cudaMemcpyDeviceToHost, o->stream));
    OBS_CUDA(o, cudaStreamSynchronize(o->stream));
    OBS_CUDA(o, cudaGetLastError());
    double data_var = 0.0;                                   // nansum(nanvar(data, axis = 0)), selected columns
    for (int j = 0; j < nb; j++) {
        if (!sel[j]) continue;
        const double var = cs[2 * (size_t) nb + j] / cs[(size_t) nb + j];
        if (var == var) data_var += var;
    }
    const float scale = (float) sqrt(data_var);
    double m = 0.0, c = 0.0;
    for (int i = 0; i < nb; i++) {
        w[i] = w[i] / scale;
        if (w[i] == w[i]) { m += w[i]; c += 1.0; }
    }
    m /= c;
    double q = 0.0;
    for (int i = 0; i < nb; i++) if (w[i] == w[i]) q += ((double) w[i] - m) * ((double) w[i] - m);
    if (explained_variance_ratio) *explained_variance_ratio = q / c;
    if (steps_out) *steps_out = steps;
    for (int j = 0; j < nb; j++) {
        if (ev1) ev1[j] = sel[j] ? v[j] : NAN;
        if (pc1) pc1[j] = w[j];
    }
    return GC_OK;
}

int gc_obs_nci(gc_obs *o, float *out)
{
    if (!o || !out) return GC_ERR_INVALID;
    OBS_CUDA(o, cudaSetDevice(o->device));
    k_obs_nci<<<(o->n_bins + 127) / 128, 128, 0, o->stream>>>(o->d_M, o->n_bins, o->d_chrom_of_bin, o->d_w);
    OBS_CUDA(o, cudaMemcpyAsync(out, o->d_w, (size_t) o->n_bins * sizeof(float), cudaMemcpyDeviceToHost, o->stream));
    OBS_CUDA(o, cudaStreamSynchronize(o->stream));
    return GC_OK;
}

int gc_obs_type_enrichment(gc_obs *o, const int32_t *bin_type, int32_t n_types, int32_t min_separation,
                           double *mean_contact)
{
    if (!o || !bin_type || n_types < 1 || n_types > 8 || !mean_contact) return o ? ofail(o, GC_ERR_INVALID, "bad args") : GC_ERR_INVALID;
    OBS_CUDA(o, cudaSetDevice(o->device));
    int *d_type = nullptr;
    double *d_acc = nullptr;
    OBS_CUDA(o, cudaMalloc(&d_type, (size_t) o->n_bins * sizeof(int)));
    cudaError_t e = cudaMalloc(&d_acc, 128 * sizeof(double));
    if (e == cudaSuccess) e = cudaMemsetAsync(d_acc, 0, 128 * sizeof(double), o->stream);
    if (e == cudaSuccess) e = cudaMemcpyAsync(d_type, bin_type, (size_t) o->n_bins * sizeof(int), cudaMemcpyHostToDevice, o->stream);
    double h[128];
    if (e == cudaSuccess) {
        k_obs_type_enrich<<<o->n_bins, 256, 0, o->stream>>>(o->d_M, o->n_bins, o->d_chrom_of_bin, d_type, n_types,
                                                          min_separation, d_acc, d_acc + 64);
        e = cudaMemcpyAsync(h, d_acc, sizeof h, cudaMemcpyDeviceToHost, o->stream);
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(o->stream);
    cudaFree(d_type);
    if (d_acc) cudaFree(d_acc);
    OBS_CUDA(o, e);
    for (int a = 0; a < n_types; a++) {
        for (int b = 0; b < n_types; b++) {
            const int k = std::min(a, b) * n_types + std::max(a, b);
            mean_contact[a * n_types + b] = h[k] / h[64 + k];
        }
    }
    return GC_OK;
}

} // extern "C"

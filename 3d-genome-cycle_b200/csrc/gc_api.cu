// gc_api.cu -- implementation of the C-ABI in include/gc_b200.h on top of the
// sm_100a kernels in gc_kernels.cuh / gc_step.cuh / gc_query.cuh.
//
// There is NO CPU fallback here: every entry point either runs the CUDA path
// or fails with GC_ERR_CUDA.

#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <unordered_map>
#include <utility>
#include <vector>

#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_reduce.cuh>

#include "../../include/gc_b200.h"
#include "gc_query.cuh"
#include "gc_tile.cuh"
#include "gc_cells.cuh"
#include "gc_contact.cuh"
#include "gc_decomp.cuh"
#include "gc_transition.cuh"
#include "gc_peak.cuh"

using namespace gc;

// ---------------------------------------------------------------------------
// context
// ---------------------------------------------------------------------------
struct HostBond { int i, j, bterm; };
struct HostTriple { int i, j, k, tterm; };
struct HostTether { int bead, pole; float spring, length; int term; };

struct gc_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    int n = 0, n_rep = 1;
    std::string error;

    Params hp{};            // passed BY VALUE to every kernel (__grid_constant__)
    bool params_dirty = true;
    bool topo_dirty = true;
    bool attrs_set = false;
    int next_term = 0;

    // host-side topology before CSR construction
    std::vector<HostBond> bonds;
    std::vector<HostTriple> triples;
    std::vector<HostTether> tethers;
    std::vector<int> well_targets;

    // device allocations (freed in gc_destroy)
    std::vector<void *> allocs;
    std::vector<void *> ipc_mapped;          // peer buffers opened with cudaIpcOpenMemHandle
    bool decomp_all_current = false;         // every bead was stamped current and no step has run since
    float *d_mob = nullptr;
    float2 *d_ab = nullptr;
    int *d_flags = nullptr, *d_bond_off = nullptr, *d_triple_off = nullptr, *d_tether_off = nullptr;
    BondEntry *d_bond_ent = nullptr;
    TripleEntry *d_triple_ent = nullptr;
    TetherEntry *d_tether_ent = nullptr;
    int *d_well_targets = nullptr;
    float *d_frame = nullptr;   // gc_get_frame staging
    size_t frame_cap = 0;
    // whole-ensemble host transfers (gc_*_positions_all_f32, gc_frame_begin / _end): one
    // page-locked host buffer + one device buffer of packed f32, a copy stream and an event
    float *h_stage = nullptr, *d_stage = nullptr;
    size_t stage_cap = 0;
    float *h_frame = nullptr;
    size_t h_frame_cap = 0, frame_pending = 0;   // floats of the frame in flight (0 = none)
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t copy_event = nullptr;
    float *d_noise = nullptr;
    size_t noise_cap = 0;
    double *d_partials = nullptr;
    unsigned long long *d_counter = nullptr;
    void *d_scratch = nullptr;
    size_t scratch_cap = 0;
    // parked temporary buffers (dev_alloc / dev_free)
    std::vector<std::pair<void *, size_t>> cache;
    size_t cache_bytes = 0;
    std::unordered_map<void *, size_t> alloc_bytes;

    long long run_step = 0;     // host mirror of RepState::step
    int parity = 0;             // host mirror of RepState::parity (replicas move in lockstep)
    long long launches = 0;
    bool pdl = true;                // programmatic dependent launch between the step's kernels
    // CUDA-graph replay of the fixed-timestep step loop (gc_run)
    cudaGraphExec_t step_graph = nullptr;
    int step_graph_steps = 0;       // steps per replay
    // small systems (<= SMALL_MAX_BEADS beads, default cell budget): one-block rebuild,
    // fused with the step epilogue in fixed-dt runs (gc_kernels.cuh, gc_step.cuh)
    bool small = false;
    long long step_graph_launches = 0;
    Params step_graph_hp;           // kernel arguments baked into the graph
    long long step_graph_interval = -1;
    double skin = 0.0;
    int tile_grid_x = 1;        // blocks per replica of the pair kernel
    // Replica groups of the step graph: the replicas are independent trajectories, so the
    // graph runs them as `groups` parallel branches (one capture stream each).  The
    // latency-bound kernels of one group (cell scan, epilogue, count / fill) then overlap
    // the pair kernel of the others instead of leaving the GPU idle.
    int groups = 1;
    int group_grid_x = 1;       // pair-kernel blocks of one group
    bool dense_pair = false;    // launches being recorded use the 9-blocks-per-SM pair kernel
    std::vector<cudaStream_t> group_streams;
    std::vector<cudaEvent_t> group_events;   // [0] fork, [1 + g] join of group g


    // result of the last contact query (sorted keys / counts on the device)
    unsigned long long *res_keys = nullptr;
    unsigned int *res_vals = nullptr;
    size_t res_count = 0;

    // per-kernel event timing (gc_profile_*)
    struct ProfKernel {
        std::string name;
        std::vector<cudaEvent_t> ev;   // begin/end pairs
    };
    bool profiling = false;
    std::vector<ProfKernel> prof;
    size_t prof_events = 0;
    int prof_open = -1;
};

namespace {

int fail(gc_ctx *ctx, int code, const char *fmt, ...)
{
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (ctx) ctx->error = buf;
    return code;
}

#define GC_CUDA(ctx, expr)                                                                    \
    do {                                                                                      \
        cudaError_t e_ = (expr);                                                              \
        if (e_ != cudaSuccess) {                                                              \
            return fail(ctx, GC_ERR_CUDA, "%s failed: %s (%s:%d)", #expr,                     \
                        cudaGetErrorString(e_), __FILE__, __LINE__);                          \
        }                                                                                     \
    } while (0)

#define GC_CHECK_CTX(ctx)                                                                     \
    do {                                                                                      \
        if (!(ctx)) return GC_ERR_INVALID;                                                    \
        cudaError_t e_ = cudaSetDevice((ctx)->device);                                        \
        if (e_ != cudaSuccess)                                                                \
            return fail(ctx, GC_ERR_CUDA, "cudaSetDevice: %s", cudaGetErrorString(e_));       \
    } while (0)

// Temporary device buffers (contact reduction, frames, queries) come and go at output
// cadence; cudaMalloc / cudaFree synchronise the device and were seen to stall for hundreds
// of milliseconds, so released blocks of >= 256 KiB are parked in the context (at most 16
// blocks / 8 GiB) and handed out again to requests of at least half their size.
constexpr size_t CACHE_MIN_BYTES = 256u << 10, CACHE_MAX_BLOCKS = 16;
constexpr size_t CACHE_MAX_BYTES = (size_t) 8 << 30;

template<typename T>
int dev_alloc(gc_ctx *ctx, T **out, size_t count, int fill_byte = 0)
{
    void *p = nullptr;
    size_t bytes = std::max<size_t>(count, 1) * sizeof(T);
    int best = -1;
    for (int k = 0; k < (int) ctx->cache.size(); k++) {
        const size_t have = ctx->cache[k].second;
        if (have >= bytes && have <= 2 * bytes &&
            (best < 0 || have < ctx->cache[best].second)) {
            best = k;
        }
    }
    if (best >= 0) {
        p = ctx->cache[best].first;
        ctx->cache_bytes -= ctx->cache[best].second;
        ctx->cache.erase(ctx->cache.begin() + best);          // still listed in ctx->allocs
    } else {
        GC_CUDA(ctx, cudaMalloc(&p, bytes));
        ctx->allocs.push_back(p);
        ctx->alloc_bytes[p] = bytes;
    }
    GC_CUDA(ctx, cudaMemsetAsync(p, fill_byte, bytes, ctx->stream));
    *out = (T *) p;
    return GC_OK;
}

void dev_free(gc_ctx *ctx, void *p)
{
    if (!p) return;
    auto sz = ctx->alloc_bytes.find(p);
    const size_t bytes = sz != ctx->alloc_bytes.end() ? sz->second : 0;
    if (bytes >= CACHE_MIN_BYTES && ctx->cache.size() < CACHE_MAX_BLOCKS &&
        ctx->cache_bytes + bytes <= CACHE_MAX_BYTES) {
        // later work on the stream may still read it: order the reuse after it
        cudaStreamSynchronize(ctx->stream);
        ctx->cache.emplace_back(p, bytes);
        ctx->cache_bytes += bytes;
        return;
    }
    auto it = std::find(ctx->allocs.begin(), ctx->allocs.end(), p);
    if (it != ctx->allocs.end()) ctx->allocs.erase(it);
    if (sz != ctx->alloc_bytes.end()) ctx->alloc_bytes.erase(sz);
    cudaFree(p);
}

template<typename T>
int upload(gc_ctx *ctx, T **dst, const std::vector<T> &src)
{
    if (*dst) dev_free(ctx, *dst);
    *dst = nullptr;
    int rc = dev_alloc(ctx, dst, src.size());
    if (rc) return rc;
    if (!src.empty()) {
        GC_CUDA(ctx, cudaMemcpyAsync(*dst, src.data(), src.size() * sizeof(T),
                                     cudaMemcpyHostToDevice, ctx->stream));
        GC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));   // src may be a temporary
    }
    return GC_OK;
}

dim3 bead_grid(const gc_ctx *ctx, int bs) { return dim3((ctx->n + bs - 1) / bs, ctx->n_rep); }

int new_term(gc_ctx *ctx)
{
    if (ctx->next_term >= GC_MAX_TERMS) return fail(ctx, GC_ERR_INVALID, "too many terms");
    int t = ctx->next_term++;
    ctx->hp.enabled |= 1u << t;
    ctx->params_dirty = true;
    return t;
}

} // namespace

extern "C" {

const char *gc_version(void) { return "gc_b200 0.1 sm_100a"; }

const char *gc_last_error(const gc_ctx *ctx) { return ctx ? ctx->error.c_str() : "null context"; }

int gc_create(int32_t n_beads, int32_t n_replicas, int32_t device, gc_ctx **out)
{
    if (!out || n_beads < 0 || n_replicas < 1) return GC_ERR_INVALID;
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) {
        // Fail loudly: this library has no CPU path.
        fprintf(stderr, "gc_b200: no CUDA device available (%s)\n",
                e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
        return GC_ERR_CUDA;
    }
    if (device < 0 || device >= count) return GC_ERR_INVALID;

    gc_ctx *ctx = new gc_ctx;
    ctx->device = device;
    ctx->n = n_beads;
    ctx->n_rep = n_replicas;
    *out = ctx;   // returned even on failure so the caller can read the error
    GC_CUDA(ctx, cudaSetDevice(device));
    GC_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));

    Params &P = ctx->hp;
    memset(&P, 0, sizeof P);
    P.n = n_beads;
    P.n_rep = n_replicas;
    {
        long long want = std::max<long long>(8LL * n_beads, 32768);
        want = std::min<long long>(want, 1LL << 23);
        want = (want + 3) & ~3LL;                // k_scan reads the counts as int4
        P.ncell_max = (int) want;
    }
    // cell edge = list radius.  (Half-size cells, 125 instead of 27 per neighbourhood, were
    // measured 3x slower -- 408 vs 137 us per pair launch, profiles/r2_notes.md -- and the knob is gone.)
    P.cells_per_cut = 1;
    {
        int sms = 148;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
        // resident blocks per SM of k_pair_cells (60 registers, 26 KB shared memory:
        // __launch_bounds__(128, GC_CELL_MINBLOCKS)); measured best at 8, profiles/r1_notes.md
        int per_sm = GC_CELL_MINBLOCKS;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_pair_cells<false, GC_CELL_MINBLOCKS>, CW * 32, 0);
        if (const char *env = getenv("GC_TILE_BLOCKS_PER_SM")) per_sm = std::max(1, atoi(env));
        ctx->tile_grid_x = sms * per_sm;        // 1-D: blocks serve all replicas
        ctx->pdl = getenv("GC_NO_PDL") == nullptr;
        // replica groups (see gc_ctx): up to 8 branches; their pair kernels share the SMs,
        // ~16 blocks per SM in total (measured: profiles/r1_notes.md)
        int groups = std::min(n_replicas, 8);
        if (const char *env = getenv("GC_GROUPS")) groups = std::max(1, std::min(atoi(env), n_replicas));
        int per_group = std::max(1, std::min(per_sm, 16 / std::max(groups, 1)));
        if (const char *env = getenv("GC_GROUP_BLOCKS_PER_SM")) per_group = std::max(1, atoi(env));
        ctx->groups = std::max(groups, 1);
        ctx->group_grid_x = ctx->groups > 1 ? sms * per_group : ctx->tile_grid_x;
        if (const char *env = getenv("GC_GROUP_GRID")) ctx->group_grid_x = std::max(1, atoi(env));
        ctx->small = n_beads <= SMALL_MAX_BEADS && P.ncell_max <= SMALL_MAX_CELLS &&
                     getenv("GC_NO_SMALL") == nullptr;
    }
    P.skin = 0.0f;
    P.kT = 1.0f;
    P.dt = 1.0f;
    P.pair.mode = PAIR_NONE;
    P.well.term = -1;
    P.wall.term = -1;

    size_t RN = (size_t) n_replicas * (size_t) n_beads;
    int rc = 0;
    for (int k = 0; k < 2 && !rc; k++) {
        rc = dev_alloc(ctx, &P.pos[k], RN);
        if (!rc) rc = dev_alloc(ctx, &P.spos[k], RN);
    }
    if (!rc) rc = dev_alloc(ctx, &P.sref, RN);
    if (!rc) rc = dev_alloc(ctx, &P.scell, RN);
    if (!rc) rc = dev_alloc(ctx, &P.sab, RN);
    P.max_layers = std::min(1024, std::max(64, n_beads / 4));
    P.n_units = (n_beads + 1) / 2 + 1 + P.max_layers;
    P.unit_plane = (size_t) n_replicas * (size_t) P.n_units;
    if (!rc) rc = dev_alloc(ctx, &P.punit, 3 * P.unit_plane);
    if (!rc) rc = dev_alloc(ctx, &P.members, RN);
    if (!rc) rc = dev_alloc(ctx, &P.cell_of, RN);
    if (!rc) rc = dev_alloc(ctx, &P.cell_list, RN);
    P.scan_blocks = (P.ncell_max + SCAN_TILE - 1) / SCAN_TILE;
    if (!rc) rc = dev_alloc(ctx, &P.scan_status, (size_t) n_replicas * P.scan_blocks);
    if (!rc) rc = dev_alloc(ctx, &P.force, RN);
    if (!rc) rc = dev_alloc(ctx, &P.pforce, RN);
    if (!rc) rc = dev_alloc(ctx, &P.cell_count, (size_t) n_replicas * P.ncell_max);
    if (!rc) rc = dev_alloc(ctx, &P.cell_start, (size_t) n_replicas * (P.ncell_max + 1));
    if (!rc) rc = dev_alloc(ctx, &P.st, (size_t) n_replicas);
    if (!rc) rc = dev_alloc(ctx, &ctx->d_partials, (size_t) (n_beads + BS - 1) / BS + 1);
    if (!rc) rc = dev_alloc(ctx, &ctx->d_counter, 4);
    if (!rc) rc = dev_alloc(ctx, &ctx->d_mob, (size_t) n_beads);
    if (!rc) rc = dev_alloc(ctx, &ctx->d_ab, (size_t) n_beads);
    if (rc) return rc;

    // initial per-replica state
    std::vector<RepState> st((size_t) n_replicas);
    for (int r = 0; r < n_replicas; r++) {
        memset(&st[r], 0, sizeof(RepState));
        st[r].core_scale = 1.0;
        st[r].bond_scale = 1.0;
        st[r].semiaxes[0] = st[r].semiaxes[1] = st[r].semiaxes[2] = 1.0;
        st[r].seed = 0x9E3779B97F4A7C15ull * (unsigned long long) (r + 1);
        st[r].need_rebuild = 1;
    }
    GC_CUDA(ctx, cudaMemcpyAsync(P.st, st.data(), st.size() * sizeof(RepState),
                                 cudaMemcpyHostToDevice, ctx->stream));
    GC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return GC_OK;
}

int gc_destroy(gc_ctx *ctx)
{
    if (!ctx) return GC_ERR_INVALID;
    cudaSetDevice(ctx->device);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    for (auto &k : ctx->prof) {
        for (cudaEvent_t e : k.ev) cudaEventDestroy(e);
    }
    for (void *m : ctx->ipc_mapped) cudaIpcCloseMemHandle(m);
    for (void *p : ctx->allocs) cudaFree(p);
    if (ctx->copy_stream) { cudaStreamSynchronize(ctx->copy_stream); cudaStreamDestroy(ctx->copy_stream); }
    if (ctx->copy_event) cudaEventDestroy(ctx->copy_event);
    if (ctx->h_stage) cudaFreeHost(ctx->h_stage);
    if (ctx->h_frame) cudaFreeHost(ctx->h_frame);
    if (ctx->step_graph) cudaGraphExecDestroy(ctx->step_graph);
    for (cudaStream_t gs : ctx->group_streams) cudaStreamDestroy(gs);
    for (cudaEvent_t ge : ctx->group_events) cudaEventDestroy(ge);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
    return GC_OK;
}

// ---------------------------------------------------------------------------
// beads
// ---------------------------------------------------------------------------

int gc_set_bead_attrs(gc_ctx *ctx, const double *mobility, const double *a_factor,
                      const double *b_factor)
{
    GC_CHECK_CTX(ctx);
    if (!mobility) return fail(ctx, GC_ERR_INVALID, "mobility is required");
    std::vector<float> mob((size_t) ctx->n);
    std::vector<float2> ab((size_t) ctx->n);
    for (int i = 0; i < ctx->n; i++) {
        mob[i] = (float) mobility[i];
        ab[i] = make_float2(a_factor ? (float) a_factor[i] : 0.0f,
                            b_factor ? (float) b_factor[i] : 0.0f);
    }
    GC_CUDA(ctx, cudaMemcpyAsync(ctx->d_mob, mob.data(), mob.size() * sizeof(float),
                                 cudaMemcpyHostToDevice, ctx->stream));
    GC_CUDA(ctx, cudaMemcpyAsync(ctx->d_ab, ab.data(), ab.size() * sizeof(float2),
                                 cudaMemcpyHostToDevice, ctx->stream));
    GC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->hp.mob = ctx->d_mob;
    ctx->hp.ab = ctx->d_ab;
    ctx->attrs_set = true;
    ctx->params_dirty = true;
    return GC_OK;
}

int gc_set_positions_f32(gc_ctx *ctx, int32_t replica, const float *xyz)
{
    GC_CHECK_CTX(ctx);
    if (!xyz || replica < -1 || replica >= ctx->n_rep) return fail(ctx, GC_ERR_INVALID, "bad args");
    std::vector<float4> buf((size_t) ctx->n);
    for (int i = 0; i < ctx->n; i++) {
        buf[i] = make_float4(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2], 0.0f);
    }
    int r0 = replica < 0 ? 0 : replica, r1 = replica < 0 ? ctx->n_rep : replica + 1;
    for (int r = r0; r < r1; r++) {
        GC_CUDA(ctx, cudaMemcpyAsync(ctx->hp.pos[ctx->parity] + (size_t) r * ctx->n, buf.data(),
                                     buf.size() * sizeof(float4), cudaMemcpyHostToDevice,
                                     ctx->stream));
    }
    GC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return GC_OK;
}

int gc_set_positions(gc_ctx *ctx, int32_t replica, const double *xyz)
{
    if (!ctx || !xyz) return GC_ERR_INVALID;
    std::vector<float> f((size_t) ctx->n * 3);
    for (size_t k = 0; k < f.size(); k++) f[k] = (float) xyz[k];
    return gc_set_positions_f32(ctx, replica, f.data());
}

int gc_get_positions_f32(gc_ctx *ctx, int32_t replica, float *xyz)
{
    GC_CHECK_CTX(ctx);
    if (!xyz || replica < 0 || replica >= ctx->n_rep) return fail(ctx, GC_ERR_INVALID, "bad args");
    std::vector<float4> buf((size_t) ctx->n);
    GC_CUDA(ctx, cudaMemcpyAsync(buf.data(), ctx->hp.pos[ctx->parity] + (size_t) replica * ctx->n,
                                 buf.size() * sizeof(float4), cudaMemcpyDeviceToHost, ctx->stream));
    GC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    for (int i = 0; i < ctx->n; i++) {
        xyz[3 * i] = buf[i].x;
        xyz[3 * i + 1] = buf[i].y;
        xyz[3 * i + 2] = buf[i].z;
    }
    return GC_OK;
}

int gc_get_positions(gc_ctx *ctx, int32_t replica, double *xyz)
{
    if (!ctx || !xyz) return GC_ERR_INVALID;
    std::vector<float> f((size_t) ctx->n * 3);
    int rc = gc_get_positions_f32(ctx, replica, f.data());
    if (rc) return rc;
    for (size_t k = 0; k < f.size(); k++) xyz[k] = (double) f[k];
    return GC_OK;
}

} // extern "C"

namespace {

// read-modify-write of the per-replica state from the host
template<typename Fn>
int modify_states(gc_ctx *ctx, int replica, Fn fn)
{
    if (replica < -1 || replica >= ctx->n_rep) return fail(ctx, GC_ERR_INVALID, "bad replica");
    std::vector<RepState> st((size_t) ctx->n_rep);
    GC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    GC_CUDA(ctx, cudaMemcpy(st.data(), ctx->hp.st, st.size() * sizeof(RepState),
                            cudaMemcpyDeviceToHost));
    int r0 = replica < 0 ? 0 : replica, r1 = replica < 0 ? ctx->n_rep : replica + 1;
    for (int r = r0; r < r1; r++) fn(st[r], r);
    GC_CUDA(ctx, cudaMemcpy(ctx->hp.st, st.data(), st.size() * sizeof(RepState),
                            cudaMemcpyHostToDevice));
    return GC_OK;
}

} // namespace

extern "C" {

int gc_set_seeds(gc_ctx *ctx, const uint64_t *seeds)
{
    GC_CHECK_CTX(ctx);
    if (!seeds) return fail(ctx, GC_ERR_INVALID, "seeds is null");
    return modify_states(ctx, -1, [&](RepState &st, int r) { st.seed = seeds[r]; st.epoch = 0; });
}

// ---------------------------------------------------------------------------
// terms
// ---------------------------------------------------------------------------

int gc_add_pair_softcore_ab(gc_ctx *ctx, double eps_a, double sigma_a, double eps_b,
                            double sigma_b, int32_t scale_by_core)
{
    GC_CHECK_CTX(ctx);
    if (ctx->hp.pair.mode != PAIR_NONE) return fail(ctx, GC_ERR_STATE, "pair term already set");
    if (!(sigma_a > 0) || !(sigma_b > 0)) return fail(ctx, GC_ERR_INVALID, "bad diameters");
    int t = new_term(ctx);
    if (t < 0) return t;
    ctx->hp.pair = {PAIR_AB, t, scale_by_core, 0, eps_a, sigma_a, eps_b, sigma_b};
    return t;
}

int gc_add_pair_softcore(gc_ctx *ctx, double eps, double sigma)
{
    GC_CHECK_CTX(ctx);
    if (ctx->hp.pair.mode != PAIR_NONE) return fail(ctx, GC_ERR_STATE, "pair term already set");
    if (!(sigma > 0)) return fail(ctx, GC_ERR_INVALID, "bad diameter");
    int t = new_term(ctx);
    if (t < 0) return t;
    ctx->hp.pair = {PAIR_PLAIN, t, 0, 0, eps, sigma, 0.0, sigma};
    return t;
}

int gc_add_pair_softwell(gc_ctx *ctx, double eps, double decay, double cutoff,
                         int32_t n_targets, const int32_t *targets)
{
    GC_CHECK_CTX(ctx);
    if (ctx->hp.well.term >= 0) return fail(ctx, GC_ERR_STATE, "droplet term already set");
    if (n_targets < 0 || (n_targets && !targets)) return fail(ctx, GC_ERR_INVALID, "bad targets");
    for (int k = 0; k < n_targets; k++) {
        if (targets[k] < 0 || targets[k] >= ctx->n) return fail(ctx, GC_ERR_INVALID, "bad target");
    }
    int t = new_term(ctx);
    if (t < 0) return t;
    ctx->well_targets.assign(targets, targets + n_targets);
    ctx->hp.well.term = t;
    ctx->hp.well.n_targets = n_targets;
    ctx->hp.well.eps = eps;
    ctx->hp.well.decay = decay;
    ctx->hp.well.cutoff = cutoff;
    ctx->topo_dirty = true;
    return t;
}

int gc_add_bonds(gc_ctx *ctx, const gc_bond_params *params, int32_t n, const int32_t *pairs)
{
    GC_CHECK_CTX(ctx);
    if (!params || n < 0 || (n && !pairs)) return fail(ctx, GC_ERR_INVALID, "bad args");
    if (params->kind < 0 || params->kind > 2) return fail(ctx, GC_ERR_INVALID, "bad bond kind");
    if (ctx->hp.n_bond_terms >= GC_MAX_BOND_TERMS) return fail(ctx, GC_ERR_INVALID, "too many bond terms");
    for (int k = 0; k < 2 * n; k++) {
        if (pairs[k] < 0 || pairs[k] >= ctx->n) return fail(ctx, GC_ERR_INVALID, "bond index out of range");
    }
    int t = new_term(ctx);
    if (t < 0) return t;
    int b = ctx->hp.n_bond_terms++;
    ctx->hp.bonds[b] = {params->kind, params->mix_ab, params->scale, t,
                        params->K_a, params->K_b, params->l_a, params->l_b};
    for (int k = 0; k < n; k++) ctx->bonds.push_back({pairs[2 * k], pairs[2 * k + 1], b});
    ctx->topo_dirty = true;
    return t;
}

int gc_add_triples(gc_ctx *ctx, double bending_energy, int32_t n, const int32_t *triples)
{
    GC_CHECK_CTX(ctx);
    if (n < 0 || (n && !triples)) return fail(ctx, GC_ERR_INVALID, "bad args");
    if (ctx->hp.n_triple_terms >= GC_MAX_TRIPLE_TERMS) return fail(ctx, GC_ERR_INVALID, "too many triple terms");
    for (int k = 0; k < 3 * n; k++) {
        if (triples[k] < 0 || triples[k] >= ctx->n) return fail(ctx, GC_ERR_INVALID, "triple index out of range");
    }
    int t = new_term(ctx);
    if (t < 0) return t;
    int tt = ctx->hp.n_triple_terms++;
    ctx->hp.triples[tt] = {t, 0, bending_energy};
    for (int k = 0; k < n; k++) {
        ctx->triples.push_back({triples[3 * k], triples[3 * k + 1], triples[3 * k + 2], tt});
    }
    ctx->topo_dirty = true;
    return t;
}

int gc_set_wall(gc_ctx *ctx, const gc_wall_params *w)
{
    GC_CHECK_CTX(ctx);
    if (!w) return fail(ctx, GC_ERR_INVALID, "null params");
    if (ctx->hp.wall.term >= 0) return fail(ctx, GC_ERR_STATE, "wall already set");
    for (int k = 0; k < 3; k++) {
        if (!(w->semiaxes[k] > 0)) return fail(ctx, GC_ERR_INVALID, "bad semiaxes");
    }
    int t = new_term(ctx);
    if (t < 0) return t;
    WallParams &W = ctx->hp.wall;
    W.term = t;
    W.scale_by_core = w->scale_by_core;
    W.ode_enabled = w->ode_enabled;
    W.eps_a = w->eps_a; W.sigma_a = w->sigma_a; W.eps_b = w->eps_b; W.sigma_b = w->sigma_b;
    W.wall_a = w->wall_a; W.wall_b = w->wall_b;
    W.outward_spring = w->outward_spring;
    W.ode_mobility = w->ode_mobility;
    for (int k = 0; k < 3; k++) W.ode_spring[k] = w->ode_spring[k];
    int rc = modify_states(ctx, -1, [&](RepState &st, int) {
        for (int k = 0; k < 3; k++) st.semiaxes[k] = w->semiaxes[k];
    });
    return rc ? rc : t;
}

int gc_add_point_source(gc_ctx *ctx, int32_t kind, const double point[3], double p0, double p1)
{
    GC_CHECK_CTX(ctx);
    if (!point || kind < 0 || kind > 1) return fail(ctx, GC_ERR_INVALID, "bad args");
    if (ctx->hp.n_sources >= GC_MAX_SOURCES) return fail(ctx, GC_ERR_INVALID, "too many sources");
    int t = new_term(ctx);
    if (t < 0) return t;
    SourceTerm &S = ctx->hp.sources[ctx->hp.n_sources++];
    S.kind = kind;
    S.term = t;
    for (int k = 0; k < 3; k++) S.point[k] = point[k];
    S.p0 = p0;
    S.p1 = p1;
    return t;
}

int gc_add_kfiber(gc_ctx *ctx, const double pole[3], int32_t n, const int32_t *beads,
                  const double *spring, const double *length)
{
    GC_CHECK_CTX(ctx);
    if (!pole || n < 0 || (n && (!beads || !spring || !length))) return fail(ctx, GC_ERR_INVALID, "bad args");
    if (ctx->hp.n_poles >= GC_MAX_POLES) return fail(ctx, GC_ERR_INVALID, "too many poles");
    for (int k = 0; k < n; k++) {
        if (beads[k] < 0 || beads[k] >= ctx->n) return fail(ctx, GC_ERR_INVALID, "bead out of range");
    }
    int t = new_term(ctx);
    if (t < 0) return t;
    int p = ctx->hp.n_poles++;
    for (int k = 0; k < 3; k++) ctx->hp.poles[p][k] = pole[k];
    for (int k = 0; k < n; k++) {
        ctx->tethers.push_back({beads[k], p, (float) spring[k], (float) length[k], t});
    }
    ctx->topo_dirty = true;
    return t;
}

int gc_set_term_enabled(gc_ctx *ctx, int32_t term, int32_t enabled)
{
    GC_CHECK_CTX(ctx);
    if (term < 0 || term >= ctx->next_term) return fail(ctx, GC_ERR_INVALID, "unknown term");
    if (enabled) ctx->hp.enabled |= 1u << term;
    else ctx->hp.enabled &= ~(1u << term);
    ctx->params_dirty = true;
    return GC_OK;
}

// ---------------------------------------------------------------------------
// schedule / context
// ---------------------------------------------------------------------------

int gc_set_schedule(gc_ctx *ctx, const gc_schedule *s)
{
    GC_CHECK_CTX(ctx);
    if (!s) return fail(ctx, GC_ERR_INVALID, "null schedule");
    ctx->hp.sched = {s->enabled, 0, s->core_init, s->core_tau, s->bond_init, s->bond_tau};
    ctx->params_dirty = true;
    return GC_OK;
}

int gc_set_context(gc_ctx *ctx, int32_t replica, const gc_context *in)
{
    GC_CHECK_CTX(ctx);
    if (!in) return fail(ctx, GC_ERR_INVALID, "null context");
    return modify_states(ctx, replica, [&](RepState &st, int) {
        st.time = in->time;
        st.core_scale = in->core_scale;
        st.bond_scale = in->bond_scale;
        st.contact_distance = in->contact_distance;
        st.contact_distance_prev = in->contact_distance;
        for (int k = 0; k < 3; k++) st.semiaxes[k] = in->semiaxes[k];
    });
}

int gc_get_context(gc_ctx *ctx, int32_t replica, gc_context *out)
{
    GC_CHECK_CTX(ctx);
    if (!out || replica < 0 || replica >= ctx->n_rep) return fail(ctx, GC_ERR_INVALID, "bad args");
    RepState st;
    GC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    GC_CUDA(ctx, cudaMemcpy(&st, ctx->hp.st + replica, sizeof st, cudaMemcpyDeviceToHost));
    out->time = st.time;
    out->core_scale = st.core_scale;
    out->bond_scale = st.bond_scale;
    out->contact_distance = st.contact_distance;
    out->reaction_force = st.reaction;
    out->last_timestep = st.dt_last;
    out->step = st.step;
    out->rebuilds = st.rebuilds;
    for (int k = 0; k < 3; k++) {
        out->semiaxes[k] = st.semiaxes[k];
        out->axial_reaction[k] = st.axial[k];
    }
    return GC_OK;
}

int gc_set_step(gc_ctx *ctx, int32_t replica, int64_t step)
{
    GC_CHECK_CTX(ctx);
    if (step < 0) return fail(ctx, GC_ERR_INVALID, "step must be >= 0");
    int rc = modify_states(ctx, replica, [&](RepState &st, int) {
        st.step = step;
        st.time = (double) step * ctx->hp.dt_cfg;
    });
    if (rc) return rc;
    ctx->run_step = step;      // host mirror (contact cadence, graph alignment): replicas move in lockstep
    // a decomposed context marks "current on this rank" with the step number: resuming at
    // another step (after gc_set_positions with all N beads) re-stamps them all
    if (ctx->hp.decomp) {
        k_decomp_stamp_all<<<(ctx->n + 255) / 256, 256, 0, ctx->stream>>>(ctx->hp);
        ctx->decomp_all_current = true;
    }
    return GC_OK;
}

int gc_get_contexts(gc_ctx *ctx, gc_context *out)
{
    GC_CHECK_CTX(ctx);
    if (!out) return fail(ctx, GC_ERR_INVALID, "bad args");
    std::vector<RepState> st((size_t) ctx->n_rep);
    GC_CUDA(ctx, cudaMemcpyAsync(st.data(), ctx->hp.st, st.size() * sizeof(RepState),
                                 cudaMemcpyDeviceToHost, ctx->stream));
    GC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    for (int r = 0; r < ctx->n_rep; r++) {
        gc_context &o = out[r];
        const RepState &s = st[r];
        o.time = s.time;
        o.core_scale = s.core_scale;
        o.bond_scale = s.bond_scale;
        o.contact_distance = s.contact_distance;
        o.reaction_force = s.reaction;
        o.last_timestep = s.dt_last;
        o.step = s.step;
        o.rebuilds = s.rebuilds;
        for (int k = 0; k < 3; k++) {
            o.semiaxes[k] = s.semiaxes[k];
            o.axial_reaction[k] = s.axial[k];
        }
    }
    return GC_OK;
}

int gc_set_skin(gc_ctx *ctx, double skin_fraction)
{
    GC_CHECK_CTX(ctx);
    if (!(skin_fraction >= 0.0)) return fail(ctx, GC_ERR_INVALID, "skin must be >= 0");
    ctx->hp.skin = (float) skin_fraction;
    ctx->params_dirty = true;
    return GC_OK;
}

} // extern "C"

// ---------------------------------------------------------------------------
// launch plumbing
// ---------------------------------------------------------------------------
namespace {

constexpr size_t PROF_MAX_EVENTS = 40000;

void prof_begin(gc_ctx *ctx, const char *name)
{
    ctx->prof_open = -1;
    if (!ctx->profiling || ctx->prof_events + 2 > PROF_MAX_EVENTS) return;
    int k = 0;
    for (; k < (int) ctx->prof.size(); k++) {
        if (ctx->prof[k].name == name) break;
    }
    if (k == (int) ctx->prof.size()) ctx->prof.push_back({name, {}});
    cudaEvent_t a = nullptr, b = nullptr;
    if (cudaEventCreate(&a) != cudaSuccess || cudaEventCreate(&b) != cudaSuccess) return;
    ctx->prof[k].ev.push_back(a);
    ctx->prof[k].ev.push_back(b);
    ctx->prof_events += 2;
    ctx->prof_open = k;
    cudaEventRecord(a, ctx->stream);
}

void prof_end(gc_ctx *ctx)
{
    if (ctx->prof_open < 0) return;
    cudaEventRecord(ctx->prof[ctx->prof_open].ev.back(), ctx->stream);
    ctx->prof_open = -1;
}

void prof_clear(gc_ctx *ctx)
{
    for (auto &k : ctx->prof) {
        for (cudaEvent_t e : k.ev) cudaEventDestroy(e);
    }
    ctx->prof.clear();
    ctx->prof_events = 0;
}

// Every launch carries the programmatic-stream-serialization attribute (unless
// GC_NO_PDL is set): the kernels open with GC_PDL_ENTRY, so kernel N + 1 is scheduled
// while kernel N drains and the kernel-to-kernel gap (7 boundaries per step) shrinks.
template<typename... KArgs, typename... Args>
inline void launch_k(gc_ctx *ctx, void (*kernel)(KArgs...), dim3 grid, dim3 block, Args &&...args)
{
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = 0;
    cfg.stream = ctx->stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = ctx->pdl ? 1 : 0;
    cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}

#define LAUNCH(ctx, kernel, grid, block, ...)                                                 \
    do {                                                                                      \
        prof_begin(ctx, #kernel);                                                             \
        launch_k(ctx, kernel, dim3(grid), dim3(block), __VA_ARGS__);                          \
        prof_end(ctx);                                                                        \
        (ctx)->launches++;                                                                    \
    } while (0)

// Per-bead adjacency in CSR form + parameter upload.
int finalize(gc_ctx *ctx)
{
    if (!ctx->attrs_set) return fail(ctx, GC_ERR_STATE, "gc_set_bead_attrs has not been called");
    Params &P = ctx->hp;
    const int n = ctx->n;
    if (ctx->topo_dirty) {
        std::vector<int> flags((size_t) n, 0);
        std::vector<int> boff((size_t) n + 1, 0), toff((size_t) n + 1, 0), koff((size_t) n + 1, 0);
        // Harmonic bonds with K == 0 exert exactly zero force and energy (the
        // reference's loop bonds, SURVEY 8a note on a12): keep the term id but do
        // not iterate them on the device.
        std::vector<HostBond> live;
        live.reserve(ctx->bonds.size());
        for (auto &b : ctx->bonds) {
            const BondTerm &bt = P.bonds[b.bterm];
            if (bt.kind == GC_BOND_HARMONIC && bt.K_a == 0.0 && bt.K_b == 0.0) continue;
            live.push_back(b);
        }
        for (auto &b : live) { boff[b.i + 1]++; boff[b.j + 1]++; }
        for (auto &t : ctx->triples) { toff[t.i + 1]++; toff[t.j + 1]++; toff[t.k + 1]++; }
        for (auto &k : ctx->tethers) koff[k.bead + 1]++;
        for (int i = 0; i < n; i++) {
            boff[i + 1] += boff[i];
            toff[i + 1] += toff[i];
            koff[i + 1] += koff[i];
        }
        std::vector<BondEntry> bent((size_t) boff[n]);
        std::vector<TripleEntry> tent((size_t) toff[n]);
        std::vector<TetherEntry> kent((size_t) koff[n]);
        {
            std::vector<int> cur(boff.begin(), boff.end() - 1);
            for (auto &b : live) {
                bent[cur[b.i]++] = {b.j, b.bterm};
                bent[cur[b.j]++] = {b.i, b.bterm};
            }
        }
        {
            std::vector<int> cur(toff.begin(), toff.end() - 1);
            for (auto &t : ctx->triples) {
                int members[3] = {t.i, t.j, t.k};
                for (int role = 0; role < 3; role++) {
                    tent[cur[members[role]]++] = {t.i, t.j, t.k, role | (t.tterm << 2)};
                    flags[members[role]] |= 1;
                }
            }
        }
        {
            std::vector<int> cur(koff.begin(), koff.end() - 1);
            for (auto &k : ctx->tethers) {
                kent[cur[k.bead]++] = {k.pole, k.spring, k.length, k.term};
                flags[k.bead] |= 2;
            }
        }
        for (int j : ctx->well_targets) flags[j] |= 4;

        int rc = 0;
        if (!rc) rc = upload(ctx, &ctx->d_flags, flags);
        if (!rc) rc = upload(ctx, &ctx->d_bond_off, boff);
        if (!rc) rc = upload(ctx, &ctx->d_triple_off, toff);
        if (!rc) rc = upload(ctx, &ctx->d_tether_off, koff);
        if (!rc) rc = upload(ctx, &ctx->d_bond_ent, bent);
        if (!rc) rc = upload(ctx, &ctx->d_triple_ent, tent);
        if (!rc) rc = upload(ctx, &ctx->d_tether_ent, kent);
        if (!rc) rc = upload(ctx, &ctx->d_well_targets, ctx->well_targets);
        if (rc) return rc;
        P.flags = ctx->d_flags;
        P.bond_off = ctx->d_bond_off;
        P.bond_ent = ctx->d_bond_ent;
        P.triple_off = ctx->d_triple_off;
        P.triple_ent = ctx->d_triple_ent;
        P.tether_off = ctx->d_tether_off;
        P.tether_ent = ctx->d_tether_ent;
        P.well.targets = ctx->d_well_targets;
        ctx->topo_dirty = false;
        ctx->params_dirty = true;
    }
    ctx->params_dirty = false;
    return GC_OK;
}

// bounding box + grid plan + forced rebuild of every replica's cell list
int plan_and_rebuild(gc_ctx *ctx, double radius_override)
{
    const int R = ctx->n_rep;
    LAUNCH(ctx, k_reset_reductions, (R + 63) / 64, 64, ctx->hp);
    int bx = std::min((ctx->n + 255) / 256, 1024);
    LAUNCH(ctx, k_bbox, dim3(std::max(bx, 1), R), 256, ctx->hp);
    LAUNCH(ctx, k_plan, (R + 63) / 64, 64, ctx->hp, radius_override);
    return GC_OK;
}

// small path: one block per replica instead of the four kernels
inline bool small_path(const gc_ctx *ctx) { return ctx->small && !ctx->hp.decomp; }

void launch_rebuild(gc_ctx *ctx)
{
    if (small_path(ctx)) {
        LAUNCH(ctx, k_rebuild_small, ctx->n_rep, SMALL_BS, ctx->hp);
        return;
    }
    dim3 g = bead_grid(ctx, 256);
    // a decomposed run strides over the beads current on the rank
    if (ctx->hp.decomp) g.x = std::min<unsigned>(g.x, 148 * 8);
    LAUNCH(ctx, k_count, g, 256, ctx->hp);
    LAUNCH(ctx, k_scan, dim3(ctx->hp.scan_blocks, ctx->n_rep), SCAN_BS, ctx->hp);
    LAUNCH(ctx, k_fill, g, 256, ctx->hp);
    LAUNCH(ctx, k_gather, g, 256, ctx->hp);
}

int check_device_errors(gc_ctx *ctx)
{
    GC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    GC_CUDA(ctx, cudaGetLastError());
    std::vector<RepState> st((size_t) ctx->n_rep);
    GC_CUDA(ctx, cudaMemcpy(st.data(), ctx->hp.st, st.size() * sizeof(RepState),
                            cudaMemcpyDeviceToHost));
    for (int r = 0; r < ctx->n_rep; r++) {
        if (st[r].error == 1) {
            return fail(ctx, GC_ERR_NUMERIC, "replica %d: non-finite position or force at step %lld",
                        r, st[r].step);
        }
        if (st[r].error == 2) {
            return fail(ctx, GC_ERR_CAPACITY, "replica %d: contact table full (capacity %llu)", r,
                        (unsigned long long) ctx->hp.contact.capacity);
        }
        if (st[r].error == 3 || (st[r].error >= 30 && st[r].error < 40)) {
            const char *what = st[r].error == 30 ? "a bead a neighbour needs under the new grid was not in the halo "
                                                   "shipped under the old one (peer mode)"
                             : st[r].error == 32 ? "a bead moved past a neighbour's whole slab"
                             : st[r].error == 33 ? "a slab is thinner than the halo"
                             : st[r].error == 34 ? "a bonded partner is not current on the rank that needs it"
                                                 : "halo violated";
            return fail(ctx, GC_ERR_STATE, "decomposed run, step %lld, %d-layer halo: %s "
                        "(gc_decomp_set_halo, or fewer ranks)", st[r].step, ctx->hp.dhalo, what);
        }
        if (st[r].error == 4) {
            return fail(ctx, GC_ERR_CAPACITY, "decomposed run, step %lld: halo buffer full (%d beads; "
                        "gc_decomp_set_halo), a rank owns more than gc_decomp_set_owned_bound allows, "
                        "or the ranks' bead counts do not add up", st[r].step,
                        ctx->hp.dcap);
        }
    }
    return GC_OK;
}

__global__ void k_restart(const __grid_constant__ Params P, double dt)
{
    GC_PDL_ENTRY();
    int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= P.n_rep) return;
    RepState &st = P.st[r];
    st.step = 0;
    st.dt_last = dt;
    st.dt_prev = dt;
    st.epoch += 1;        // a new logical run draws a new noise stream
}

// the pair kernel in the shape this launch needs (names kept short: the LAUNCH macro
// stringifies them for the profiler)
void launch_pair(gc_ctx *ctx, unsigned mask, int use_enabled)
{
    constexpr auto k_pair_cells_plain = k_pair_cells<true, GC_CELL_MINBLOCKS>;
    constexpr auto k_pair_cells_ab = k_pair_cells<false, GC_CELL_MINBLOCKS>;
    constexpr auto k_pair_cells_plain_dense = k_pair_cells<true, MINBLOCKS_DENSE>;
    constexpr auto k_pair_cells_ab_dense = k_pair_cells<false, MINBLOCKS_DENSE>;
    const bool plain = ctx->hp.pair.mode == PAIR_PLAIN;
    if (plain && ctx->dense_pair) LAUNCH(ctx, k_pair_cells_plain_dense, ctx->tile_grid_x, CW * 32, ctx->hp, mask, use_enabled);
    else if (plain) LAUNCH(ctx, k_pair_cells_plain, ctx->tile_grid_x, CW * 32, ctx->hp, mask, use_enabled);
    else if (ctx->dense_pair) LAUNCH(ctx, k_pair_cells_ab_dense, ctx->tile_grid_x, CW * 32, ctx->hp, mask, use_enabled);
    else LAUNCH(ctx, k_pair_cells_ab, ctx->tile_grid_x, CW * 32, ctx->hp, mask, use_enabled);
}

// one force evaluation (+ update when fused) for every replica
// fuse_rebuild (fixed dt, small path): the epilogue also builds the NEXT step's cell list
void launch_step(gc_ctx *ctx, int mode, unsigned mask, int use_enabled, bool fuse_rebuild = false)
{
    // (the macro stringifies its kernel argument for the profiler: keep the names short)
    launch_pair(ctx, mask, use_enabled);
    dim3 gf = bead_grid(ctx, BS);
    const int gr = ctx->n_rep;
    if (mode == STEP_FUSED) {
        LAUNCH(ctx, k_finish<STEP_FUSED>, gf, BS, ctx->hp, mask, use_enabled);
        if (fuse_rebuild) {
            LAUNCH(ctx, k_finalize_rebuild, gr, SMALL_BS, ctx->hp);
        } else {
            LAUNCH(ctx, k_finalize_step, gr, 32, ctx->hp, 0);
        }
    } else {
        LAUNCH(ctx, k_finish<STEP_FORCE>, gf, BS, ctx->hp, mask, use_enabled);
        LAUNCH(ctx, k_finalize_force, gr, 32, ctx->hp);
    }
}

void launch_contact_update(gc_ctx *ctx)
{
    LAUNCH(ctx, k_contact_cells, ctx->tile_grid_x, CW * 32, ctx->hp, (int) RADIUS_CONTACT_PREV);
}

// The kernel arguments of replicas [r0, r0 + nr): every per-replica array starts at r0, so
// the kernels (which index replicas from 0) run unchanged on the sub-range.
Params params_window(const Params &P, int r0, int nr)
{
    Params W = P;
    const size_t o = (size_t) r0 * (size_t) P.n;
    W.n_rep = nr;
    for (int k = 0; k < 2; k++) { W.pos[k] += o; W.spos[k] += o; }
    W.sref += o;
    W.scell += o;
    W.sab += o;
    W.punit += (size_t) r0 * (size_t) P.n_units;      // the plane stride stays
    W.members += o;
    W.cell_of += o;
    W.cell_list += o;
    W.force += o;
    W.pforce += o;
    W.cell_count += (size_t) r0 * P.ncell_max;
    W.cell_start += (size_t) r0 * (P.ncell_max + 1);
    W.scan_status += (size_t) r0 * P.scan_blocks;
    W.st += r0;
    if (W.noise) W.noise += (size_t) r0 * (size_t) P.noise_steps * (size_t) P.n * 3;
    if (W.contact.bucket) W.contact.bucket += o * (size_t) P.contact.bucket_slots;
    if (W.contact.keys) W.contact.keys += (size_t) r0 * P.contact.capacity;
    if (W.contact.vals) W.contact.vals += (size_t) r0 * P.contact.capacity;
    if (W.contact.used) W.contact.used += r0;
    return W;
}

bool ensure_group_streams(gc_ctx *ctx)
{
    const int G = ctx->groups;
    while ((int) ctx->group_streams.size() < G) {
        cudaStream_t s = nullptr;
        if (cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking) != cudaSuccess) return false;
        ctx->group_streams.push_back(s);
    }
    while ((int) ctx->group_events.size() < G + 1) {
        cudaEvent_t e = nullptr;
        if (cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess) return false;
        ctx->group_events.push_back(e);
    }
    return true;
}

void drop_step_graph(gc_ctx *ctx)
{
    if (ctx->step_graph) cudaGraphExecDestroy(ctx->step_graph);
    ctx->step_graph = nullptr;
    ctx->step_graph_steps = 0;
}

// The fixed-timestep loop launches the same 7 kernels with the same by-value Params
// every step (all per-step state lives in RepState on the device), so a block of steps
// is captured once and replayed: one graph launch instead of 7 B stream launches.  This
// matters where the step is launch-bound (one replica, the ~600-bead mitotic stages).
// With a contact cadence the block is one interval long and starts with the contact
// update of the step just finished; otherwise it is 32 steps.
bool ensure_step_graph(gc_ctx *ctx, long long interval)
{
    static const bool disabled = getenv("GC_NO_GRAPH") != nullptr;
    if (disabled || ctx->profiling) return false;
    if (interval > 64) return false;
    if (ctx->step_graph && ctx->step_graph_interval == interval &&
        memcmp(&ctx->step_graph_hp, &ctx->hp, sizeof(Params)) == 0) {
        return true;
    }
    drop_step_graph(ctx);
    const int B = interval > 0 ? (int) interval : 32;
    const long long before = ctx->launches;
    cudaGraph_t graph = nullptr;
    if (cudaStreamBeginCapture(ctx->stream, cudaStreamCaptureModeThreadLocal) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    const bool fuse = small_path(ctx);    // the cell list is current when the block starts
    auto record_block = [&]() {
        for (int k = 0; k < B; k++) {
            if (!fuse) launch_rebuild(ctx);
            if (k == 0 && interval > 0) launch_contact_update(ctx);
            launch_step(ctx, STEP_FUSED, 0u, 1, fuse);
        }
    };
    const int G = (ctx->groups > 1 && ensure_group_streams(ctx)) ? ctx->groups : 1;
    if (G > 1) {
        // fork: one branch per replica group, each on its own capture stream with the
        // kernel arguments windowed to its replicas; join back into the launch stream
        const Params full = ctx->hp;
        const int R = ctx->n_rep, grid_full = ctx->tile_grid_x;
        cudaStream_t main_stream = ctx->stream;
        cudaEventRecord(ctx->group_events[0], main_stream);
        for (int g = 0; g < G; g++) {
            const int r0 = (int) ((long long) R * g / G), r1 = (int) ((long long) R * (g + 1) / G);
            cudaStreamWaitEvent(ctx->group_streams[g], ctx->group_events[0], 0);
            ctx->stream = ctx->group_streams[g];
            ctx->hp = params_window(full, r0, r1 - r0);
            ctx->n_rep = r1 - r0;
            ctx->tile_grid_x = ctx->group_grid_x;
            ctx->dense_pair = getenv("GC_NO_DENSE_PAIR") == nullptr;
            record_block();
            cudaEventRecord(ctx->group_events[1 + g], ctx->group_streams[g]);
        }
        ctx->stream = main_stream;
        ctx->hp = full;
        ctx->n_rep = R;
        ctx->tile_grid_x = grid_full;
        ctx->dense_pair = false;
        for (int g = 0; g < G; g++) cudaStreamWaitEvent(main_stream, ctx->group_events[1 + g], 0);
    } else {
        record_block();
    }
    cudaError_t e = cudaStreamEndCapture(ctx->stream, &graph);
    ctx->step_graph_launches = ctx->launches - before;
    ctx->launches = before;
    if (e == cudaSuccess) e = cudaGraphInstantiate(&ctx->step_graph, graph, 0);
    if (graph) cudaGraphDestroy(graph);
    if (e != cudaSuccess) {
        cudaGetLastError();
        ctx->step_graph = nullptr;
        return false;
    }
    ctx->step_graph_steps = B;
    ctx->step_graph_interval = interval;
    memcpy(&ctx->step_graph_hp, &ctx->hp, sizeof(Params));
    return true;
}

} // namespace

extern "C" {

// ---------------------------------------------------------------------------
// dynamics
// ---------------------------------------------------------------------------

int gc_run(gc_ctx *ctx, const gc_run_config *c)
{
    GC_CHECK_CTX(ctx);
    if (!c || c->steps < 0 || !(c->timestep > 0) || !(c->temperature >= 0)) {
        return fail(ctx, GC_ERR_INVALID, "bad run config");
    }
    if (c->noise && !c->restart) {
        return fail(ctx, GC_ERR_INVALID, "injected noise requires restart = 1");
    }
    Params &P = ctx->hp;
    if (P.decomp && c->steps > 0) {
        return fail(ctx, GC_ERR_STATE, "decomposed context: drive steps with gc_decomp_step_*");
    }
    const bool adaptive = c->spacestep != 0.0;
    if (P.kT != (float) c->temperature || P.dt_cfg != c->timestep ||
        P.spacestep != (float) c->spacestep) {
        P.kT = (float) c->temperature;
        P.dt = (float) c->timestep;
        P.dt_cfg = c->timestep;
        P.spacestep = (float) c->spacestep;
        ctx->params_dirty = true;
    }

    // injected noise: double [R][steps][N][3] -> float on device
    const float *noise_dev = nullptr;
    if (c->noise) {
        size_t count = (size_t) ctx->n_rep * (size_t) c->steps * (size_t) ctx->n * 3;
        if (count > ctx->noise_cap) {
            if (ctx->d_noise) dev_free(ctx, ctx->d_noise);
            ctx->d_noise = nullptr;
            int rc = dev_alloc(ctx, &ctx->d_noise, count);
            if (rc) return rc;
            ctx->noise_cap = count;
        }
        std::vector<float> f(count);
        for (size_t k = 0; k < count; k++) f[k] = (float) c->noise[k];
        GC_CUDA(ctx, cudaMemcpyAsync(ctx->d_noise, f.data(), count * sizeof(float),
                                     cudaMemcpyHostToDevice, ctx->stream));
        GC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        noise_dev = ctx->d_noise;
    }
    if (P.noise != noise_dev || P.noise_steps != c->steps) {
        P.noise = noise_dev;
        P.noise_steps = c->steps;
        P.noise_step0 = 0;
        ctx->params_dirty = true;
    }

    int rc = finalize(ctx);
    if (rc) return rc;

    if (c->restart) {
        LAUNCH(ctx, k_restart, (ctx->n_rep + 63) / 64, 64, ctx->hp, c->timestep);
        ctx->run_step = 0;
    }
    if (P.decomp) {
        // restart: every rank holds all N positions (gc_set_positions), all are current.
        // Otherwise only the parameters were refreshed: the grid of the last step stands,
        // and a bounding box over this rank's partly stale positions would be wrong.
        // (unless every bead is still current: right after a restart or gc_set_step, where a
        //  changed context -- core_scale sets the cell edge -- must reach the plan)
        if (!c->restart && !ctx->decomp_all_current) return GC_OK;
        if (c->restart) LAUNCH(ctx, k_decomp_stamp_all, (ctx->n + 255) / 256, 256, ctx->hp);
        ctx->decomp_all_current = true;
    }
    plan_and_rebuild(ctx, 0.0);

    const long long interval = P.contact.enabled ? P.contact.interval : 0;
    const dim3 g = bead_grid(ctx, BS);
    const bool graphable = !adaptive && !c->noise && c->steps >= 64 && ensure_step_graph(ctx, interval);
    const int B = graphable ? ctx->step_graph_steps : 0;
    // small path, fixed dt: every step's epilogue leaves the cell list of the next step
    const bool fuse = small_path(ctx) && !adaptive;
    if (fuse && c->steps > 0) launch_rebuild(ctx);
    for (long long i = 0; i < c->steps;) {
        // a whole block, aligned to the contact cadence (the block opens with the contact
        // update of the step just finished)
        if (graphable && i > 0 && c->steps - i >= B &&
            (interval == 0 || ctx->run_step % interval == 0)) {
            GC_CUDA(ctx, cudaGraphLaunch(ctx->step_graph, ctx->stream));
            ctx->launches += ctx->step_graph_launches;
            ctx->run_step += B;
            if (B & 1) ctx->parity ^= 1;
            i += B;
            continue;
        }
        if (!fuse) launch_rebuild(ctx);
        // contact update of the driver callback for the step just finished
        // (SIM/stage_interphase/simulation_driver_interphase.cpp:38-40)
        if (interval > 0 && ctx->run_step > 0 && ctx->run_step % interval == 0 && i > 0) {
            launch_contact_update(ctx);
        }
        if (adaptive) {
            launch_step(ctx, STEP_FORCE, 0u, 1);
            LAUNCH(ctx, k_integrate, g, BS, ctx->hp);
            LAUNCH(ctx, k_finalize_step, ctx->n_rep, 32, ctx->hp, 1);
        } else {
            launch_step(ctx, STEP_FUSED, 0u, 1, fuse);
        }
        ctx->run_step++;
        ctx->parity ^= 1;
        i++;
    }
    // the callback of the last step of this chunk
    if (interval > 0 && c->steps > 0 && ctx->run_step % interval == 0) {
        if (!fuse) launch_rebuild(ctx);
        launch_contact_update(ctx);
    }
    return check_device_errors(ctx);
}

int gc_get_frame(gc_ctx *ctx, int32_t replica, int32_t fraction_bits, float *xyz)
{
    GC_CHECK_CTX(ctx);
    if (!xyz || replica < -1 || replica >= ctx->n_rep || fraction_bits < 1 || fraction_bits > 24) {
        return fail(ctx, GC_ERR_INVALID, "bad args");
    }
    if (ctx->frame_pending) return fail(ctx, GC_ERR_STATE, "a frame is in flight (gc_frame_end first)");
    const int r0 = replica < 0 ? 0 : replica, nrep = replica < 0 ? ctx->n_rep : 1;
    const size_t count = (size_t) nrep * ctx->n * 3;
    if (count == 0) return GC_OK;
    if (count > ctx->frame_cap) {
        if (ctx->d_frame) dev_free(ctx, ctx->d_frame);
        ctx->d_frame = nullptr;
        ctx->frame_cap = 0;
        int rc = dev_alloc(ctx, &ctx->d_frame, count);
        if (rc) return rc;
        ctx->frame_cap = count;
    }
    const int blocks = (int) std::min<size_t>((count / 3 + 255) / 256, 148 * 8);
    LAUNCH(ctx, k_frame, std::max(blocks, 1), 256, ctx->hp, r0, nrep, (int) fraction_bits, ctx->d_frame);
    GC_CUDA(ctx, cudaMemcpyAsync(xyz, ctx->d_frame, count * sizeof(float), cudaMemcpyDeviceToHost,
                                 ctx->stream));
    GC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return GC_OK;
}

} // extern "C"

namespace {

// page-locked host + device staging of `floats` packed f32, copy stream, event
int ensure_stage(gc_ctx *ctx, size_t floats)
{
    if (!ctx->copy_stream) {
        GC_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
        GC_CUDA(ctx, cudaEventCreateWithFlags(&ctx->copy_event, cudaEventDisableTiming));
    }
    if (floats <= ctx->stage_cap) return GC_OK;
    GC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (ctx->h_stage) cudaFreeHost(ctx->h_stage);
    if (ctx->d_stage) dev_free(ctx, ctx->d_stage);
    ctx->h_stage = nullptr;
    ctx->d_stage = nullptr;
    ctx->stage_cap = 0;
    GC_CUDA(ctx, cudaMallocHost(&ctx->h_stage, floats * sizeof(float)));
    int rc = dev_alloc(ctx, &ctx->d_stage, floats);
    if (rc) return rc;
    ctx->stage_cap = floats;
    return GC_OK;
}

// true when `p` is page-locked host memory (then copies need no staging pass)
bool is_pinned(const void *p)
{
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    return a.type == cudaMemoryTypeHost;
}

int ensemble_blocks(const gc_ctx *ctx)
{
    const size_t total = (size_t) ctx->n_rep * ctx->n;
    return (int) std::max<size_t>(1, std::min<size_t>((total + 255) / 256, 148 * 8));
}

} // namespace

extern "C" {

int gc_set_positions_all_f32(gc_ctx *ctx, const float *xyz)
{
    GC_CHECK_CTX(ctx);
    if (!xyz) return fail(ctx, GC_ERR_INVALID, "bad args");
    const size_t floats = (size_t) ctx->n_rep * ctx->n * 3;
    if (floats == 0) return GC_OK;
    int rc = ensure_stage(ctx, floats);
    if (rc) return rc;
    const float *src = xyz;
    if (!is_pinned(xyz)) {
        GC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));      // the staging buffer is free again
        memcpy(ctx->h_stage, xyz, floats * sizeof(float));
        src = ctx->h_stage;
    }
    GC_CUDA(ctx, cudaMemcpyAsync(ctx->d_stage, src, floats * sizeof(float), cudaMemcpyHostToDevice,
                                 ctx->stream));
    LAUNCH(ctx, k_unpack_positions, ensemble_blocks(ctx), 256, ctx->hp, (const float *) ctx->d_stage);
    // asynchronous: later work of this context is ordered behind it on the stream; a
    // pinned caller buffer must stay untouched until the next synchronising call
    return GC_OK;
}

int gc_get_positions_all_f32(gc_ctx *ctx, float *xyz)
{
    GC_CHECK_CTX(ctx);
    if (!xyz) return fail(ctx, GC_ERR_INVALID, "bad args");
    const size_t floats = (size_t) ctx->n_rep * ctx->n * 3;
    if (floats == 0) return GC_OK;
    int rc = ensure_stage(ctx, floats);
    if (rc) return rc;
    LAUNCH(ctx, k_pack_positions, ensemble_blocks(ctx), 256, ctx->hp, ctx->d_stage);
    const bool direct = is_pinned(xyz);
    GC_CUDA(ctx, cudaMemcpyAsync(direct ? xyz : ctx->h_stage, ctx->d_stage, floats * sizeof(float),
                                 cudaMemcpyDeviceToHost, ctx->stream));
    GC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (!direct) memcpy(xyz, ctx->h_stage, floats * sizeof(float));
    return GC_OK;
}

int gc_host_alloc(size_t bytes, void **out)
{
    if (!out) return GC_ERR_INVALID;
    *out = nullptr;
    return cudaMallocHost(out, std::max<size_t>(bytes, 1)) == cudaSuccess ? GC_OK : GC_ERR_CUDA;
}

int gc_host_free(void *p)
{
    if (!p) return GC_OK;
    return cudaFreeHost(p) == cudaSuccess ? GC_OK : GC_ERR_CUDA;
}

int gc_frame_begin(gc_ctx *ctx, int32_t replica, int32_t fraction_bits)
{
    GC_CHECK_CTX(ctx);
    if (replica < -1 || replica >= ctx->n_rep || fraction_bits < 1 || fraction_bits > 24) {
        return fail(ctx, GC_ERR_INVALID, "bad args");
    }
    if (ctx->frame_pending) return fail(ctx, GC_ERR_STATE, "a frame is already in flight");
    const int r0 = replica < 0 ? 0 : replica, nrep = replica < 0 ? ctx->n_rep : 1;
    const size_t count = (size_t) nrep * ctx->n * 3;
    if (count == 0) return GC_OK;
    int rc = ensure_stage(ctx, 1);                               // copy stream + event
    if (rc) return rc;
    if (count > ctx->frame_cap) {
        if (ctx->d_frame) dev_free(ctx, ctx->d_frame);
        ctx->d_frame = nullptr;
        ctx->frame_cap = 0;
        rc = dev_alloc(ctx, &ctx->d_frame, count);
        if (rc) return rc;
        ctx->frame_cap = count;
    }
    if (count > ctx->h_frame_cap) {
        if (ctx->h_frame) cudaFreeHost(ctx->h_frame);
        ctx->h_frame = nullptr;
        ctx->h_frame_cap = 0;
        GC_CUDA(ctx, cudaMallocHost(&ctx->h_frame, count * sizeof(float)));
        ctx->h_frame_cap = count;
    }
    const int blocks = (int) std::min<size_t>((count / 3 + 255) / 256, 148 * 8);
    LAUNCH(ctx, k_frame, std::max(blocks, 1), 256, ctx->hp, r0, nrep, (int) fraction_bits, ctx->d_frame);
    // the copy leaves on its own stream: steps queued on the launch stream after this call
    // (they never touch d_frame) overlap the device-to-host transfer
    GC_CUDA(ctx, cudaEventRecord(ctx->copy_event, ctx->stream));
    GC_CUDA(ctx, cudaStreamWaitEvent(ctx->copy_stream, ctx->copy_event, 0));
    GC_CUDA(ctx, cudaMemcpyAsync(ctx->h_frame, ctx->d_frame, count * sizeof(float),
                                 cudaMemcpyDeviceToHost, ctx->copy_stream));
    ctx->frame_pending = count;
    return GC_OK;
}

int gc_frame_end(gc_ctx *ctx, float *xyz)
{
    GC_CHECK_CTX(ctx);
    if (!xyz) return fail(ctx, GC_ERR_INVALID, "bad args");
    if (!ctx->frame_pending) return fail(ctx, GC_ERR_STATE, "no frame in flight");
    GC_CUDA(ctx, cudaStreamSynchronize(ctx->copy_stream));
    memcpy(xyz, ctx->h_frame, ctx->frame_pending * sizeof(float));
    ctx->frame_pending = 0;
    return GC_OK;
}

int gc_debug_noise(gc_ctx *ctx, int32_t replica, int64_t step, float *out)
{
    GC_CHECK_CTX(ctx);
    if (!out || replica < 0 || replica >= ctx->n_rep) return fail(ctx, GC_ERR_INVALID, "bad args");
    int rc = finalize(ctx);
    if (rc) return rc;
    float *d = nullptr;
    rc = dev_alloc(ctx, &d, (size_t) ctx->n * 3);
    if (rc) return rc;
    LAUNCH(ctx, k_debug_noise, (ctx->n + 255) / 256, 256, ctx->hp, replica, (long long) step, d);
    GC_CUDA(ctx, cudaMemcpyAsync(out, d, (size_t) ctx->n * 3 * sizeof(float),
                                 cudaMemcpyDeviceToHost, ctx->stream));
    GC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    dev_free(ctx, d);
    return GC_OK;
}

int gc_compute_force(gc_ctx *ctx, int32_t replica, uint32_t term_mask, double *out)
{
    GC_CHECK_CTX(ctx);
    if (!out || replica < 0 || replica >= ctx->n_rep) return fail(ctx, GC_ERR_INVALID, "bad args");
    if (ctx->hp.decomp) return fail(ctx, GC_ERR_STATE, "decomposed context: forces are partial per rank");
    int rc = finalize(ctx);
    if (rc) return rc;
    plan_and_rebuild(ctx, 0.0);
    launch_rebuild(ctx);
    launch_step(ctx, STEP_FORCE, term_mask, 0);
    std::vector<float4> f((size_t) ctx->n);
    GC_CUDA(ctx, cudaMemcpyAsync(f.data(), ctx->hp.force + (size_t) replica * ctx->n,
                                 f.size() * sizeof(float4), cudaMemcpyDeviceToHost, ctx->stream));
    rc = check_device_errors(ctx);
    if (rc) return rc;
    for (int i = 0; i < ctx->n; i++) {
        out[3 * i] = f[i].x;
        out[3 * i + 1] = f[i].y;
        out[3 * i + 2] = f[i].z;
    }
    return GC_OK;
}

int gc_compute_energy(gc_ctx *ctx, int32_t replica, uint32_t term_mask, double *out)
{
    GC_CHECK_CTX(ctx);
    if (!out || replica < 0 || replica >= ctx->n_rep) return fail(ctx, GC_ERR_INVALID, "bad args");
    int rc = finalize(ctx);
    if (rc) return rc;
    plan_and_rebuild(ctx, 0.0);
    launch_rebuild(ctx);
    int nb = (ctx->n + BS - 1) / BS;
    LAUNCH(ctx, k_energy, std::max(nb, 1), BS, ctx->hp, replica, term_mask, ctx->d_partials);
    std::vector<double> part((size_t) std::max(nb, 1));
    GC_CUDA(ctx, cudaMemcpyAsync(part.data(), ctx->d_partials, part.size() * sizeof(double),
                                 cudaMemcpyDeviceToHost, ctx->stream));
    rc = check_device_errors(ctx);
    if (rc) return rc;
    double e = 0;
    for (int b = 0; b < nb; b++) e += part[b];
    *out = e;
    return GC_OK;
}

int gc_compute_energy_all(gc_ctx *ctx, uint32_t term_mask, double *out)
{
    GC_CHECK_CTX(ctx);
    if (!out) return fail(ctx, GC_ERR_INVALID, "bad args");
    int rc = finalize(ctx);
    if (rc) return rc;
    plan_and_rebuild(ctx, 0.0);
    launch_rebuild(ctx);
    const int nb = std::max((ctx->n + BS - 1) / BS, 1), R = ctx->n_rep;
    double *d_part = nullptr;
    rc = dev_alloc(ctx, &d_part, (size_t) R * nb);
    if (rc) return rc;
    for (int r = 0; r < R; r++) {
        LAUNCH(ctx, k_energy, nb, BS, ctx->hp, r, term_mask, d_part + (size_t) r * nb);
    }
    std::vector<double> part((size_t) R * nb);
    GC_CUDA(ctx, cudaMemcpyAsync(part.data(), d_part, part.size() * sizeof(double),
                                 cudaMemcpyDeviceToHost, ctx->stream));
    rc = check_device_errors(ctx);
    dev_free(ctx, d_part);
    if (rc) return rc;
    for (int r = 0; r < R; r++) {
        double e = 0;
        for (int b = 0; b < nb; b++) e += part[(size_t) r * nb + b];   // block order: deterministic
        out[r] = e;
    }
    return GC_OK;
}

int gc_neighbor_pairs(gc_ctx *ctx, int32_t replica, double radius, uint32_t *pairs,
                      size_t capacity, size_t *count)
{
    GC_CHECK_CTX(ctx);
    if (!count || replica < 0 || replica >= ctx->n_rep || radius == 0 || radius != radius) {
        return fail(ctx, GC_ERR_INVALID, "bad args");
    }
    int rc = finalize(ctx);
    if (rc) return rc;
    if (radius > 0) {
        plan_and_rebuild(ctx, radius);
        launch_rebuild(ctx);
    } else {
        // diagnostic: enumerate through the cell list as the last step left it
        // (tests the skin / rebuild invariant); |radius| must be <= rc_valid
        radius = -radius;
    }
    uint2 *d_out = nullptr;
    rc = dev_alloc(ctx, &d_out, capacity);
    if (rc) return rc;
    GC_CUDA(ctx, cudaMemsetAsync(ctx->d_counter, 0, sizeof(unsigned long long), ctx->stream));
    LAUNCH(ctx, k_pairs<PAIRS_LIST>, bead_grid(ctx, SBEADS), SBS, ctx->hp, radius, (int) RADIUS_ARG,
           replica, d_out, (unsigned long long) capacity, ctx->d_counter);
    unsigned long long found = 0;
    GC_CUDA(ctx, cudaMemcpyAsync(&found, ctx->d_counter, sizeof found, cudaMemcpyDeviceToHost,
                                 ctx->stream));
    rc = check_device_errors(ctx);
    if (rc) { dev_free(ctx, d_out); return rc; }
    *count = (size_t) found;
    size_t m = std::min<size_t>(found, capacity);
    std::vector<uint2> host(m);
    if (m && pairs) {
        GC_CUDA(ctx, cudaMemcpy(host.data(), d_out, m * sizeof(uint2), cudaMemcpyDeviceToHost));
        std::sort(host.begin(), host.end(), [](uint2 a, uint2 b) {
            return a.x != b.x ? a.x < b.x : a.y < b.y;
        });
        for (size_t k = 0; k < m; k++) {
            pairs[2 * k] = host[k].x;
            pairs[2 * k + 1] = host[k].y;
        }
    }
    dev_free(ctx, d_out);
    if (found > capacity) return fail(ctx, GC_ERR_CAPACITY, "pair buffer too small: %llu pairs", found);
    return GC_OK;
}

// ---------------------------------------------------------------------------
// contact map
// ---------------------------------------------------------------------------

int gc_set_contact_map(gc_ctx *ctx, double distance, int64_t interval, int32_t scale_by_core,
                       size_t table_capacity)
{
    GC_CHECK_CTX(ctx);
    if (!(distance >= 0) || interval < 0) return fail(ctx, GC_ERR_INVALID, "bad args");
    ContactParams &C = ctx->hp.contact;
    size_t cap = table_capacity ? table_capacity : std::max<size_t>(64 * (size_t) ctx->n, 1 << 16);
    size_t pow2 = 1;
    while (pow2 < cap) pow2 <<= 1;
    int slots = 128;
    if (const char *env = getenv("GC_CONTACT_BUCKET")) slots = std::max(1, atoi(env));
    while (slots & (slots - 1)) slots++;               // power of two
    if (!C.keys || C.capacity != pow2 || C.bucket_slots != slots) {
        if (C.keys) dev_free(ctx, C.keys);
        if (C.vals) dev_free(ctx, C.vals);
        if (C.used) dev_free(ctx, C.used);
        if (C.bucket) dev_free(ctx, C.bucket);
        C.keys = nullptr; C.vals = nullptr; C.used = nullptr; C.bucket = nullptr;
        int rc = dev_alloc(ctx, &C.keys, pow2 * ctx->n_rep, 0xff);
        if (!rc) rc = dev_alloc(ctx, &C.vals, pow2 * ctx->n_rep);
        if (!rc) rc = dev_alloc(ctx, &C.used, (size_t) ctx->n_rep);
        if (!rc) rc = dev_alloc(ctx, &C.bucket, (size_t) ctx->n_rep * ctx->n * slots, 0xff);
        if (rc) return rc;
        C.capacity = pow2;
        C.bucket_slots = slots;
    }
    C.base_distance = distance;
    C.interval = interval;
    C.scale_by_core = scale_by_core;
    C.enabled = distance > 0;
    ctx->params_dirty = true;
    // contact distance in force now (SIM/stage_interphase/simulation_driver_interphase.cpp:75)
    return modify_states(ctx, -1, [&](RepState &st, int) {
        st.contact_distance = scale_by_core ? distance * st.core_scale : distance;
        st.contact_distance_prev = st.contact_distance;
    });
}

int gc_contacts_update(gc_ctx *ctx)
{
    GC_CHECK_CTX(ctx);
    if (!ctx->hp.contact.enabled) return fail(ctx, GC_ERR_STATE, "contact map is not configured");
    int rc = finalize(ctx);
    if (rc) return rc;
    plan_and_rebuild(ctx, 0.0);
    launch_rebuild(ctx);
    LAUNCH(ctx, k_contact_cells, ctx->tile_grid_x, CW * 32, ctx->hp, (int) RADIUS_CONTACT);
    return check_device_errors(ctx);
}

int gc_contacts_clear(gc_ctx *ctx)
{
    GC_CHECK_CTX(ctx);
    ContactParams &C = ctx->hp.contact;
    if (!C.keys) return GC_OK;
    GC_CUDA(ctx, cudaMemsetAsync(C.keys, 0xff, C.capacity * ctx->n_rep * sizeof(unsigned long long),
                                 ctx->stream));
    GC_CUDA(ctx, cudaMemsetAsync(C.bucket, 0xff, (size_t) ctx->n_rep * ctx->n * C.bucket_slots *
                                                     sizeof(unsigned long long), ctx->stream));
    GC_CUDA(ctx, cudaMemsetAsync(C.vals, 0, C.capacity * ctx->n_rep * sizeof(unsigned int),
                                 ctx->stream));
    GC_CUDA(ctx, cudaMemsetAsync(C.used, 0, ctx->n_rep * sizeof(unsigned int), ctx->stream));
    GC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    // a "table full" condition ends with the clear
    return modify_states(ctx, -1, [&](RepState &st, int) {
        if (st.error == 2) st.error = 0;
    });
}

} // extern "C"

namespace {

void free_contact_result(gc_ctx *ctx)
{
    dev_free(ctx, ctx->res_keys);
    dev_free(ctx, ctx->res_vals);
    ctx->res_keys = nullptr;
    ctx->res_vals = nullptr;
    ctx->res_count = 0;
}

// Sorts (key, count) pairs by key = (i << 32 | j) -- the order contact_map::accumulate
// produces (SIM/stage_interphase/contact_map.cpp:75-82) -- and, if `reduce`, sums the
// counts of equal keys (ensemble sum).  Takes ownership of k_in / v_in; the result
// lives in ctx->res_* until the next contact query.
int sort_reduce_contacts(gc_ctx *ctx, unsigned long long *k_in, unsigned int *v_in, size_t m,
                         bool reduce)
{
    unsigned long long *k_out = nullptr, *d_runs = nullptr;
    unsigned int *v_out = nullptr;
    unsigned char *tmp = nullptr;
    auto cleanup = [&] {
        dev_free(ctx, k_in); dev_free(ctx, v_in); dev_free(ctx, k_out); dev_free(ctx, v_out);
        dev_free(ctx, d_runs); dev_free(ctx, tmp);
    };
    free_contact_result(ctx);
    if (m == 0) { cleanup(); return GC_OK; }
    if (m > (size_t) INT32_MAX) { cleanup(); return fail(ctx, GC_ERR_CAPACITY, "too many contact keys"); }
    int rc = dev_alloc(ctx, &k_out, m);
    if (!rc) rc = dev_alloc(ctx, &v_out, m);
    if (!rc) rc = dev_alloc(ctx, &d_runs, 1);
    if (rc) { cleanup(); return rc; }
    size_t tmp_bytes = 0, tmp2 = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, k_in, k_out, v_in, v_out, (int) m, 0, 64,
                                    ctx->stream);
    if (reduce) {
        cub::DeviceReduce::ReduceByKey(nullptr, tmp2, k_out, k_in, v_out, v_in, d_runs, cub::Sum(),
                                       (int) m, ctx->stream);
    }
    tmp_bytes = std::max(tmp_bytes, tmp2);
    rc = dev_alloc(ctx, &tmp, tmp_bytes);
    if (rc) { cleanup(); return rc; }
    cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, k_in, k_out, v_in, v_out, (int) m, 0, 64,
                                    ctx->stream);
    ctx->launches += 8;
    size_t n_out = m;
    if (reduce) {
        cub::DeviceReduce::ReduceByKey(tmp, tmp_bytes, k_out, k_in, v_out, v_in, d_runs, cub::Sum(),
                                       (int) m, ctx->stream);
        ctx->launches += 2;
        unsigned long long runs = 0;
        cudaError_t e = cudaMemcpyAsync(&runs, d_runs, sizeof runs, cudaMemcpyDeviceToHost, ctx->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
        if (e != cudaSuccess) {
            cleanup();
            return fail(ctx, GC_ERR_CUDA, "contact reduce: %s", cudaGetErrorString(e));
        }
        n_out = (size_t) runs;
        std::swap(k_in, k_out);
        std::swap(v_in, v_out);
    }
    ctx->res_keys = k_out;
    ctx->res_vals = v_out;
    ctx->res_count = n_out;
    k_out = nullptr;
    v_out = nullptr;
    cudaStreamSynchronize(ctx->stream);
    cleanup();
    return GC_OK;
}

// table slots of one replica (or all) -> sorted (and, for all, summed) device arrays
int reduce_contact_tables(gc_ctx *ctx, int replica)
{
    ContactParams &C = ctx->hp.contact;
    free_contact_result(ctx);
    if (!C.keys) return GC_OK;
    int rc = finalize(ctx);
    if (rc) return rc;
    std::vector<unsigned int> used((size_t) ctx->n_rep);
    GC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    GC_CUDA(ctx, cudaMemcpy(used.data(), C.used, used.size() * sizeof(unsigned int),
                            cudaMemcpyDeviceToHost));
    size_t m = 0;
    for (int r = 0; r < ctx->n_rep; r++) {
        if (replica < 0 || r == replica) m += used[r];
    }
    if (m == 0) return GC_OK;
    unsigned long long *k_in = nullptr;
    unsigned int *v_in = nullptr;
    rc = dev_alloc(ctx, &k_in, m);
    if (!rc) rc = dev_alloc(ctx, &v_in, m);
    if (rc) { dev_free(ctx, k_in); dev_free(ctx, v_in); return rc; }
    GC_CUDA(ctx, cudaMemsetAsync(ctx->d_counter, 0, sizeof(unsigned long long), ctx->stream));
    LAUNCH(ctx, k_contact_compact, 148 * 8, 256, ctx->hp, replica, k_in, v_in, ctx->d_counter);
    return sort_reduce_contacts(ctx, k_in, v_in, m, replica < 0);
}

} // namespace

extern "C" {

int gc_contacts_fetch(gc_ctx *ctx, int32_t replica, uint32_t *ijv, size_t capacity, size_t *count)
{
    GC_CHECK_CTX(ctx);
    if (!count || replica < -1 || replica >= ctx->n_rep) return fail(ctx, GC_ERR_INVALID, "bad args");
    *count = 0;
    int rc = reduce_contact_tables(ctx, replica);
    if (rc) return rc;
    const size_t n_out = ctx->res_count;
    *count = n_out;
    if (ijv && n_out) {
        size_t take = std::min(n_out, capacity);
        std::vector<unsigned long long> hk(take);
        std::vector<unsigned int> hv(take);
        GC_CUDA(ctx, cudaMemcpyAsync(hk.data(), ctx->res_keys, take * sizeof(unsigned long long),
                                     cudaMemcpyDeviceToHost, ctx->stream));
        GC_CUDA(ctx, cudaMemcpyAsync(hv.data(), ctx->res_vals, take * sizeof(unsigned int),
                                     cudaMemcpyDeviceToHost, ctx->stream));
        GC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        for (size_t k = 0; k < take; k++) {
            ijv[3 * k] = (uint32_t) (hk[k] >> 32);
            ijv[3 * k + 1] = (uint32_t) hk[k];
            ijv[3 * k + 2] = hv[k];
        }
    }
    if (ijv && n_out > capacity) return fail(ctx, GC_ERR_CAPACITY, "contact buffer too small: %zu", n_out);
    return GC_OK;
}

int gc_contacts_reduce_device(gc_ctx *ctx, int32_t replica, uint64_t **keys, uint32_t **counts,
                              size_t *n)
{
    GC_CHECK_CTX(ctx);
    if (!keys || !counts || !n || replica < -1 || replica >= ctx->n_rep) {
        return fail(ctx, GC_ERR_INVALID, "bad args");
    }
    int rc = reduce_contact_tables(ctx, replica);
    if (rc) return rc;
    *keys = (uint64_t *) ctx->res_keys;
    *counts = ctx->res_vals;
    *n = ctx->res_count;
    return GC_OK;
}

int gc_contacts_merge_device(gc_ctx *ctx, const uint64_t *keys_in, const uint32_t *counts_in,
                             size_t n_in, uint64_t **keys, uint32_t **counts, size_t *n)
{
    GC_CHECK_CTX(ctx);
    if (!keys || !counts || !n || (n_in && (!keys_in || !counts_in))) {
        return fail(ctx, GC_ERR_INVALID, "bad args");
    }
    unsigned long long *k_in = nullptr;
    unsigned int *v_in = nullptr;
    int rc = dev_alloc(ctx, &k_in, n_in);
    if (!rc) rc = dev_alloc(ctx, &v_in, n_in);
    if (rc) { dev_free(ctx, k_in); dev_free(ctx, v_in); return rc; }
    // the inputs may be the previous result buffers: copy before they are released
    GC_CUDA(ctx, cudaMemcpyAsync(k_in, keys_in, n_in * sizeof(uint64_t), cudaMemcpyDeviceToDevice,
                                 ctx->stream));
    GC_CUDA(ctx, cudaMemcpyAsync(v_in, counts_in, n_in * sizeof(uint32_t), cudaMemcpyDeviceToDevice,
                                 ctx->stream));
    rc = sort_reduce_contacts(ctx, k_in, v_in, n_in, true);
    if (rc) return rc;
    *keys = (uint64_t *) ctx->res_keys;
    *counts = ctx->res_vals;
    *n = ctx->res_count;
    return GC_OK;
}

// ---------------------------------------------------------------------------
// observables, misc
// ---------------------------------------------------------------------------

int gc_radius_of_gyration(gc_ctx *ctx, int32_t replica, int32_t n_ranges, const int32_t *ranges,
                          double *out_rg2)
{
    GC_CHECK_CTX(ctx);
    if (replica < 0 || replica >= ctx->n_rep || n_ranges < 0 || (n_ranges && (!ranges || !out_rg2))) {
        return fail(ctx, GC_ERR_INVALID, "bad args");
    }
    if (n_ranges == 0) return GC_OK;
    for (int g = 0; g < n_ranges; g++) {
        if (ranges[2 * g] < 0 || ranges[2 * g + 1] > ctx->n || ranges[2 * g] > ranges[2 * g + 1]) {
            return fail(ctx, GC_ERR_INVALID, "bad range %d", g);
        }
    }
    int rc = finalize(ctx);
    if (rc) return rc;
    int *d_ranges = nullptr;
    double *d_out = nullptr;
    rc = dev_alloc(ctx, &d_ranges, (size_t) 2 * n_ranges);
    if (!rc) rc = dev_alloc(ctx, &d_out, (size_t) n_ranges);
    if (rc) return rc;
    GC_CUDA(ctx, cudaMemcpyAsync(d_ranges, ranges, (size_t) 2 * n_ranges * sizeof(int),
                                 cudaMemcpyHostToDevice, ctx->stream));
    LAUNCH(ctx, k_rg2, n_ranges, 256, ctx->hp, replica, n_ranges, d_ranges, d_out);
    GC_CUDA(ctx, cudaMemcpyAsync(out_rg2, d_out, (size_t) n_ranges * sizeof(double),
                                 cudaMemcpyDeviceToHost, ctx->stream));
    GC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    dev_free(ctx, d_ranges);
    dev_free(ctx, d_out);
    return GC_OK;
}

} // extern "C"

// ---------------------------------------------------------------------------
// Allocation-free variant of gc_radius_of_gyration
// (no counterpart in the reference).  The scratch of k_rg2 -- the chain ranges and
// the per-chain results -- lives in a file-scope device array instead of cudaMalloc'ed
// buffers: 64 KiB held by every GPU that loads the library, until exit.  In host code the
// array's NAME is only a host-side stand-in; the device address comes from
// cudaGetSymbolAddress, once per device, with that device current.
// ---------------------------------------------------------------------------
struct RgScratch {
    int ranges[2 * GC_RG_NOALLOC_MAX_RANGES];
    double out[GC_RG_NOALLOC_MAX_RANGES];
};
static __device__ __align__(256) RgScratch g_rg_scratch;

namespace {
constexpr int RG_MAX_DEVICES = 64;
RgScratch *g_rg_scratch_ptr[RG_MAX_DEVICES] = {};

// device address of g_rg_scratch on ctx's device (which GC_CHECK_CTX has made current)
int rg_scratch_address(gc_ctx *ctx, RgScratch **out)
{
    if (ctx->device < 0 || ctx->device >= RG_MAX_DEVICES) return fail(ctx, GC_ERR_INVALID, "device index %d", ctx->device);
    if (!g_rg_scratch_ptr[ctx->device]) {
        void *p = nullptr;
        GC_CUDA(ctx, cudaGetSymbolAddress(&p, g_rg_scratch));
        g_rg_scratch_ptr[ctx->device] = static_cast<RgScratch *>(p);
    }
    *out = g_rg_scratch_ptr[ctx->device];
    return GC_OK;
}
} // namespace

extern "C" {

int gc_radius_of_gyration_noalloc(gc_ctx *ctx, int32_t replica, int32_t n_ranges, const int32_t *ranges,
                                  double *out_rg2)
{
    GC_CHECK_CTX(ctx);
    if (replica < 0 || replica >= ctx->n_rep || n_ranges < 0 || (n_ranges && (!ranges || !out_rg2))) {
        return fail(ctx, GC_ERR_INVALID, "bad args");
    }
    if (n_ranges > GC_RG_NOALLOC_MAX_RANGES) {
        return fail(ctx, GC_ERR_CAPACITY, "gc_radius_of_gyration_noalloc: %d ranges exceed the static scratch of %d "
                    "(use gc_radius_of_gyration)", n_ranges, GC_RG_NOALLOC_MAX_RANGES);
    }
    if (n_ranges == 0) return GC_OK;
    for (int g = 0; g < n_ranges; g++) {
        if (ranges[2 * g] < 0 || ranges[2 * g + 1] > ctx->n || ranges[2 * g] > ranges[2 * g + 1]) {
            return fail(ctx, GC_ERR_INVALID, "bad range %d", g);
        }
    }
    RgScratch *scratch = nullptr;
    int rc = rg_scratch_address(ctx, &scratch);
    if (rc) return rc;
    // the array keeps the last call's contents: the results this call reads back start at 0
    GC_CUDA(ctx, cudaMemsetAsync(scratch->out, 0, (size_t) n_ranges * sizeof(double), ctx->stream));
    GC_CUDA(ctx, cudaMemcpyAsync(scratch->ranges, ranges, (size_t) 2 * n_ranges * sizeof(int),
                                 cudaMemcpyHostToDevice, ctx->stream));
    k_rg2<<<n_ranges, 256, 0, ctx->stream>>>(ctx->hp, replica, n_ranges, scratch->ranges, scratch->out);
    ctx->launches++;
    GC_CUDA(ctx, cudaGetLastError());
    GC_CUDA(ctx, cudaMemcpyAsync(out_rg2, scratch->out, (size_t) n_ranges * sizeof(double),
                                 cudaMemcpyDeviceToHost, ctx->stream));
    GC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return GC_OK;
}

int gc_debug_grid(gc_ctx *ctx, int32_t replica, double out[16])
{
    GC_CHECK_CTX(ctx);
    if (!out || replica < 0 || replica >= ctx->n_rep) return fail(ctx, GC_ERR_INVALID, "bad args");
    RepState st;
    GC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    GC_CUDA(ctx, cudaMemcpy(&st, ctx->hp.st + replica, sizeof st, cudaMemcpyDeviceToHost));
    for (int k = 0; k < 16; k++) out[k] = 0;
    for (int k = 0; k < 3; k++) {
        out[k] = st.dims[k];
        out[4 + k] = st.origin[k];
    }
    out[3] = st.ncell;
    out[7] = st.cell_size;
    out[8] = st.rc_valid;
    float md2;
    memcpy(&md2, &st.maxdisp2_bits, 4);
    out[9] = std::sqrt((double) md2);
    out[10] = st.need_rebuild;
    out[11] = st.parity;
    out[12] = st.sc.dt_cur;
    return GC_OK;
}

int gc_measure_fp32_peak(int32_t device, double *tflops)
{
    if (!tflops) return GC_ERR_INVALID;
    *tflops = 0.0;
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || device < 0 || device >= count) return GC_ERR_CUDA;
    // scalar FFMA chains, or (GC_PEAK_PACKED=1) packed fma.rn.f32x2 chains
    const char *env = getenv("GC_PEAK_PACKED");
    return measure_fp32_peak(device, tflops, env && atoi(env)) == cudaSuccess ? GC_OK : GC_ERR_CUDA;
}

int gc_profile_enable(gc_ctx *ctx, int32_t enabled)
{
    GC_CHECK_CTX(ctx);
    GC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (enabled) prof_clear(ctx);
    ctx->profiling = enabled != 0;
    return GC_OK;
}

int gc_profile_read(gc_ctx *ctx, int32_t index, char *name, size_t name_capacity,
                    int64_t *launches, double *total_ms)
{
    GC_CHECK_CTX(ctx);
    if (index < 0 || index >= (int) ctx->prof.size()) return GC_ERR_INVALID;
    GC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    const auto &k = ctx->prof[index];
    double total = 0.0;
    for (size_t e = 0; e + 1 < k.ev.size(); e += 2) {
        float ms = 0.0f;
        GC_CUDA(ctx, cudaEventElapsedTime(&ms, k.ev[e], k.ev[e + 1]));
        total += ms;
    }
    if (name && name_capacity) snprintf(name, name_capacity, "%s", k.name.c_str());
    if (launches) *launches = (int64_t) (k.ev.size() / 2);
    if (total_ms) *total_ms = total;
    return GC_OK;
}

int gc_launch_count(gc_ctx *ctx, int64_t *out)
{
    if (!ctx || !out) return GC_ERR_INVALID;
    *out = ctx->launches;
    return GC_OK;
}

// ---------------------------------------------------------------------------
// spatial decomposition (gc_decomp.cuh)
// ---------------------------------------------------------------------------
namespace {

void decomp_peer_close(gc_ctx *ctx)
{
    Params &P = ctx->hp;
    for (void *m : ctx->ipc_mapped) cudaIpcCloseMemHandle(m);
    ctx->ipc_mapped.clear();
    for (int b = 0; b < 2; b++) P.dpeer_lo[b] = P.dpeer_hi[b] = nullptr;
    if (P.dpeer) ctx->params_dirty = true;
    P.dpeer = 0;
}

void decomp_release(gc_ctx *ctx)
{
    Params &P = ctx->hp;
    decomp_peer_close(ctx);
    if (P.dstamp) dev_free(ctx, P.dstamp);
    if (P.dsend_lo) dev_free(ctx, P.dsend_lo);
    if (P.dsend_hi) dev_free(ctx, P.dsend_hi);
    for (int b = 0; b < 2; b++) {
        if (P.drecv_lo[b]) dev_free(ctx, const_cast<float4 *>(P.drecv_lo[b]));
        if (P.drecv_hi[b]) dev_free(ctx, const_cast<float4 *>(P.drecv_hi[b]));
        P.drecv_lo[b] = P.drecv_hi[b] = nullptr;
    }
    if (P.dhdr) dev_free(ctx, P.dhdr);
    if (P.dhdr_all) dev_free(ctx, const_cast<double *>(P.dhdr_all));
    P.dstamp = nullptr;
    P.dsend_lo = P.dsend_hi = nullptr;
    P.dhdr = nullptr;
    P.dhdr_all = nullptr;
    P.decomp = 0;
    ctx->params_dirty = true;
}

// two send buffers (in peer mode only their counters are used) and 2 x 2 receive buffers
int decomp_alloc_halo(gc_ctx *ctx, int cap)
{
    Params &P = ctx->hp;
    if (P.dpeer) return fail(ctx, GC_ERR_STATE, "detach the peer buffers before resizing (gc_decomp_peer_detach)");
    float4 *b[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    for (int k = 0; k < 6; k++) {
        int rc = dev_alloc(ctx, &b[k], (size_t) cap + 1);
        if (rc) return rc;
    }
    // the received halos are part of the next step's work list (decomp_current_bead): carry
    // them over when the buffers are resized mid-run
    const float4 *old_recv[4] = {P.drecv_lo[0], P.drecv_lo[1], P.drecv_hi[0], P.drecv_hi[1]};
    for (int k = 0; k < 4; k++) {
        if (!old_recv[k]) continue;
        int count = 0;
        GC_CUDA(ctx, cudaMemcpyAsync(&count, old_recv[k], sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
        GC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        if (count > cap) {
            for (int j = 0; j < 6; j++) dev_free(ctx, b[j]);
            return fail(ctx, GC_ERR_CAPACITY, "halo capacity %d is below the %d beads just received", cap, count);
        }
        GC_CUDA(ctx, cudaMemcpyAsync(b[2 + k], old_recv[k], ((size_t) count + 1) * sizeof(float4),
                                     cudaMemcpyDeviceToDevice, ctx->stream));
    }
    GC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (P.dsend_lo) dev_free(ctx, P.dsend_lo);
    if (P.dsend_hi) dev_free(ctx, P.dsend_hi);
    for (int k = 0; k < 4; k++) {
        if (old_recv[k]) dev_free(ctx, const_cast<float4 *>(old_recv[k]));
    }
    P.dsend_lo = b[0];
    P.dsend_hi = b[1];
    P.drecv_lo[0] = b[2];
    P.drecv_lo[1] = b[3];
    P.drecv_hi[0] = b[4];
    P.drecv_hi[1] = b[5];
    P.dcap = cap;
    ctx->params_dirty = true;
    return GC_OK;
}

} // namespace

int gc_decomp_configure(gc_ctx *ctx, int32_t rank, int32_t world)
{
    GC_CHECK_CTX(ctx);
    if (world < 1 || rank < 0 || rank >= world) return fail(ctx, GC_ERR_INVALID, "bad rank / world");
    if (world > GC_MAX_RANKS) return fail(ctx, GC_ERR_INVALID, "at most %d ranks", GC_MAX_RANKS);
    if (ctx->n_rep != 1) return fail(ctx, GC_ERR_STATE, "decomposition needs n_replicas = 1");
    decomp_release(ctx);
    if (world == 1) return GC_OK;
    Params &P = ctx->hp;
    int *stamp = nullptr;
    double *hdr = nullptr, *hdr_all = nullptr;
    int rc = dev_alloc(ctx, &stamp, (size_t) ctx->n);
    if (!rc) rc = dev_alloc(ctx, &hdr, (size_t) DHDR);
    if (!rc) rc = dev_alloc(ctx, &hdr_all, (size_t) DHDR * world);
    if (rc) return rc;
    GC_CUDA(ctx, cudaMemsetAsync(stamp, 0, (size_t) ctx->n * sizeof(int), ctx->stream));
    GC_CUDA(ctx, cudaMemsetAsync(hdr, 0, DHDR * sizeof(double), ctx->stream));
    GC_CUDA(ctx, cudaMemsetAsync(hdr_all, 0, (size_t) DHDR * world * sizeof(double), ctx->stream));
    P.dstamp = stamp;
    P.dhdr = hdr;
    P.dhdr_all = hdr_all;
    // a rank never ships more than it owns; an equal-volume slab of a uniformly filled
    // nucleus owns about N / world beads, 2x that is a generous default
    rc = decomp_alloc_halo(ctx, std::max(1024, (int) std::min<long long>(ctx->n, 2LL * ((ctx->n + world - 1) / world))));
    if (rc) return rc;
    P.decomp = 1;
    P.drank = rank;
    P.dworld = world;
    P.dhalo = std::max(2, P.cells_per_cut);
    P.down_bound = ctx->n;
    return GC_OK;
}

int gc_decomp_set_halo(gc_ctx *ctx, int32_t layers, int32_t capacity)
{
    GC_CHECK_CTX(ctx);
    Params &P = ctx->hp;
    if (!P.decomp) return fail(ctx, GC_ERR_STATE, "gc_decomp_configure has not been called");
    if (layers < 1 || capacity < 0) return fail(ctx, GC_ERR_INVALID, "bad halo layers / capacity");
    P.dhalo = layers;
    ctx->params_dirty = true;
    if (capacity > 0 && capacity != P.dcap) return decomp_alloc_halo(ctx, capacity);
    return GC_OK;
}

int gc_decomp_set_owned_bound(gc_ctx *ctx, int32_t beads)
{
    GC_CHECK_CTX(ctx);
    if (!ctx->hp.decomp) return fail(ctx, GC_ERR_STATE, "gc_decomp_configure has not been called");
    ctx->hp.down_bound = beads > 0 ? std::min(beads, ctx->n) : ctx->n;
    ctx->params_dirty = true;
    return GC_OK;
}

int gc_decomp_buffers(gc_ctx *ctx, void **send_lo, void **send_hi, void **recv_lo, void **recv_hi,
                      void **header, void **headers_all, size_t *halo_elems)
{
    GC_CHECK_CTX(ctx);
    if (!send_lo || !send_hi || !recv_lo || !recv_hi || !header || !headers_all || !halo_elems) {
        return fail(ctx, GC_ERR_INVALID, "bad args");
    }
    const Params &P = ctx->hp;
    if (!P.decomp) return fail(ctx, GC_ERR_STATE, "gc_decomp_configure has not been called");
    *send_lo = P.dsend_lo;
    *send_hi = P.dsend_hi;
    *recv_lo = const_cast<float4 *>(P.drecv_lo[0]);
    *recv_hi = const_cast<float4 *>(P.drecv_hi[0]);
    *header = P.dhdr;
    *headers_all = const_cast<double *>(P.dhdr_all);
    *halo_elems = (size_t) P.dcap + 1;
    return GC_OK;
}

// ---- peer mode: neighbours write the halo straight into this rank's receive buffers ------
int gc_decomp_recv_buffers(gc_ctx *ctx, void **out4)
{
    GC_CHECK_CTX(ctx);
    const Params &P = ctx->hp;
    if (!out4) return fail(ctx, GC_ERR_INVALID, "bad args");
    if (!P.decomp) return fail(ctx, GC_ERR_STATE, "gc_decomp_configure has not been called");
    out4[0] = const_cast<float4 *>(P.drecv_lo[0]);
    out4[1] = const_cast<float4 *>(P.drecv_lo[1]);
    out4[2] = const_cast<float4 *>(P.drecv_hi[0]);
    out4[3] = const_cast<float4 *>(P.drecv_hi[1]);
    return GC_OK;
}

int gc_decomp_peer_export(gc_ctx *ctx, void *blob)
{
    GC_CHECK_CTX(ctx);
    static_assert(sizeof(cudaIpcMemHandle_t) == GC_DECOMP_IPC_HANDLE_BYTES, "IPC handle size");
    void *buf[4];
    int rc = gc_decomp_recv_buffers(ctx, buf);
    if (rc) return rc;
    GC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    for (int k = 0; k < 4; k++) {
        cudaIpcMemHandle_t h;
        GC_CUDA(ctx, cudaIpcGetMemHandle(&h, buf[k]));
        memcpy((char *) blob + (size_t) k * sizeof(h), &h, sizeof(h));
    }
    return GC_OK;
}

int gc_decomp_peer_attach(gc_ctx *ctx, void *lower_recv_hi0, void *lower_recv_hi1, void *upper_recv_lo0,
                          void *upper_recv_lo1)
{
    GC_CHECK_CTX(ctx);
    Params &P = ctx->hp;
    if (!P.decomp) return fail(ctx, GC_ERR_STATE, "gc_decomp_configure has not been called");
    if ((P.drank > 0 && (!lower_recv_hi0 || !lower_recv_hi1)) ||
        (P.drank + 1 < P.dworld && (!upper_recv_lo0 || !upper_recv_lo1))) {
        return fail(ctx, GC_ERR_INVALID, "a neighbour's receive buffers are missing");
    }
    P.dpeer_lo[0] = (float4 *) lower_recv_hi0;
    P.dpeer_lo[1] = (float4 *) lower_recv_hi1;
    P.dpeer_hi[0] = (float4 *) upper_recv_lo0;
    P.dpeer_hi[1] = (float4 *) upper_recv_lo1;
    P.dpeer = 1;
    ctx->params_dirty = true;
    return GC_OK;
}

int gc_decomp_peer_attach_ipc(gc_ctx *ctx, const void *lower_blob, const void *upper_blob)
{
    GC_CHECK_CTX(ctx);
    Params &P = ctx->hp;
    if (!P.decomp) return fail(ctx, GC_ERR_STATE, "gc_decomp_configure has not been called");
    decomp_peer_close(ctx);
    void *ptr[4] = {nullptr, nullptr, nullptr, nullptr};   // lower hi[0], hi[1], upper lo[0], lo[1]
    for (int k = 0; k < 4; k++) {
        const bool lower = k < 2;
        if (lower ? P.drank == 0 : P.drank + 1 >= P.dworld) continue;
        const void *blob = lower ? lower_blob : upper_blob;
        if (!blob) return fail(ctx, GC_ERR_INVALID, "a neighbour's export blob is missing");
        // blob = recv_lo[0], recv_lo[1], recv_hi[0], recv_hi[1] of that rank
        cudaIpcMemHandle_t h;
        memcpy(&h, (const char *) blob + (size_t) (lower ? 2 + k : k - 2) * sizeof(h), sizeof(h));
        GC_CUDA(ctx, cudaIpcOpenMemHandle(&ptr[k], h, cudaIpcMemLazyEnablePeerAccess));
        ctx->ipc_mapped.push_back(ptr[k]);
    }
    return gc_decomp_peer_attach(ctx, ptr[0], ptr[1], ptr[2], ptr[3]);
}

int gc_decomp_peer_detach(gc_ctx *ctx)
{
    GC_CHECK_CTX(ctx);
    GC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    decomp_peer_close(ctx);
    return GC_OK;
}

int gc_decomp_step_compute(gc_ctx *ctx)
{
    GC_CHECK_CTX(ctx);
    const Params &P = ctx->hp;
    if (!P.decomp) return fail(ctx, GC_ERR_STATE, "gc_decomp_configure has not been called");
    if (P.skin > 0.0f || P.spacestep != 0.0f || P.contact.enabled || P.noise) {
        return fail(ctx, GC_ERR_STATE, "decomposed runs: fixed timestep, no skin, no contact map");
    }
    if (ctx->params_dirty || ctx->topo_dirty) return fail(ctx, GC_ERR_STATE, "call gc_run(steps = 0) first");
    ctx->decomp_all_current = false;
    launch_rebuild(ctx);
    launch_pair(ctx, 0u, 1);
    LAUNCH(ctx, k_finish<STEP_FUSED>, dim3((P.down_bound + BS - 1) / BS, 1), BS, ctx->hp, 0u, 1);
    LAUNCH(ctx, k_decomp_header, 1, 32, ctx->hp);
    if (P.dpeer) {
        // the halo leaves now, straight into the neighbours' buffers; the header all-gather
        // the caller runs next is the step's only collective
        LAUNCH(ctx, k_decomp_pack, std::min((P.down_bound + 255) / 256, 148 * 8), 256, ctx->hp);
        LAUNCH(ctx, k_decomp_publish, 1, 32, ctx->hp);
    }
    return GC_OK;
}

int gc_decomp_step_pack(gc_ctx *ctx)
{
    GC_CHECK_CTX(ctx);
    const Params &P = ctx->hp;
    if (!P.decomp) return fail(ctx, GC_ERR_STATE, "gc_decomp_configure has not been called");
    LAUNCH(ctx, k_finalize_step, 1, 32, ctx->hp, 0);
    if (!P.dpeer) LAUNCH(ctx, k_decomp_pack, std::min((P.down_bound + 255) / 256, 148 * 8), 256, ctx->hp);
    ctx->run_step++;
    ctx->parity ^= 1;
    return GC_OK;
}

int gc_decomp_step_commit(gc_ctx *ctx)
{
    GC_CHECK_CTX(ctx);
    if (!ctx->hp.decomp) return fail(ctx, GC_ERR_STATE, "gc_decomp_configure has not been called");
    // nothing to launch: the next step's k_count takes the received beads in on its way
    return GC_OK;
}

int gc_decomp_host_advance(gc_ctx *ctx, int64_t steps)
{
    GC_CHECK_CTX(ctx);
    if (!ctx->hp.decomp) return fail(ctx, GC_ERR_STATE, "gc_decomp_configure has not been called");
    ctx->run_step += steps;
    if (steps & 1) ctx->parity ^= 1;
    return GC_OK;
}

int gc_decomp_owned(gc_ctx *ctx, uint8_t *mask)
{
    GC_CHECK_CTX(ctx);
    if (!mask) return fail(ctx, GC_ERR_INVALID, "bad args");
    if (!ctx->hp.decomp) return fail(ctx, GC_ERR_STATE, "gc_decomp_configure has not been called");
    unsigned char *d = nullptr;
    int rc = dev_alloc(ctx, &d, (size_t) ctx->n);
    if (rc) return rc;
    // halos received since the last step are taken in first
    k_decomp_unpack<<<dim3(std::max(1, std::min((ctx->hp.dcap + 255) / 256, 148 * 4)), 2), 256, 0, ctx->stream>>>(ctx->hp);
    k_decomp_owned<<<(ctx->n + 255) / 256, 256, 0, ctx->stream>>>(ctx->hp, d);
    cudaError_t e = cudaMemcpyAsync(mask, d, (size_t) ctx->n, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    dev_free(ctx, d);
    if (e != cudaSuccess) return fail(ctx, GC_ERR_CUDA, cudaGetErrorString(e));
    return check_device_errors(ctx);
}

// ---------------------------------------------------------------------------
// inter-stage transitions between two contexts of one device (gc_transition.cuh)
// ---------------------------------------------------------------------------
} // extern "C"

namespace {

// common checks + ordering: everything queued on src is done before dst's stream reads it
int transition_begin(gc_ctx *src, int src_replica, gc_ctx *dst, int dst_replica, int n_chains,
                     const float4 **sp, float4 **dp)
{
    if (!src || !dst) return GC_ERR_INVALID;
    if (src->device != dst->device) return fail(dst, GC_ERR_INVALID, "contexts live on different devices");
    if (src_replica < 0 || src_replica >= src->n_rep || dst_replica < 0 || dst_replica >= dst->n_rep ||
        n_chains < 0) {
        return fail(dst, GC_ERR_INVALID, "bad args");
    }
    GC_CUDA(dst, cudaSetDevice(dst->device));
    GC_CUDA(dst, cudaStreamSynchronize(src->stream));
    *sp = src->hp.pos[src->parity] + (size_t) src_replica * src->n;
    *dp = dst->hp.pos[dst->parity] + (size_t) dst_replica * dst->n;
    return GC_OK;
}

int check_ranges(gc_ctx *ctx, const int32_t *ranges, int n_chains, int n_beads, const char *what)
{
    for (int c = 0; c < n_chains; c++) {
        if (ranges[2 * c] < 0 || ranges[2 * c + 1] < ranges[2 * c] || ranges[2 * c + 1] > n_beads) {
            return fail(ctx, GC_ERR_INVALID, "%s range %d = [%d, %d) outside [0, %d)", what, c,
                        ranges[2 * c], ranges[2 * c + 1], n_beads);
        }
    }
    return GC_OK;
}

int upload_ints(gc_ctx *ctx, const int32_t *host, size_t count, int **dev)
{
    *dev = nullptr;
    int rc = dev_alloc(ctx, dev, std::max<size_t>(count, 1));
    if (rc) return rc;
    if (count) {
        GC_CUDA(ctx, cudaMemcpyAsync(*dev, host, count * sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
    }
    return GC_OK;
}

} // namespace

extern "C" {

int gc_transition_interphase(gc_ctx *src, int32_t src_replica, gc_ctx *dst, int32_t dst_replica,
                             int32_t n_chains, const int32_t *src_ranges, const int32_t *dst_ranges,
                             int32_t n_copies, const int32_t *copy_pairs, int32_t fraction_bits)
{
    const float4 *sp;
    float4 *dp;
    int rc = transition_begin(src, src_replica, dst, dst_replica, n_chains, &sp, &dp);
    if (rc) return rc;
    if (!src_ranges || !dst_ranges || n_copies < 0 || (n_copies && !copy_pairs)) {
        return fail(dst, GC_ERR_INVALID, "bad args");
    }
    if ((rc = check_ranges(dst, src_ranges, n_chains, src->n, "source")) != 0) return rc;
    if ((rc = check_ranges(dst, dst_ranges, n_chains, dst->n, "destination")) != 0) return rc;
    for (int k = 0; k < 2 * n_copies; k++) {
        if (copy_pairs[k] < 0 || copy_pairs[k] >= dst->n) return fail(dst, GC_ERR_INVALID, "copy pair out of range");
    }
    int max_knots = 1;
    for (int c = 0; c < n_chains; c++) max_knots = std::max(max_knots, src_ranges[2 * c + 1] - src_ranges[2 * c]);
    const size_t smem = (size_t) 15 * max_knots * sizeof(double);
    if (smem > 200u * 1024u) return fail(dst, GC_ERR_CAPACITY, "source chain of %d beads: spline does not fit shared memory", max_knots);
    if (n_chains == 0) return GC_OK;
    GC_CUDA(dst, cudaFuncSetAttribute(k_transition_spline, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
    int *d_src = nullptr, *d_dst = nullptr, *d_cp = nullptr;
    if ((rc = upload_ints(dst, src_ranges, 2 * (size_t) n_chains, &d_src)) != 0) return rc;
    if ((rc = upload_ints(dst, dst_ranges, 2 * (size_t) n_chains, &d_dst)) != 0) return rc;
    if ((rc = upload_ints(dst, copy_pairs, 2 * (size_t) n_copies, &d_cp)) != 0) return rc;
    k_transition_spline<<<n_chains, 256, smem, dst->stream>>>(sp, dp, d_src, d_dst, max_knots, (int) fraction_bits);
    dst->launches++;
    if (n_copies) {
        k_transition_copy<<<(n_copies + 127) / 128, 128, 0, dst->stream>>>(dp, d_cp, n_copies);
        dst->launches++;
    }
    GC_CUDA(dst, cudaStreamSynchronize(dst->stream));
    GC_CUDA(dst, cudaGetLastError());
    dev_free(dst, d_src);
    dev_free(dst, d_dst);
    dev_free(dst, d_cp);
    return GC_OK;
}

int gc_transition_prometaphase(gc_ctx *src, int32_t src_replica, gc_ctx *dst, int32_t dst_replica,
                               int32_t n_chains, const int32_t *src_ranges, const int32_t *target_ranges,
                               const int32_t *sister_ranges, int32_t coarse_graining,
                               const double displacement[3], int32_t fraction_bits)
{
    const float4 *sp;
    float4 *dp;
    int rc = transition_begin(src, src_replica, dst, dst_replica, n_chains, &sp, &dp);
    if (rc) return rc;
    if (!src_ranges || !target_ranges || coarse_graining < 1 || (sister_ranges && !displacement)) {
        return fail(dst, GC_ERR_INVALID, "bad args");
    }
    if ((rc = check_ranges(dst, src_ranges, n_chains, src->n, "source")) != 0) return rc;
    if ((rc = check_ranges(dst, target_ranges, n_chains, dst->n, "target")) != 0) return rc;
    if (sister_ranges && (rc = check_ranges(dst, sister_ranges, n_chains, dst->n, "sister")) != 0) return rc;
    int max_len = 0;
    for (int c = 0; c < n_chains; c++) {
        const int len = target_ranges[2 * c + 1] - target_ranges[2 * c];
        if (sister_ranges && sister_ranges[2 * c + 1] - sister_ranges[2 * c] < len) {
            return fail(dst, GC_ERR_INVALID, "sister chain %d shorter than its target", c);
        }
        max_len = std::max(max_len, len);
    }
    if (n_chains == 0 || max_len == 0) return GC_OK;
    int *d_src = nullptr, *d_t = nullptr, *d_s = nullptr;
    if ((rc = upload_ints(dst, src_ranges, 2 * (size_t) n_chains, &d_src)) != 0) return rc;
    if ((rc = upload_ints(dst, target_ranges, 2 * (size_t) n_chains, &d_t)) != 0) return rc;
    if (sister_ranges && (rc = upload_ints(dst, sister_ranges, 2 * (size_t) n_chains, &d_s)) != 0) return rc;
    const double dx = displacement ? displacement[0] : 0.0, dy = displacement ? displacement[1] : 0.0;
    const double dz = displacement ? displacement[2] : 0.0;
    k_transition_coarse<<<dim3((max_len + 127) / 128, n_chains), 128, 0, dst->stream>>>(
        sp, src->n, dp, n_chains, d_src, d_t, sister_ranges ? d_s : nullptr, (int) coarse_graining, dx, dy, dz,
        (int) fraction_bits);
    dst->launches++;
    GC_CUDA(dst, cudaStreamSynchronize(dst->stream));
    GC_CUDA(dst, cudaGetLastError());
    dev_free(dst, d_src);
    dev_free(dst, d_t);
    if (d_s) dev_free(dst, d_s);
    return GC_OK;
}

int gc_transition_cycle(gc_ctx *src, int32_t src_replica, gc_ctx *dst, int32_t dst_replica,
                        int32_t n_chains, const int32_t *src_ranges, const int32_t *dst_ranges,
                        const double displacement[3], int32_t fraction_bits)
{
    const float4 *sp;
    float4 *dp;
    int rc = transition_begin(src, src_replica, dst, dst_replica, n_chains, &sp, &dp);
    if (rc) return rc;
    if (!src_ranges || !dst_ranges || !displacement) return fail(dst, GC_ERR_INVALID, "bad args");
    if ((rc = check_ranges(dst, src_ranges, n_chains, src->n, "source")) != 0) return rc;
    if ((rc = check_ranges(dst, dst_ranges, n_chains, dst->n, "destination")) != 0) return rc;
    int max_len = 0;
    for (int c = 0; c < n_chains; c++) max_len = std::max(max_len, dst_ranges[2 * c + 1] - dst_ranges[2 * c]);
    if (n_chains == 0 || max_len == 0) return GC_OK;
    int *d_src = nullptr, *d_dst = nullptr;
    if ((rc = upload_ints(dst, src_ranges, 2 * (size_t) n_chains, &d_src)) != 0) return rc;
    if ((rc = upload_ints(dst, dst_ranges, 2 * (size_t) n_chains, &d_dst)) != 0) return rc;
    k_transition_shift<<<dim3((max_len + 127) / 128, n_chains), 128, 0, dst->stream>>>(
        sp, dp, n_chains, d_src, d_dst, displacement[0], displacement[1], displacement[2], (int) fraction_bits);
    dst->launches++;
    GC_CUDA(dst, cudaStreamSynchronize(dst->stream));
    GC_CUDA(dst, cudaGetLastError());
    dev_free(dst, d_src);
    dev_free(dst, d_dst);
    return GC_OK;
}

int gc_synchronize(gc_ctx *ctx)
{
    GC_CHECK_CTX(ctx);
    return check_device_errors(ctx);
}

int gc_stream(gc_ctx *ctx, void **out)
{
    if (!ctx || !out) return GC_ERR_INVALID;
    *out = (void *) ctx->stream;
    return GC_OK;
}

} // extern "C"

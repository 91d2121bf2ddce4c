// gc_oracle_port.cpp -- "port" oracle: a plain FP64 C++ restatement of the
// reference's hot-path algorithm with NO dependency on the reference tree, so
// that it can be built and run on the GPU box.  TEST INFRASTRUCTURE ONLY (see
// gc_oracle.h).  Pinned against (a) micromd's own known-answer tests
// (tests/test_oracle.py, constants from MDT/potential/*.cc and
// MDT/forcefield/test_ellipsoid_surface_forcefield.cc) and (b) the compiled
// reference oracle oracle/_ref/libgc_ref.so on seeded inputs, wherever
// /root/reference is present.
//
// MD/ = submodules/github.com/snsinfu/micromd/include/md/, SIM/ = src/simulation/.
// Deliberate differences from the reference (none changes a result beyond FP64
// rounding): pair forces are summed over the exact pair set at dcut instead of
// a Verlet list at 1.5 dcut (all potentials vanish beyond dcut); the neighbor
// search uses a sorted dense grid instead of the linear-hash buckets; without
// injected noise the normals come from mt19937_64 + Box-Muller, not
// sfc64 + ziggurat (statistical equivalence only).

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <map>
#include <random>
#include <unordered_map>
#include <utility>
#include <vector>

#include "gc_oracle.h"

namespace {

struct vec3 {
    double x = 0, y = 0, z = 0;
};
inline vec3 operator+(vec3 a, vec3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline vec3 operator-(vec3 a, vec3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline vec3 operator*(double s, vec3 a) { return {s * a.x, s * a.y, s * a.z}; }
inline double dot(vec3 a, vec3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline double norm2(vec3 a) { return dot(a, a); }

// ---- potentials ------------------------------------------------------------
// MD/misc/math.hpp:18-44
inline double ipow(double x, int n)
{
    double r = 1;
    for (int k = 0; k < n; k++) r *= x;
    return r;
}

// power_sqrt<N>(x) = x^(N/2) (MD/misc/math.hpp:37-44)
inline double ipow_half(double x, int n)
{
    return n % 2 == 0 ? ipow(x, n / 2) : ipow(std::sqrt(x), n);
}

// MD/potential/softcore_potential.hpp:22-62
struct softcore {
    int P, Q;
    double energy, diameter;
    double e(vec3 r) const
    {
        double k2 = 1 / (diameter * diameter);
        double u2 = k2 * norm2(r);
        double g = 1 - ipow_half(u2, P);
        return g < 0 ? 0 : energy * ipow(g, Q);
    }
    vec3 f(vec3 r) const
    {
        double k2 = 1 / (diameter * diameter);
        double u2 = k2 * norm2(r);
        double v = ipow_half(u2, P - 2);
        double g = 1 - v * u2;
        if (g < 0) return {};
        return (P * Q * energy * k2 * ipow(g, Q - 1) * v) * r;
    }
};

// a * softcore<2,3> + b * softcore<8,3>
// (SIM/stage_interphase/simulation_driver_forcefield.cpp:30-52,
//  MD/potential/sum_potential.hpp:19-52, scaled_potential.hpp:19-57)
struct ab_softcore {
    double a, b;
    softcore pa, pb;
    double e(vec3 r) const { return a * pa.e(r) + b * pb.e(r); }
    vec3 f(vec3 r) const { return a * pa.f(r) + b * pb.f(r); }
};

// MD/potential/softwell_potential.hpp:25-51 (P = 6) + cutoff_potential.hpp:18-53
struct softwell6 {
    double energy, decay, cutoff;   // cutoff <= 0: none
    bool cut(vec3 r) const { return cutoff > 0 && norm2(r) >= cutoff * cutoff; }
    double e(vec3 r) const
    {
        if (cut(r)) return 0;
        double u2 = norm2(r) / (decay * decay);
        return -energy / (1 + u2 * u2 * u2);
    }
    vec3 f(vec3 r) const
    {
        if (cut(r)) return {};
        double k2 = 1 / (decay * decay);
        double u2 = k2 * norm2(r);
        double up2 = u2 * u2;
        double v1 = 1 + up2 * u2;
        return ((-6 * energy * k2) * up2 / (v1 * v1)) * r;
    }
};

// MD/potential/{semispring,harmonic,spring}_potential.hpp
struct bond_pot {
    int kind;
    double K, l;
    double e(vec3 r) const
    {
        if (kind == GC_BOND_HARMONIC) return 0.5 * K * norm2(r);
        double u = std::sqrt(norm2(r)) - l;
        if (kind == GC_BOND_SEMISPRING && u <= 0) return 0;
        return 0.5 * K * u * u;
    }
    vec3 f(vec3 r) const
    {
        if (kind == GC_BOND_HARMONIC) return (-K) * r;
        double r1 = std::sqrt(norm2(r));
        if (kind == GC_BOND_SEMISPRING ? r1 <= l : r1 == 0) return {};
        return (K * l / r1 - K) * r;
    }
};

// SIM/common/potentials/force_flux_potential.hpp:19-38
struct flux_pot {
    double f0, b;
    double e(vec3 r) const { return f0 * b * std::atan2(b, std::sqrt(norm2(r))); }
    vec3 f(vec3 r) const
    {
        double r2 = norm2(r), r1 = std::sqrt(r2);
        double b2 = b * b;
        return (f0 * b2 / (b2 * r1 + r1 * r2)) * r;
    }
};

// MD/potential/cosine_bending_potential.hpp:23-62
struct bending {
    double e0;
    double e(vec3 rij, vec3 rjk) const
    {
        double a = norm2(rij), b = norm2(rjk);
        if (a * b == 0) return 0;
        return e0 - e0 * dot(rij, rjk) / std::sqrt(a * b);
    }
    void f(vec3 rij, vec3 rjk, vec3 out[3]) const
    {
        double a = norm2(rij), b = norm2(rjk);
        double dd = std::sqrt(a * b);
        out[0] = out[1] = out[2] = vec3{};
        if (dd == 0) return;
        double d = dot(rij, rjk);
        double ed = e0 / dd;
        vec3 fij = ed * (rjk - (d / a) * rij);
        vec3 fjk = ed * (rij - (d / b) * rjk);
        out[0] = fij;
        out[1] = fjk - fij;
        out[2] = (-1.0) * fjk;
    }
};

// ---- neighbor search ---------------------------------------------------------
// Contract of md::neighbor_searcher<open_box>::search
// (MD/misc/neighbor_searcher.hpp:75-100): every unordered pair with
// |xi - xj|^2 <= dcut^2 (inclusive) exactly once, as (min, max).
void search_pairs(const std::vector<vec3> &x, const std::vector<uint32_t> &subset, double dcut,
                  std::vector<std::pair<uint32_t, uint32_t>> &out)
{
    out.clear();
    size_t n = subset.size();
    if (n < 2 || !(dcut > 0)) return;
    double dcut2 = dcut * dcut;
    // cell coordinates as in search_grid<open_box>::locate_bucket
    // (MD/misc/nsearch_detail/search_grid.hpp:231-244): trunc(2^20 + x / dcut)
    struct item { uint64_t key; uint32_t idx; };
    auto coord = [&](double v) { return (uint64_t) (uint32_t) (1048576.0 + v / dcut); };
    std::vector<item> items(n);
    for (size_t k = 0; k < n; k++) {
        vec3 p = x[subset[k]];
        items[k] = {(coord(p.z) << 42) | (coord(p.y) << 21) | coord(p.x), subset[k]};
    }
    std::sort(items.begin(), items.end(), [](const item &a, const item &b) {
        return a.key != b.key ? a.key < b.key : a.idx < b.idx;
    });
    std::unordered_map<uint64_t, std::pair<uint32_t, uint32_t>> cells;
    cells.reserve(n);
    for (size_t k = 0; k < n;) {
        size_t e = k;
        while (e < n && items[e].key == items[k].key) e++;
        cells[items[k].key] = {(uint32_t) k, (uint32_t) e};
        k = e;
    }
    for (auto &kv : cells) {
        uint64_t key = kv.first;
        int64_t cx = key & 0x1fffff, cy = (key >> 21) & 0x1fffff, cz = key >> 42;
        for (int dz = -1; dz <= 1; dz++)
            for (int dy = -1; dy <= 1; dy++)
                for (int dx = -1; dx <= 1; dx++) {
                    uint64_t nk = ((uint64_t) (cz + dz) << 42) | ((uint64_t) (cy + dy) << 21) |
                                  (uint64_t) (cx + dx);
                    if (nk < key) continue;   // each cell pair once
                    auto it = cells.find(nk);
                    if (it == cells.end()) continue;
                    for (uint32_t a = kv.second.first; a < kv.second.second; a++) {
                        uint32_t b0 = nk == key ? a + 1 : it->second.first;
                        for (uint32_t b = b0; b < it->second.second; b++) {
                            uint32_t i = items[a].idx, j = items[b].idx;
                            if (norm2(x[i] - x[j]) > dcut2) continue;
                            out.emplace_back(std::min(i, j), std::max(i, j));
                        }
                    }
                }
    }
    std::sort(out.begin(), out.end());
}

enum term_kind { T_PAIR_AB, T_PAIR_PLAIN, T_WELL, T_BONDS, T_TRIPLES, T_WALL, T_SOURCE, T_KFIBER };

struct term {
    term_kind kind;
    bool enabled = true;
    // pair
    double eps_a = 0, sigma_a = 0, eps_b = 0, sigma_b = 0;
    bool scale_by_core = false;
    // well
    double decay = 0, cutoff = 0;
    std::vector<uint32_t> targets;
    // bonds / triples / kfiber
    gc_bond_params bp{};
    std::vector<int32_t> idx;       // flattened pairs / triples / beads
    std::vector<double> K, L;
    double bending_energy = 0;
    // source / pole
    int source_kind = 0;
    vec3 point;
    double p0 = 0, p1 = 0;
};

} // namespace

struct gco_sys {
    size_t n = 0;
    std::vector<vec3> x;
    std::vector<double> mob, a, b;
    std::vector<term> terms;
    gc_context ctx{};
    gc_schedule sched{};
    gc_wall_params wall{};
    bool has_wall = false;
    vec3 axial;            // wall statistics of the last force evaluation
    double reaction = 0;
    std::map<uint64_t, uint32_t> contacts;
    double contact_base = 0, contact_distance = 0;
    int64_t contact_interval = 0;
    bool contact_scaled = false;
    std::vector<uint32_t> all;   // 0..n-1
};

namespace {

double core_of(const gco_sys *s, bool scaled) { return scaled ? s->ctx.core_scale : 1.0; }

ab_softcore pair_pot(const gco_sys *s, const term &t, size_t i, size_t j)
{
    double sc = core_of(s, t.scale_by_core);
    if (t.kind == T_PAIR_PLAIN) {
        return {1.0, 0.0, {2, 3, t.eps_a, t.sigma_a}, {8, 3, 0.0, t.sigma_a}};
    }
    return {0.5 * (s->a[i] + s->a[j]), 0.5 * (s->b[i] + s->b[j]),
            {2, 3, t.eps_a, t.sigma_a * sc}, {8, 3, t.eps_b, t.sigma_b * sc}};
}

double pair_dcut(const gco_sys *s, const term &t)
{
    if (t.kind == T_WELL) return t.cutoff;
    if (t.kind == T_PAIR_PLAIN) return t.sigma_a;
    return std::max(t.sigma_a, t.sigma_b) * core_of(s, t.scale_by_core);
}

// interphase bond lambdas: SIM/stage_interphase/simulation_driver_forcefield.cpp:56-160
bond_pot bond_of(const gco_sys *s, const term &t, size_t i, size_t j)
{
    double K = t.bp.K_a, l = t.bp.l_a;
    if (t.bp.mix_ab) {
        double a = 0.5 * (s->a[i] + s->a[j]);
        double b = 0.5 * (s->b[i] + s->b[j]);
        K = a * t.bp.K_a + b * t.bp.K_b;
        l = a * t.bp.l_a + b * t.bp.l_b;
    }
    if (t.bp.scale) {
        double sc = s->ctx.bond_scale;
        K = K * (1 / (sc * sc));
        l = l * sc;
    }
    return {t.bp.kind, K, l};
}

// MD/forcefield/ellipsoid_surface_forcefield.hpp:55-90
struct ell_eval {
    bool undefined;
    vec3 delta, strain;
    double implicit;
};

ell_eval evaluate_point(const double semi[3], vec3 pt)
{
    vec3 q = {1 / (semi[0] * semi[0]), 1 / (semi[1] * semi[1]), 1 / (semi[2] * semi[2])};
    vec3 dual = {q.x * pt.x, q.y * pt.y, q.z * pt.z};
    ell_eval ev{};
    if (norm2(dual) == 0) {
        ev.undefined = true;
        return ev;
    }
    double implicit = dot(dual, pt) - 1;
    double scale = implicit / (2 * norm2(dual));
    ev.undefined = false;
    ev.delta = scale * dual;
    ev.strain = scale * q;
    ev.implicit = implicit;
    return ev;
}

// Evaluates one term.  F may be null (energy only).  Returns the energy.
double eval_term(gco_sys *s, term &t, std::vector<vec3> *F)
{
    const auto &x = s->x;
    double E = 0;
    switch (t.kind) {
    case T_PAIR_AB:
    case T_PAIR_PLAIN: {
        // MD/forcefield/neighbor_pairwise_forcefield.hpp:48-84
        std::vector<std::pair<uint32_t, uint32_t>> pairs;
        search_pairs(x, s->all, pair_dcut(s, t), pairs);
        for (auto &pr : pairs) {
            auto pot = pair_pot(s, t, pr.first, pr.second);
            vec3 r = x[pr.first] - x[pr.second];
            E += pot.e(r);
            if (F) {
                vec3 f = pot.f(r);
                (*F)[pr.first] = (*F)[pr.first] + f;
                (*F)[pr.second] = (*F)[pr.second] - f;
            }
        }
        break;
    }
    case T_WELL: {
        std::vector<std::pair<uint32_t, uint32_t>> pairs;
        search_pairs(x, t.targets, t.cutoff, pairs);
        softwell6 pot{t.eps_a, t.decay, t.cutoff};
        for (auto &pr : pairs) {
            vec3 r = x[pr.first] - x[pr.second];
            E += pot.e(r);
            if (F) {
                vec3 f = pot.f(r);
                (*F)[pr.first] = (*F)[pr.first] + f;
                (*F)[pr.second] = (*F)[pr.second] - f;
            }
        }
        break;
    }
    case T_BONDS:
        // MD/forcefield/bonded_pairwise_forcefield.hpp:56-93
        for (size_t k = 0; k + 1 < t.idx.size(); k += 2) {
            size_t i = t.idx[k], j = t.idx[k + 1];
            auto pot = bond_of(s, t, i, j);
            vec3 r = x[i] - x[j];
            E += pot.e(r);
            if (F) {
                vec3 f = pot.f(r);
                (*F)[i] = (*F)[i] + f;
                (*F)[j] = (*F)[j] - f;
            }
        }
        break;
    case T_TRIPLES: {
        // MD/forcefield/bonded_triplewise_forcefield.hpp:54-100
        bending pot{t.bending_energy};
        for (size_t k = 0; k + 2 < t.idx.size(); k += 3) {
            size_t i = t.idx[k], j = t.idx[k + 1], m = t.idx[k + 2];
            vec3 rij = x[i] - x[j], rjk = x[j] - x[m];
            E += pot.e(rij, rjk);
            if (F) {
                vec3 f[3];
                pot.f(rij, rjk, f);
                (*F)[i] = (*F)[i] + f[0];
                (*F)[j] = (*F)[j] + f[1];
                (*F)[m] = (*F)[m] + f[2];
            }
        }
        break;
    }
    case T_WALL: {
        // MD/forcefield/ellipsoid_surface_forcefield.hpp:138-211; inward then
        // outward force fields of the driver fused (they act on disjoint beads),
        // statistics summed as _compute_packing_reaction does
        // (SIM/stage_interphase/simulation_driver_forcefield.cpp:239-243)
        const gc_wall_params &w = s->wall;
        double sc = core_of(s, w.scale_by_core);
        if (F) {
            s->axial = vec3{};
            s->reaction = 0;
        }
        for (size_t i = 0; i < s->n; i++) {
            ell_eval ev = evaluate_point(s->ctx.semiaxes, x[i]);
            if (ev.undefined) continue;
            vec3 basic;
            if (ev.implicit < 0) {
                ab_softcore pot{0.5 * (s->a[i] + w.wall_a), 0.5 * (s->b[i] + w.wall_b),
                                {2, 3, w.eps_a, w.sigma_a / 2 * sc},
                                {8, 3, w.eps_b, w.sigma_b / 2 * sc}};
                E += pot.e(ev.delta);
                if (F) basic = pot.f(ev.delta);
            } else {
                bond_pot pot{GC_BOND_HARMONIC, w.outward_spring, 0};
                E += pot.e(ev.delta);
                if (F) basic = pot.f(ev.delta);
            }
            if (!F) continue;
            double d2 = norm2(ev.delta);
            vec3 iso = (dot(basic, ev.delta) / d2) * ev.delta;   // project()
            vec3 aniso = basic - iso;
            vec3 df = aniso - iso;
            vec3 strain_force = {df.x * ev.strain.x, df.y * ev.strain.y, df.z * ev.strain.z};
            vec3 force = iso + strain_force;
            (*F)[i] = (*F)[i] + force;
            vec3 normal = ((ev.implicit < 0 ? -1.0 : 1.0) / std::sqrt(d2)) * ev.delta;
            s->reaction -= dot(force, normal);
            s->axial.x -= normal.x > 0 ? force.x : -force.x;
            s->axial.y -= normal.y > 0 ? force.y : -force.y;
            s->axial.z -= normal.z > 0 ? force.z : -force.z;
        }
        break;
    }
    case T_SOURCE:
        // MD/forcefield/point_source_forcefield.hpp:52-110
        for (size_t i = 0; i < s->n; i++) {
            vec3 r = x[i] - t.point;
            if (t.source_kind == GC_SOURCE_SEMISPRING) {
                bond_pot pot{GC_BOND_SEMISPRING, t.p0, t.p1};
                E += pot.e(r);
                if (F) (*F)[i] = (*F)[i] + pot.f(r);
            } else {
                flux_pot pot{t.p0, t.p1};
                E += pot.e(r);
                if (F) (*F)[i] = (*F)[i] + pot.f(r);
            }
        }
        break;
    case T_KFIBER:
        // SIM/common/forcefield/kinetochore_fiber_forcefield.cpp:56-84
        for (size_t k = 0; k < t.idx.size(); k++) {
            size_t i = t.idx[k];
            bond_pot pot{GC_BOND_SPRING, t.K[k], t.L[k]};
            vec3 r = x[i] - t.point;
            E += pot.e(r);
            if (F) (*F)[i] = (*F)[i] + pot.f(r);
        }
        break;
    }
    return E;
}

void compute_all(gco_sys *s, uint32_t mask, std::vector<vec3> &F)
{
    F.assign(s->n, vec3{});
    for (size_t t = 0; t < s->terms.size(); t++) {
        if ((mask >> t & 1u) && s->terms[t].enabled) eval_term(s, s->terms[t], &F);
    }
}

int push(gco_sys *s, term t)
{
    s->terms.push_back(std::move(t));
    return int(s->terms.size()) - 1;
}

} // namespace

extern "C" {

const char *gco_kind(void) { return "port"; }

gco_sys *gco_create(int32_t n_beads)
{
    auto *s = new gco_sys;
    s->n = size_t(n_beads);
    s->x.resize(s->n);
    s->mob.assign(s->n, 1.0);
    s->a.assign(s->n, 0.0);
    s->b.assign(s->n, 0.0);
    s->all.resize(s->n);
    for (size_t i = 0; i < s->n; i++) s->all[i] = uint32_t(i);
    s->ctx.core_scale = 1;
    s->ctx.bond_scale = 1;
    return s;
}

void gco_destroy(gco_sys *s) { delete s; }

void gco_set_bead_attrs(gco_sys *s, const double *mobility, const double *a, const double *b)
{
    for (size_t i = 0; i < s->n; i++) {
        s->mob[i] = mobility[i];
        s->a[i] = a ? a[i] : 0;
        s->b[i] = b ? b[i] : 0;
    }
}

void gco_set_positions(gco_sys *s, const double *xyz)
{
    for (size_t i = 0; i < s->n; i++) s->x[i] = {xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]};
}

void gco_get_positions(gco_sys *s, double *xyz)
{
    for (size_t i = 0; i < s->n; i++) {
        xyz[3 * i] = s->x[i].x;
        xyz[3 * i + 1] = s->x[i].y;
        xyz[3 * i + 2] = s->x[i].z;
    }
}

int gco_add_pair_softcore_ab(gco_sys *s, double eps_a, double sigma_a, double eps_b,
                             double sigma_b, int32_t scale_by_core)
{
    term t;
    t.kind = T_PAIR_AB;
    t.eps_a = eps_a; t.sigma_a = sigma_a; t.eps_b = eps_b; t.sigma_b = sigma_b;
    t.scale_by_core = scale_by_core != 0;
    return push(s, t);
}

int gco_add_pair_softcore(gco_sys *s, double eps, double sigma)
{
    term t;
    t.kind = T_PAIR_PLAIN;
    t.eps_a = eps; t.sigma_a = sigma;
    return push(s, t);
}

int gco_add_pair_softwell(gco_sys *s, double eps, double decay, double cutoff,
                          int32_t n_targets, const int32_t *targets)
{
    term t;
    t.kind = T_WELL;
    t.eps_a = eps; t.decay = decay; t.cutoff = cutoff;
    for (int k = 0; k < n_targets; k++) t.targets.push_back(uint32_t(targets[k]));
    return push(s, t);
}

int gco_add_bonds(gco_sys *s, const gc_bond_params *params, int32_t n, const int32_t *pairs)
{
    if (params->kind < 0 || params->kind > 2) return -1;
    term t;
    t.kind = T_BONDS;
    t.bp = *params;
    t.idx.assign(pairs, pairs + 2 * size_t(n));
    return push(s, t);
}

int gco_add_triples(gco_sys *s, double bending_energy, int32_t n, const int32_t *triples)
{
    term t;
    t.kind = T_TRIPLES;
    t.bending_energy = bending_energy;
    t.idx.assign(triples, triples + 3 * size_t(n));
    return push(s, t);
}

int gco_set_wall(gco_sys *s, const gc_wall_params *params)
{
    s->wall = *params;
    s->has_wall = true;
    for (int k = 0; k < 3; k++) s->ctx.semiaxes[k] = params->semiaxes[k];
    term t;
    t.kind = T_WALL;
    return push(s, t);
}

int gco_add_point_source(gco_sys *s, int32_t kind, const double point[3], double p0, double p1)
{
    if (kind < 0 || kind > 1) return -1;
    term t;
    t.kind = T_SOURCE;
    t.source_kind = kind;
    t.point = {point[0], point[1], point[2]};
    t.p0 = p0; t.p1 = p1;
    return push(s, t);
}

int gco_add_kfiber(gco_sys *s, const double pole[3], int32_t n, const int32_t *beads,
                   const double *spring, const double *length)
{
    term t;
    t.kind = T_KFIBER;
    t.point = {pole[0], pole[1], pole[2]};
    t.idx.assign(beads, beads + n);
    t.K.assign(spring, spring + n);
    t.L.assign(length, length + n);
    return push(s, t);
}

void gco_set_term_enabled(gco_sys *s, int32_t term, int32_t enabled)
{
    s->terms.at(size_t(term)).enabled = enabled != 0;
}

void gco_set_schedule(gco_sys *s, const gc_schedule *schedule) { s->sched = *schedule; }

void gco_set_context(gco_sys *s, const gc_context *in)
{
    s->ctx.time = in->time;
    s->ctx.core_scale = in->core_scale;
    s->ctx.bond_scale = in->bond_scale;
    for (int k = 0; k < 3; k++) s->ctx.semiaxes[k] = in->semiaxes[k];
    s->contact_distance = in->contact_distance;
}

void gco_get_context(gco_sys *s, gc_context *out)
{
    s->ctx.axial_reaction[0] = s->axial.x;
    s->ctx.axial_reaction[1] = s->axial.y;
    s->ctx.axial_reaction[2] = s->axial.z;
    s->ctx.reaction_force = s->reaction;
    s->ctx.contact_distance = s->contact_distance;
    *out = s->ctx;
}

void gco_compute_force(gco_sys *s, uint32_t term_mask, double *out)
{
    std::vector<vec3> F;
    compute_all(s, term_mask, F);
    for (size_t i = 0; i < s->n; i++) {
        out[3 * i] = F[i].x;
        out[3 * i + 1] = F[i].y;
        out[3 * i + 2] = F[i].z;
    }
}

double gco_compute_energy(gco_sys *s, uint32_t term_mask)
{
    double E = 0;
    for (size_t t = 0; t < s->terms.size(); t++) {
        if ((term_mask >> t & 1u) && s->terms[t].enabled) E += eval_term(s, s->terms[t], nullptr);
    }
    return E;
}

size_t gco_neighbor_pairs(gco_sys *s, double radius, uint32_t *pairs, size_t capacity)
{
    std::vector<std::pair<uint32_t, uint32_t>> found;
    search_pairs(s->x, s->all, radius, found);
    for (size_t k = 0; k < found.size() && k < capacity; k++) {
        pairs[2 * k] = found[k].first;
        pairs[2 * k + 1] = found[k].second;
    }
    return found.size();
}

// Verlet radius = 1.5 dcut (MD/forcefield/detail/neighbor_list.hpp:178-183)
size_t gco_verlet_pairs(gco_sys *s, int32_t term, uint32_t *pairs, size_t capacity)
{
    auto &t = s->terms.at(size_t(term));
    std::vector<std::pair<uint32_t, uint32_t>> found;
    search_pairs(s->x, t.kind == T_WELL ? t.targets : s->all, 1.5 * pair_dcut(s, t), found);
    for (size_t k = 0; k < found.size() && k < capacity; k++) {
        pairs[2 * k] = found[k].first;
        pairs[2 * k + 1] = found[k].second;
    }
    return found.size();
}

void gco_set_contact_map(gco_sys *s, double distance, int64_t interval, int32_t scale_by_core)
{
    s->contact_base = distance;
    s->contact_interval = interval;
    s->contact_scaled = scale_by_core != 0;
    s->contact_distance = scale_by_core ? distance * s->ctx.core_scale : distance;
}

// SIM/stage_interphase/contact_map.cpp:33-63
void gco_contacts_update(gco_sys *s)
{
    std::vector<std::pair<uint32_t, uint32_t>> found;
    search_pairs(s->x, s->all, s->contact_distance, found);
    for (auto &pr : found) s->contacts[(uint64_t(pr.first) << 32) | pr.second] += 1;
}

// SIM/stage_interphase/contact_map.cpp:66-85 (sorted by (i << 32 | j))
size_t gco_contacts_fetch(gco_sys *s, uint32_t *ijv, size_t capacity)
{
    size_t k = 0;
    for (auto &kv : s->contacts) {
        if (k < capacity) {
            ijv[3 * k] = uint32_t(kv.first >> 32);
            ijv[3 * k + 1] = uint32_t(kv.first);
            ijv[3 * k + 2] = kv.second;
        }
        k++;
    }
    return s->contacts.size();
}

void gco_contacts_clear(gco_sys *s) { s->contacts.clear(); }

size_t gco_reference_normals(uint64_t, size_t, double *) { return 0; }

} // extern "C"

namespace {

// MD/simulation/detail/brownian_timestepper.hpp:72-91
double solve_brownian_timestep(double distance, double mobility, double force2, double temperature)
{
    const double epsilon = 1e-6, random_walk_factor = 2.55;
    double a = mobility * mobility * force2;
    double b = random_walk_factor * mobility * temperature;
    double c = distance * distance;
    if (a * c < epsilon * b * b) return c / b;
    return (-b + std::sqrt(b * b + 4 * a * c)) / (2 * a);
}

} // namespace

extern "C" {

// md::simulate_brownian_dynamics (MD/simulation/brownian_dynamics.hpp:49-72) +
// brownian_simulator::simulate_step (detail/brownian_simulator.hpp:44-76) +
// the interphase callback (SIM/stage_interphase/simulation_driver_interphase.cpp:16-90)
void gco_run(gco_sys *s, gco_run_config *c)
{
    size_t n = s->n;
    std::vector<vec3> F(n), wprev(n);
    std::mt19937_64 rng(c->seed);
    std::normal_distribution<double> gauss;
    s->ctx.last_timestep = c->timestep;

    for (int64_t step = 1; step <= c->steps; step++) {
        compute_all(s, GC_MASK_ALL, F);
        if (c->force_out) {
            double *o = c->force_out + size_t(step - 1) * n * 3;
            for (size_t i = 0; i < n; i++) {
                o[3 * i] = F[i].x; o[3 * i + 1] = F[i].y; o[3 * i + 2] = F[i].z;
            }
        }
        // timestep: fixed, or min over beads (brownian_timestepper.hpp:109-136)
        double dt = c->timestep;
        if (c->spacestep != 0) {
            for (size_t i = 0; i < n; i++) {
                double f2 = norm2(F[i]);
                double d = c->timestep;
                if (s->mob[i] != 0 || f2 != 0) {
                    d = solve_brownian_timestep(c->spacestep, s->mob[i], f2, c->temperature);
                }
                dt = std::min(dt, d);
            }
        }
        const double *z = c->noise ? c->noise + size_t(step - 1) * n * 3 : nullptr;
        for (size_t i = 0; i < n; i++) {
            double mu_dt = dt * s->mob[i];
            double sigma = std::sqrt(2 * c->temperature * mu_dt);
            vec3 w;
            if (z) {
                w = {sigma * z[3 * i], sigma * z[3 * i + 1], sigma * z[3 * i + 2]};
            } else {
                w.x = sigma * gauss(rng);
                w.y = sigma * gauss(rng);
                w.z = sigma * gauss(rng);
            }
            s->x[i] = s->x[i] + mu_dt * F[i];
            s->x[i] = s->x[i] + 0.5 * (w + wprev[i]);
            wprev[i] = w;
        }

        // ---- callback(step) ----
        int64_t abs_step = c->step_offset + step;
        s->ctx.time = double(abs_step) * c->timestep;
        s->ctx.step = abs_step;
        if (c->energy_interval && abs_step % c->energy_interval == 0) {
            c->last_energy = gco_compute_energy(s, GC_MASK_ALL) / double(n);
        }
        if (c->traj_out) gco_get_positions(s, c->traj_out + size_t(step - 1) * n * 3);
        if (s->contact_interval && abs_step % s->contact_interval == 0) gco_contacts_update(s);
        if (c->interphase_callback) {
            if (s->sched.enabled) {
                s->ctx.core_scale = 1 - (1 - s->sched.core_init) * std::exp(-s->ctx.time / s->sched.core_tau);
                s->ctx.bond_scale = 1 - (1 - s->sched.bond_init) * std::exp(-s->ctx.time / s->sched.bond_tau);
            }
            if (s->contact_scaled) s->contact_distance = s->contact_base * s->ctx.core_scale;
            if (s->has_wall && s->wall.ode_enabled) {
                double ax[3] = {s->axial.x, s->axial.y, s->axial.z};
                for (int k = 0; k < 3; k++) {
                    double net = ax[k] - s->wall.ode_spring[k] * s->ctx.semiaxes[k];
                    s->ctx.semiaxes[k] += c->timestep * s->wall.ode_mobility * net;
                }
            }
        }
        if (c->ctx_out) {
            double *o = c->ctx_out + size_t(step - 1) * 8;
            o[0] = s->ctx.time;
            o[1] = s->ctx.semiaxes[0]; o[2] = s->ctx.semiaxes[1]; o[3] = s->ctx.semiaxes[2];
            o[4] = s->ctx.core_scale; o[5] = s->ctx.bond_scale;
            o[6] = dt; o[7] = 0;
        }
    }
}

double gco_potential_probe(int32_t kind, double p0, double p1, double p2, const double r[3],
                           double f[3])
{
    vec3 rv = {r[0], r[1], r[2]}, fv;
    double e = 0;
    switch (kind) {
    case 0: { softcore p{2, 3, p0, p1}; e = p.e(rv); fv = p.f(rv); break; }
    case 1: { softcore p{8, 3, p0, p1}; e = p.e(rv); fv = p.f(rv); break; }
    case 2: { softwell6 p{p0, p1, 0}; e = p.e(rv); fv = p.f(rv); break; }
    case 3: { bond_pot p{GC_BOND_SEMISPRING, p0, p1}; e = p.e(rv); fv = p.f(rv); break; }
    case 4: { bond_pot p{GC_BOND_HARMONIC, p0, 0}; e = p.e(rv); fv = p.f(rv); break; }
    case 5: { bond_pot p{GC_BOND_SPRING, p0, p1}; e = p.e(rv); fv = p.f(rv); break; }
    case 6: { flux_pot p{p0, p1}; e = p.e(rv); fv = p.f(rv); break; }
    case 7: { softwell6 p{p0, p1, p2}; e = p.e(rv); fv = p.f(rv); break; }
    case 8: { softcore p{int(p2) / 10, int(p2) % 10, p0, p1}; e = p.e(rv); fv = p.f(rv); break; }
    default: break;
    }
    f[0] = fv.x; f[1] = fv.y; f[2] = fv.z;
    return e;
}

double gco_bending_probe(double bending_energy, const double rij[3], const double rjk[3],
                         double f[9])
{
    bending p{bending_energy};
    vec3 a = {rij[0], rij[1], rij[2]}, b = {rjk[0], rjk[1], rjk[2]}, out[3];
    p.f(a, b, out);
    for (int k = 0; k < 3; k++) {
        f[3 * k] = out[k].x; f[3 * k + 1] = out[k].y; f[3 * k + 2] = out[k].z;
    }
    return p.e(a, b);
}

void gco_ellipsoid_probe(const double semiaxes[3], const double pt[3], double out[8])
{
    ell_eval ev = evaluate_point(semiaxes, {pt[0], pt[1], pt[2]});
    for (int k = 0; k < 8; k++) out[k] = 0;
    out[0] = ev.undefined ? 1 : 0;
    if (ev.undefined) return;
    out[1] = ev.delta.x; out[2] = ev.delta.y; out[3] = ev.delta.z;
    out[4] = ev.strain.x; out[5] = ev.strain.y; out[6] = ev.strain.z;
    out[7] = ev.implicit;
}


// ---- inter-stage transitions, restated (no dependency on the reference tree) --------------
// Not-a-knot cubic spline as submodules/github.com/snsinfu/cxx-spline/include/spline.hpp
// builds it: second derivatives M from a tridiagonal system (rows 0 and n closed by the
// not-a-knot conditions, eliminated with row pivoting), piecewise cubics anchored at the left
// knot, the boundary pieces extended outside the knot range.
namespace {

struct PortSpline {
    std::vector<double> knot, c0, c1, c2, c3;

    PortSpline(const std::vector<double> &t, const std::vector<double> &x)
    {
        const size_t ns = t.size() - 1;
        std::vector<double> h(ns), sl(ns);
        for (size_t i = 0; i < ns; i++) { h[i] = t[i + 1] - t[i]; sl[i] = (x[i + 1] - x[i]) / h[i]; }
        std::vector<double> L(ns + 1, 0.0), D(ns + 1, 0.0), U(ns + 1, 0.0), Y(ns + 1, 0.0);
        if (ns == 1) {
            D[0] = 1; D[ns] = 1;                       // a straight line
        } else {
            D[0] = h[0] - h[1]; U[0] = 2 * h[0] + h[1];
            Y[0] = -6 * h[0] / (h[0] + h[1]) * (sl[0] - sl[1]);
            D[ns] = h[ns - 1] - h[ns - 2]; L[ns] = 2 * h[ns - 1] + h[ns - 2];
            Y[ns] = 6 * h[ns - 1] / (h[ns - 1] + h[ns - 2]) * (sl[ns - 1] - sl[ns - 2]);
        }
        for (size_t i = 1; i < ns; i++) {
            L[i] = h[i - 1]; D[i] = 2 * (h[i - 1] + h[i]); U[i] = h[i]; Y[i] = 6 * (sl[i] - sl[i - 1]);
        }
        const size_t dim = ns + 1;
        for (size_t i = 0; i + 1 < dim; i++) {
            if (std::fabs(D[i]) >= std::fabs(L[i + 1])) {
                const double w = L[i + 1] / D[i];
                D[i + 1] -= w * U[i]; Y[i + 1] -= w * Y[i]; L[i + 1] = 0;
            } else {                                     // row interchange
                const double w = D[i] / L[i + 1], u = D[i + 1];
                D[i] = L[i + 1]; D[i + 1] = U[i] - w * u; L[i + 1] = U[i + 1]; U[i + 1] *= -w; U[i] = u;
                const double r = Y[i];
                Y[i] = Y[i + 1]; Y[i + 1] = r - w * Y[i + 1];
            }
        }
        Y[dim - 1] /= D[dim - 1];
        for (size_t rev = 2; rev <= dim; rev++) {
            const size_t i = dim - rev;
            if (i == dim - 2) Y[i] -= U[i] * Y[i + 1];
            else Y[i] -= U[i] * Y[i + 1] + L[i + 1] * Y[i + 2];
            Y[i] /= D[i];
        }
        for (size_t i = 0; i < ns; i++) {
            knot.push_back(t[i]);
            c0.push_back(x[i]);
            c1.push_back(sl[i] - (Y[i + 1] + 2 * Y[i]) * h[i] / 6);
            c2.push_back(Y[i] / 2);
            c3.push_back((Y[i + 1] - Y[i]) / (6 * h[i]));
        }
        knot.push_back(t[ns]);
    }

    double operator()(double t) const
    {
        const size_t ns = c0.size();
        size_t i = 0;
        if (t >= knot[ns]) i = ns - 1;
        else if (t > knot[0]) { while (i + 1 < ns && t >= knot[i + 1]) i++; }
        const double d = t - knot[i];
        return ((c3[i] * d + c2[i]) * d + c1[i]) * d + c0[i];
    }
};

} // namespace

void gco_resample_chain(int32_t n, const double *chain, int32_t new_n, double *out)
{
    std::vector<double> ts(n), v(n);
    for (int32_t k = 0; k < n; k++) ts[k] = (0.5 + double(k)) / double(n);
    for (int c = 0; c < 3; c++) {
        for (int32_t k = 0; k < n; k++) v[k] = chain[3 * k + c];
        PortSpline sp(ts, v);
        for (int32_t i = 0; i < new_n; i++) out[3 * i + c] = sp((0.5 + double(i)) / double(new_n));
    }
}

void gco_coarse_grain(const double *positions, int64_t source_start, int64_t source_len, int64_t cg,
                      int64_t coarse_len, const double disp[3], double *target, double *sister)
{
    for (int64_t offset = 0; offset < coarse_len; offset++) {
        const int64_t start = source_start + cg * offset;
        const int64_t end = std::min(start + cg, start + source_len);
        double sum[3] = {0, 0, 0}, count = 0;
        for (int64_t k = start; k < end; k++) {
            for (int c = 0; c < 3; c++) sum[c] += positions[3 * k + c];
            count++;
        }
        for (int c = 0; c < 3; c++) {
            target[3 * offset + c] = sum[c] / count;
            sister[3 * offset + c] = sum[c] / count + disp[c];
        }
    }
}

void gco_quantize_f32(size_t n, const double *in, int32_t bits, float *out)
{
    for (size_t k = 0; k < n; k++) {
        int e;
        const double m = std::frexp(in[k], &e);
        out[k] = (float) std::ldexp(std::nearbyint(std::ldexp(m, bits)), e - bits);
    }
}

} // extern "C"

// hb_host.cc — host side of ThermalModel::new: what the north star keeps on the host.
//
// For every surface: decide massive vs no-mass layers, element counts and tstep_subdivision
// (reference: src/discretization.rs:410-544), build node masses and link conductances (:163-298),
// chunk the nodes (:144-160), distribute solar absorptances onto nodes (src/surface.rs:466-537,
// src/glazing.rs:67-286) and evaluate the wind-speed modifier (src/surface.rs:135-166).  Then group
// models into tiles (one CUDA block each) and lay the SoA device buffers out.
//
// The integer decisions are parity-critical (0 mismatches allowed), so the floating-point expressions
// that feed them keep the reference's operation order; this file is compiled with -ffp-contract=off.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <numeric>
#include <stdexcept>
#include <tuple>

#include "hb_dev.hpp"

namespace hb {
namespace {

inline double opt(double v, double dflt) { return std::isnan(v) ? dflt : v; }

[[noreturn]] void fail(const std::string& msg) { throw std::runtime_error(msg); }

// Smallest stable element size for a solid layer at sub-step dt (discretization.rs:453-465):
// positive root of dx^2 - dt/(rho cp Rs) dx - 2 dt k/(rho cp) = 0.
inline double min_stable_dx(const hb_layer& L, double dt) {
    const double b = -dt / (L.density * L.specific_heat * kMaxRs);
    const double c = -2. * dt * L.conductivity / (L.density * L.specific_heat);
    const double disc = b * b - 4. * 1. * c;
    return (-b + std::sqrt(disc)) / (2. * 1.);
}

// discretization.rs:410-544.  Search the smallest subdivision count `n` for which every solid layer can be
// cut into equal elements with min_dx <= dx <= max_dx; a layer that cannot is retried at n+1 while
// main_dt/(n+1) > min_dt and otherwise becomes no-mass (0 elements).  Gas layers always get 0.
void choose_elements(const hb_layer* layers, int n_layers, double main_dt, int32_t& tsub,
                     std::vector<int32_t>& elements) {
    for (int32_t n = 1;; ++n) {
        const double dt = main_dt / (double)n;
        const bool can_refine = main_dt / (double)(n + 1) > kMinDt;
        elements.clear();
        bool refine = false;
        for (int l = 0; l < n_layers && !refine; ++l) {
            const hb_layer& L = layers[l];
            if (L.kind == HB_LAYER_GAS) {
                elements.push_back(0);
                continue;
            }
            const double min_dx = min_stable_dx(L, dt);
            int32_t count = 0;
            bool fits = false;
            if (!(min_dx > L.thickness)) {
                const double m = std::floor(L.thickness / min_dx);
                const double dx = L.thickness / m;
                if (!(dx > kMaxDx)) {
                    fits = true;
                    count = (int32_t)m;
                }
            }
            if (fits) {
                elements.push_back(count);
            } else if (can_refine) {
                refine = true;
            } else {
                elements.push_back(0);
            }
        }
        if (!refine) {
            tsub = n;
            return;
        }
    }
}

struct Pane {  // glazing.rs:31-65
    double tau, rf, rb, af, ab;
    static Pane of(double tau, double rf, double rb) {
        if (!(tau >= 0. && tau <= 1. && rf >= 0. && rf <= 1. && rb >= 0. && rb <= 1.))
            fail("Glazing: transmittance/reflectance out of [0,1]");
        return Pane{tau, rf, rb, 1. - tau - rf, 1. - tau - rb};
    }
    // ISO 9050 combination of two panes (glazing.rs:180-218)
    Pane then(const Pane& o) const {
        const double rb2 = o.rb + (o.tau * o.tau) * rb / (1. - o.rf * rb);
        const double rf2 = rf + (tau * tau) * o.rf / (1. - rb * o.rf);
        const double t2 = tau * o.tau / (1. - rb * o.rf);
        return of(t2, rf2, rb2);
    }
};

Pane fold(const Pane* p, size_t n) {  // glazing.rs:221-230 (right fold)
    return n == 1 ? p[0] : p[0].then(fold(p + 1, n - 1));
}

// glazing.rs:67-112 + 259-286: absorbed fraction per pane, walking the construction from one side.
std::vector<double> pane_absorptions(const hb_layer* layers, int n_layers, bool from_back) {
    std::vector<Pane> panes;
    int i = 0;
    auto at = [&](int j) -> const hb_layer& { return layers[from_back ? n_layers - 1 - j : j]; };
    while (true) {
        if (i >= n_layers) fail("construction ended while collecting glazing layers");
        const hb_layer& L = at(i++);
        if (L.kind == HB_LAYER_GAS) fail("NOT expecting a gas");
        const double tau = opt(L.solar_transmittance, 0.0);
        // The reference reads front_solar_absorbtance as the leading-side absorptance in BOTH directions.
        const double a_lead = opt(L.front_solar_abs, 0.84);
        const double a_trail = opt(L.back_solar_abs, 0.84);
        panes.push_back(Pane::of(tau, 1. - tau - a_lead, 1. - tau - a_trail));
        if (tau < 1e-9) break;
        if (i >= n_layers) break;
        if (at(i++).kind != HB_LAYER_GAS) fail("Expecting a Gas");
    }
    std::vector<double> out;
    if (panes.size() == 1) {
        out.push_back(panes[0].af);
        return out;
    }
    double seen = 0.0;
    for (size_t k = 1; k < panes.size(); ++k) {
        const Pane g0 = fold(panes.data(), k);
        const Pane g1 = fold(panes.data() + k, panes.size() - k);
        const double denom = 1. - g0.rb * g1.rf;
        const double a0 = g0.af + g0.ab * g0.tau * g1.rf / denom;
        out.push_back(a0 - seen);
        seen = a0;
    }
    const Pane g0 = fold(panes.data(), panes.size() - 1);
    const Pane& last = panes.back();
    out.push_back(last.af * g0.tau / (1. - g0.rb * last.rf));
    return out;
}

double wind_modifier(double height, int terrain) {  // surface.rs:135-166
    if (height < 1e-5) return 0.0;
    double alpha = 0.22, delta = 370.;  // Urban; also the default without SiteDetails
    switch (terrain) {
        case HB_TERRAIN_COUNTRY: alpha = 0.14; delta = 270.; break;
        case HB_TERRAIN_CITY: alpha = 0.33; delta = 460.; break;
        case HB_TERRAIN_OCEAN: alpha = 0.10; delta = 210.; break;
        case HB_TERRAIN_UNSET: alpha = 0.0; delta = 0.0; break;
        default: break;  // SUBURBS, URBAN, NONE
    }
    return std::pow(270. / 10., 0.14) * std::pow(height / delta, alpha);
}

void runs(const std::vector<double>& mass, bool massive, std::vector<int32_t>& out) {  // discretization.rs:117-160
    const int n = (int)mass.size();
    int i = 0;
    while (i < n) {
        const bool is_m = mass[i] >= kMassiveThreshold;
        if (is_m != massive) {
            ++i;
            continue;
        }
        int j = i;
        while (j < n && (mass[j] >= kMassiveThreshold) == massive) ++j;
        out.push_back(i);
        out.push_back(j);
        i = j;
    }
}

}  // namespace

void plan_surface(const hb_layer* layers, int n_layers, const hb_surface& s, int terrain, double main_dt,
                  SurfacePlan& P) {
    if (n_layers <= 0) fail("Construction has no layers");
    for (int l = 0; l < n_layers; ++l) {
        const hb_layer& L = layers[l];
        if (L.kind == HB_LAYER_GAS) {
            if (l == 0) fail("Construction has a Gas as its first layer");
            if (l == n_layers - 1) fail("Construction has a Gas as its last layer");
            if (layers[l - 1].kind == HB_LAYER_GAS || layers[l + 1].kind == HB_LAYER_GAS)
                fail("Construction has two gases without a solid layer between them");
            if (L.gas < HB_GAS_AIR || L.gas > HB_GAS_XENON) fail("Substance does not have a standard gas.");
        } else if (L.kind != HB_LAYER_SOLID) {
            fail("unknown layer kind");
        }
    }
    P.cos_tilt = s.normal[0] * 0. + s.normal[1] * 0. + s.normal[2] * 1.;  // model.rs:249
    const double cavity_height = 1.;                                         // model.rs:252
    const double cavity_angle = std::acos(P.cos_tilt);                       // model.rs:253

    choose_elements(layers, n_layers, main_dt, P.tstep_subdivision, P.n_elements);

    // discretization.rs:174-291 — nodes, half-masses onto both ends of each element, link conductances.
    int n_nodes = 1;
    for (int32_t e : P.n_elements) n_nodes += e == 0 ? 1 : e;
    P.n_nodes = n_nodes;
    P.mass.assign(n_nodes, 0.0);
    P.u.assign(n_nodes, 0.0);
    P.cavities.clear();
    int node = 0;
    std::vector<int> layer_first_node(n_layers, 0);
    for (int l = 0; l < n_layers; ++l) {
        const hb_layer& L = layers[l];
        const int32_t e = P.n_elements[l];
        const int count = e == 0 ? 1 : e;
        layer_first_node[l] = node;
        if (L.kind == HB_LAYER_SOLID) {
            const double elem_mass = e == 0 ? 0.0 : L.density * L.specific_heat * (L.thickness / (double)e);
            const double link = L.conductivity / (L.thickness / (double)count);
            for (int k = 0; k < count; ++k, ++node) {
                P.mass[node] += elem_mass / 2.;
                P.mass[node + 1] += elem_mass / 2.;
                P.u[node] = link;
            }
        } else {
            CavityParams c;
            c.node = node;
            c.gas = L.gas;
            c.thickness = L.thickness;
            c.height = cavity_height;
            c.angle = cavity_angle;
            c.eout = opt(layers[l - 1].back_emissivity, kDefaultEm);
            c.ein = opt(layers[l + 1].front_emissivity, kDefaultEm);
            P.cavities.push_back(c);
            P.u[node] = -1.0;
            ++node;
        }
    }
    P.massive_chunks.clear();
    P.nomass_chunks.clear();
    runs(P.mass, true, P.massive_chunks);
    runs(P.mass, false, P.nomass_chunks);

    // surface.rs:449-461
    P.front_emissivity = opt(s.front_emissivity_override, opt(layers[0].front_emissivity, kDefaultEm));
    P.back_emissivity = opt(s.back_emissivity_override, opt(layers[n_layers - 1].back_emissivity, kDefaultEm));

    // surface.rs:466-537 — solar absorptance per node
    P.front_alpha.assign(n_nodes, 0.0);
    P.back_alpha.assign(n_nodes, 0.0);
    for (int side = 0; side < 2; ++side) {
        const bool from_back = side == 1;
        const std::vector<double> ap = pane_absorptions(layers, n_layers, from_back);
        const size_t len = ap.size();
        // The reference accepts len ∈ {1, n_layers}; we additionally accept the pane/gas/pane pattern its
        // `2*alpha_i` layer indexing is written for (see DESIGN.md "deviations").
        if (!(len == 1 || (int)len == n_layers || (int)len == (n_layers + 1) / 2))
            fail("Construction seems to have a mixture of transparent and opaque layers. This is not currently "
                 "supported.");
        std::vector<double>& dst = from_back ? P.back_alpha : P.front_alpha;
        for (size_t k = 0; k < len; ++k) {
            const int l = from_back ? n_layers - 2 * (int)k - 1 : 2 * (int)k;
            if (l < 0 || l >= n_layers || layers[l].kind != HB_LAYER_SOLID) fail("solar absorptance on a gas layer");
            const int count = P.n_elements[l] == 0 ? 1 : P.n_elements[l];
            const int first = layer_first_node[l], last = first + count;
            if (opt(layers[l].solar_transmittance, 0.0) > 0.0) {
                // accumulate in the reference's order: front side walks nodes forward, back side backward
                for (int j = 0; j <= count; ++j) dst[from_back ? last - j : first + j] += ap[k] / (double)(count + 1);
            } else {
                dst[from_back ? last : first] += ap[k];
            }
        }
    }
    P.distributed_solar = false;
    for (int i = 0; i < n_nodes; ++i) {
        if (i != 0 && P.front_alpha[i] != 0.0) P.distributed_solar = true;
        if (i != n_nodes - 1 && P.back_alpha[i] != 0.0) P.distributed_solar = true;
    }
    P.wind_speed_modifier = wind_modifier(s.centroid_z, terrain);
}

// Single pane: two nodes, both without mass, one chunk (0, 2), no cavity.
inline bool is_pane(const SurfacePlan& P) {
    return P.n_nodes == 2 && P.cavities.empty() && P.nomass_chunks.size() == 2 && P.nomass_chunks[0] == 0 &&
           P.nomass_chunks[1] == 2;
}
// Double-glazed window: pane | gas | pane, four nodes without mass, one chunk (0, 4), the cavity between nodes 1 and 2.
inline bool is_window4(const SurfacePlan& P) {
    return P.n_nodes == 4 && P.cavities.size() == 1 && P.cavities[0].node == 1 && P.nomass_chunks.size() == 2 &&
           P.nomass_chunks[0] == 0 && P.nomass_chunks[1] == 4;
}
constexpr int kBulkSubsteps = 56;   // below: tiles are scheduled for throughput, not for latency (hb_capi.cu S bins)
constexpr int kTileKey = TILE_EXT | TILE_DIST | TILE_END | TILE_GEN | TILE_PANE | TILE_SIMPLE | TILE_WARP;   // tiles are uniform in these

const std::vector<int>& wt_nodes_per_lane() {
    static const std::vector<int> nls = [] {
    std::vector<int> v;
    if (const char* e = std::getenv("HB_WT_NLS"))
        for (const char* p = e; *p;) {
            v.push_back(std::atoi(p));
            while (*p && *p != ',') ++p;
            if (*p == ',') ++p;
        }
    if (v.empty()) v = {8, 12, 16};
    std::sort(v.begin(), v.end());
    return v;
}();
    return nls;
}

// … and for all-massive walls (HB_WT_NLS_LEAN): with 24 / 32 the kernels keep the coefficients in shared memory
const std::vector<int>& wt_nodes_per_lane_lean() {
    static const std::vector<int> nls = [] {
        std::vector<int> v;
        if (const char* e = std::getenv("HB_WT_NLS_LEAN"))
            for (const char* p = e; *p;) {
                v.push_back(std::atoi(p));
                while (*p && *p != ',') ++p;
                if (*p == ',') ++p;
            }
        if (v.empty()) {
            v = wt_nodes_per_lane();
            if (!std::getenv("HB_WT_NLS")) v = {8, 12, 16, 24, 32};
        }
        std::sort(v.begin(), v.end());
        return v;
    }();
    return nls;
}

// Warp tiles cost S sub-steps of a whole warp however many of their wall slots are filled.  Within a bucket of equal
// kernel variant and sub-step count, the walls left over after the last FULL tile of a geometry move up to the next
// larger geometry (a wall fits any geometry of at least its size) when they would fill at most 3/4 of a tile there:
// 8.5 % fewer FP64 warp instructions on the 10^5-wall ensemble of BASELINE.json configs[2].
void wt_merge_leftovers_on(BatchPlan& B, const std::vector<WtGeom>& geoms, bool lean);
void wt_merge_leftovers(BatchPlan& B) {
    auto geoms_of = [](const std::vector<int>& nls) {
        std::vector<WtGeom> geoms;
        for (int G = 1; G <= 32; G *= 2)
            for (int NL : nls)
                if (G * NL <= kWtMaxNodes && wt_geom_exists(G, NL)) geoms.push_back(WtGeom{G * NL, G, NL});
        std::sort(geoms.begin(), geoms.end(), [](const WtGeom& a, const WtGeom& b) { return a.cap != b.cap ? a.cap < b.cap : a.NL > b.NL; });
        geoms.erase(std::unique(geoms.begin(), geoms.end(), [](const WtGeom& a, const WtGeom& b) { return a.cap == b.cap; }), geoms.end());
        return geoms;
    };
    // two ladders: the all-massive buckets merge on theirs, those with no-mass ends on the common one
    wt_merge_leftovers_on(B, geoms_of(wt_nodes_per_lane_lean()), true);
    wt_merge_leftovers_on(B, geoms_of(wt_nodes_per_lane()), false);
}

// (one pass over the all-massive buckets (`lean`) or over those with no-mass ends, on the ladder `geoms`)
void wt_merge_leftovers_on(BatchPlan& B, const std::vector<WtGeom>& geoms, bool lean) {
    auto index_of = [&](const ModelPlan& M) {
        for (size_t k = 0; k < geoms.size(); ++k)
            if (geoms[k].G == M.G && geoms[k].NL == M.NL) return (int)k;
        return -1;
    };
    std::map<std::pair<int, int>, std::vector<std::vector<int>>> buckets;   // (variant, S) -> geometry -> models
    for (size_t m = 0; m < B.models.size(); ++m) {
        const ModelPlan& M = B.models[m];
        if (!(M.variant & TILE_WARP) || !(M.variant & TILE_SIMPLE)) continue;   // buildings keep 32 / G >= n_surfaces
        if (lean != !(M.variant & TILE_END)) continue;
        const int k = index_of(M);
        if (k < 0) continue;
        auto& v = buckets[{M.variant & kTileKey, M.dt_subdivisions}];
        if (v.empty()) v.resize(geoms.size());
        v[k].push_back((int)m);
    }
    for (auto& kv : buckets) {
        const bool end = (kv.first.first & TILE_END) != 0;
        if (end && kv.first.second >= kBulkSubsteps && kv.second.size() && [&] {
                size_t c = 0;
                for (auto& v : kv.second) c += v.size();
                return c <= 4096;
            }())
            continue;   // a few latency-bound tiles keep their 8-node geometry
        auto& per = kv.second;
        auto usable = [&](size_t k) { return !(end && geoms[k].G < 2); };
        std::vector<int> carry;
        for (size_t k = 0; k < geoms.size(); ++k) {
            if (!usable(k)) continue;
            std::vector<int> here = per[k];
            for (int m : carry) {
                ModelPlan& M = B.models[m];
                M.G = geoms[k].G;
                M.NL = geoms[k].NL;
                M.walls = 32 / geoms[k].G;
                here.push_back(m);
            }
            carry.clear();
            per[k] = here;
            const int W = 32 / geoms[k].G;
            const int r = (int)here.size() % W;
            bool later = false;
            for (size_t q = k + 1; q < geoms.size(); ++q) later = later || (usable(q) && !per[q].empty());
            if (r > 0 && later && 4 * r <= 3 * W) {
                carry.assign(here.end() - r, here.end());
                per[k].resize(here.size() - r);
            }
        }
    }
}

// dt · λmax(C⁻¹K) over the massive chunks of a wall: what the classical RK4 of surface.rs:228-308 needs below 2.7853
// (its stability interval on the negative real axis) for the march to stay bounded.  K is the chunk's tridiagonal
// conductance matrix (discretization.rs:596-700): solid links between its nodes, plus — on the diagonal only — the link
// to whatever lies beyond either end: the film (`film` W/m²K, convective + linearised IR) at a face of the wall, the
// solid link to a no-mass neighbour, 8 W/m²K for a gas cavity.  C⁻¹K is similar to the symmetric tridiagonal
// (d_i, e_i) = ((K_ii / m_i), u_i / sqrt(m_i m_{i+1})); its largest eigenvalue lies between max d_i and the Gershgorin
// bound and is found by bisection on the Sturm count to 1e-7 relative.
double rk4_number(const SurfacePlan& P, double dt, double film) {
    constexpr double kCavityU = 8.;
    auto link = [&](int i) { return P.u[i] < 0. ? kCavityU : P.u[i]; };   // node i ↔ node i+1
    double worst = 0.;
    std::vector<double> d, e2;
    for (size_t c = 0; c + 1 < P.massive_chunks.size(); c += 2) {
        const int ini = P.massive_chunks[c], fin = P.massive_chunks[c + 1], n = fin - ini;
        if (n <= 0) continue;
        d.assign(n, 0.);
        e2.assign(n, 0.);   // e2[i] = e_i² couples i and i+1
        double lo = 0., hi = 0.;
        for (int i = 0; i < n; ++i) {
            const int g = ini + i;
            const double left = g == 0 ? film : link(g - 1), right = g == P.n_nodes - 1 ? film : link(g);
            d[i] = (left + right) / P.mass[g];
            if (i + 1 < n) e2[i] = link(g) * link(g) / (P.mass[g] * P.mass[g + 1]);
            lo = std::max(lo, d[i]);
        }
        for (int i = 0; i < n; ++i)
            hi = std::max(hi, d[i] + (i ? std::sqrt(e2[i - 1]) : 0.) + (i + 1 < n ? std::sqrt(e2[i]) : 0.));
        // all n eigenvalues are below x  <=>  every pivot of the LDLᵀ factorisation of (x·I − A) is positive
        auto all_below = [&](double x) {
            double q = x - d[0];
            if (q <= 0.) return false;
            for (int i = 1; i < n; ++i) {
                q = x - d[i] - e2[i - 1] / q;
                if (q <= 0.) return false;
            }
            return true;
        };
        while (hi - lo > 1e-7 * hi) {
            const double mid = 0.5 * (lo + hi);
            (all_below(mid) ? hi : lo) = mid;
        }
        worst = std::max(worst, hi);
    }
    return worst * dt;
}

void plan_batch(const hb_batch_desc* d, BatchPlan& B, bool build_tiles) {
    if (!d) fail("null descriptor");
    if (d->steps_per_hour <= 0) fail("steps_per_hour must be positive");
    if (d->n_models <= 0) fail("batch has no models");
    B.layers.assign(d->layers, d->layers + d->n_layers);
    B.construction_offset.assign(d->construction_offset, d->construction_offset + d->n_constructions + 1);
    const int64_t n_surf = d->model_surface_offset[d->n_models];
    const int64_t n_zone = d->model_zone_offset[d->n_models];
    B.surfaces.assign(d->surfaces, d->surfaces + n_surf);
    B.zone_volume.assign(d->zone_volume, d->zone_volume + n_zone);
    B.steps_per_hour = d->steps_per_hour;
    B.model_terrain.assign(d->n_models, HB_TERRAIN_NONE);
    if (d->model_terrain) B.model_terrain.assign(d->model_terrain, d->model_terrain + d->n_models);

    const double main_dt = 60. * 60. / (double)d->steps_per_hour;  // model.rs:240
    B.plans.resize(n_surf);
    B.surface_state_offset.resize(n_surf);
    B.surface_model.resize(n_surf);
    B.models.resize(d->n_models);
    int64_t state = 0;
    B.nodes_total = 0;
    B.updates_per_step = 0;
    for (int m = 0; m < d->n_models; ++m) {
        ModelPlan& M = B.models[m];
        M.zone_begin = d->model_zone_offset[m];
        M.n_zones = d->model_zone_offset[m + 1] - M.zone_begin;
        M.surface_begin = d->model_surface_offset[m];
        M.n_surfaces = d->model_surface_offset[m + 1] - M.surface_begin;
        if (M.n_zones < 0 || M.n_surfaces < 0) fail("offsets must be non-decreasing");
        M.state_offset = state;
        state += M.n_zones;  // one dry-bulb slot per space, zone.rs:43-50
        int32_t tsub = 1;
        bool seen_fenestration = false;
        M.nodes = 0;
        M.nmax = 0;
        for (int k = 0; k < M.n_surfaces; ++k) {
            const int64_t g = M.surface_begin + k;
            const hb_surface& s = B.surfaces[g];
            if (s.is_fenestration) seen_fenestration = true;
            else if (seen_fenestration) fail("within a model, surfaces must precede fenestrations");
            if (s.construction < 0 || s.construction >= d->n_constructions) fail("Construction not found");
            for (int side = 0; side < 2; ++side) {
                const int kind = side ? s.back_kind : s.front_kind;
                const int zone = side ? s.back_zone : s.front_zone;
                if (kind == HB_BOUNDARY_GROUND) fail("not implemented: Boundary::Ground");
                if (kind < 0 || kind > HB_BOUNDARY_GROUND) fail("unknown boundary kind");
                if (kind == HB_BOUNDARY_SPACE && (zone < 0 || zone >= M.n_zones)) fail("Space not found");
            }
            const int l0 = B.construction_offset[s.construction], l1 = B.construction_offset[s.construction + 1];
            SurfacePlan& P = B.plans[g];
            plan_surface(B.layers.data() + l0, l1 - l0, s, B.model_terrain[m], main_dt, P);
            for (size_t c = 0; c + 1 < P.nomass_chunks.size(); c += 2)
                if (P.nomass_chunks[c + 1] - P.nomass_chunks[c] > kMaxNomassChunk)
                    fail("no-mass chunk longer than the supported maximum");
            tsub = std::max(tsub, P.tstep_subdivision);  // model.rs:261-263
            B.surface_state_offset[g] = state;
            B.surface_model[g] = m;
            state += 8 + P.n_nodes;  // surface.rs:428-442
            M.nodes += P.n_nodes;
            M.nmax = std::max(M.nmax, P.n_nodes);
        }
        // model.rs:326-331
        double dt = 60. * 60. / ((double)d->steps_per_hour * (double)tsub);
        dt /= (double)kSafety;
        M.dt_subdivisions = tsub * kSafety;
        M.dt = dt;
        M.state_len = (int32_t)(state - M.state_offset);
        B.nodes_total += M.nodes;
        B.updates_per_step += M.nodes * (int64_t)M.dt_subdivisions;
    }
    B.state_len_total = state;
    if (!build_tiles) return;   // host-only planning handle: discretization and state layout only

    // ---- device ordering --------------------------------------------------------------------------------
    // Every model goes, whole, into one tile (its surfaces couple through its zones every sub-step).  Tiles are
    // uniform in lanes-per-wall G (from the model's largest wall), kernel variant and sub-step count S, so a block
    // runs one loop shape; models are sorted by (G, variant, S) and packed greedily.
    std::vector<int32_t> nmax_real_of(d->n_models, 0);
    std::vector<char> wt_building_ok(d->n_models, 0);   // walls suit the warp-tile kernel's building mode
    int64_t n_end_long = 0;   // walls with no-mass ends and many sub-steps
    for (int m = 0; m < d->n_models; ++m) {
        ModelPlan& M = B.models[m];
        if (M.nmax > kMaxNodesBig) fail("surface has more nodes than the supported maximum (1022)");
        M.variant = 0;
        nmax_real_of[m] = M.nmax;
        wt_building_ok[m] = 0;
        // `rest`: variant of the model without its single panes; a model whose other walls suit the face kernel keeps
        // its panes there too (both faces of a wall in one warp: tiles of at most 16 walls, >= 2 lanes per wall)
        int rest = 0;
        bool ends_only = true;   // every no-mass chunk is one node at an end of a wall of >= 3 nodes (panes aside)
        bool panes = false;
        int rest_wt = 0;          // … and without its double-glazed windows (the warp-tile kernel solves those too)
        for (int k = 0; k < M.n_surfaces; ++k) {
            const SurfacePlan& P = B.plans[M.surface_begin + k];
            int v = 0;
            if (!P.nomass_chunks.empty() || !P.cavities.empty()) v |= TILE_EXT;
            if (!P.cavities.empty()) v |= TILE_CAV;
            if (P.distributed_solar) v |= TILE_EXT | TILE_DIST;
            M.variant |= v;
            if (is_pane(P)) {
                panes = true;
                continue;
            }
            rest |= v;
            if (is_window4(P)) continue;
            rest_wt |= v;
            int nm_nodes = 0;   // no-mass chunks of one or two nodes at the ends, at least one massive node in between
            for (size_t c = 0; c + 1 < P.nomass_chunks.size(); c += 2) {
                const int ini = P.nomass_chunks[c], fin = P.nomass_chunks[c + 1];
                if (fin - ini > 2 || (ini != 0 && fin != P.n_nodes)) ends_only = false;
                nm_nodes += fin - ini;
            }
            if (nm_nodes > 0 && P.n_nodes - nm_nodes < 1) ends_only = false;
        }
        bool plain_faces = true;   // no film-coefficient override, no AmbientTemperature back face
        for (int k = 0; k < M.n_surfaces; ++k) {
            const hb_surface& s = B.surfaces[M.surface_begin + k];
            if (s.back_kind == HB_BOUNDARY_AMBIENT || s.front_hs_override == s.front_hs_override ||
                s.back_hs_override == s.back_hs_override)
                plain_faces = false;
        }
        wt_building_ok[m] = (rest_wt == 0 || rest_wt == TILE_EXT) && ends_only && plain_faces;
        if (panes && (rest == 0 || rest == TILE_EXT) && ends_only) {
            const int nmax2 = std::max<int>(M.nmax, kNL - 1);
            if (walls_for_model(group_of_nodes(nmax2), M.n_surfaces, M.n_zones) <= 16) {
                M.nmax = nmax2;
                M.variant = TILE_END | TILE_PANE;
            }
        } else if (M.variant == TILE_EXT && ends_only) {
            M.variant = TILE_END;
        }
        // the lean face kernels leave out film-coefficient overrides and AmbientTemperature back faces: such models
        // take the general face kernel (which handles massive ends, no-mass ends and panes alike)
        if (M.variant == 0 || (M.variant & TILE_END))
            for (int k = 0; k < M.n_surfaces; ++k) {
                const hb_surface& s = B.surfaces[M.surface_begin + k];
                if (s.back_kind == HB_BOUNDARY_AMBIENT || s.front_hs_override == s.front_hs_override ||
                    s.back_hs_override == s.back_hs_override)
                    M.variant |= TILE_END | TILE_GEN;
            }
        // one wall with its own zone behind it: the zone is advanced in the back-face lane's registers
        if ((M.variant == 0 || M.variant == TILE_END) && M.n_surfaces == 1 && M.n_zones == 1) {
            const hb_surface& s = B.surfaces[M.surface_begin];
            if (s.back_kind == HB_BOUNDARY_SPACE && s.front_kind != HB_BOUNDARY_SPACE) M.variant |= TILE_SIMPLE;
        }
        if (M.variant == (TILE_END | TILE_SIMPLE) && M.dt_subdivisions >= kBulkSubsteps) ++n_end_long;
    }
    for (int m = 0; m < d->n_models; ++m) {
        ModelPlan& M = B.models[m];
        const int nmax_real = nmax_real_of[m];
        M.G = group_of_nodes(M.nmax);
        // A wall with its own zone behind it, all-massive or with no-mass chunks at its ends only: the warp-tile kernel
        // (no block barrier; HB_NO_WTILE=1 keeps the block-tile kernels for A/B measurements)
        static const bool no_wtile = std::getenv("HB_NO_WTILE") != nullptr;
        // HB_WT_NLS="8,12,16" (diagnostic): nodes-per-lane values the planner may use (compiled: 8, 10, 12, 16)
        static const std::vector<int>& wt_nls = wt_nodes_per_lane();
        // Walls with no-mass ends never put both faces (two fixed-point loops in sequence) on one lane, and — they are
        // the walls with the most sub-steps, the longest dependent chains of a series — those with many sub-steps take
        // 8 nodes per lane (measured alone, S = 118: 2550 cycles per sub-step against 3100-3700 with 12 or 16).
        // All-massive walls take the cheapest geometry whatever their sub-step count: 8 nodes per lane does not
        // shorten their sub-step (1690 vs 1700 cycles) and costs 10 % more warp-sub-steps.
        static const std::vector<int> wt_nls_long = {8};
        // (when such walls are a few stragglers that END the series — configs[2]; where they are the bulk of the ensemble —
        // configs[4] — throughput decides and they take the cheapest geometry like everybody else)
        static const int end_long_min_g = std::getenv("HB_END_MIN_G") ? std::atoi(std::getenv("HB_END_MIN_G")) : 8;
        const bool end_long = (M.variant & TILE_END) && M.dt_subdivisions >= kBulkSubsteps && n_end_long <= 4096;
        WtGeom wg{};
        if (!no_wtile && (M.variant == TILE_SIMPLE || M.variant == (TILE_END | TILE_SIMPLE)) &&
            wt_geom_of_nodes(M.nmax, end_long ? wt_nls_long : (M.variant & TILE_END) ? wt_nls : wt_nodes_per_lane_lean(), wg, (M.variant & TILE_END) ? (end_long ? end_long_min_g : 2) : 1)) {
            M.variant |= TILE_WARP;
            M.G = wg.G;
            M.NL = wg.NL;
            M.walls = 32 / wg.G;
            continue;
        }
        // A building with ONE zone behind all its walls (every back face looks at it, no front face looks at a zone):
        // all-massive walls, no-mass ends, single panes — the warp-tile kernel sums the zone over the model's walls with
        // a segmented shuffle reduction.  All walls of the model sit in one warp: 32 / G >= n_surfaces.
        if (!no_wtile && M.n_zones == 1 && M.n_surfaces >= 2 && wt_building_ok[m]) {
            bool ok = true;
            for (int k = 0; k < M.n_surfaces && ok; ++k) {
                const hb_surface& s = B.surfaces[M.surface_begin + k];
                ok = s.back_kind == HB_BOUNDARY_SPACE && s.back_zone == 0 && s.front_kind != HB_BOUNDARY_SPACE;
            }
            WtGeom best{};
            bool found = false;
            if (ok)
                for (int G = 1; G <= 8; G *= 2)
                    for (int NL : wt_nls) {
                        const int cap = G * NL;
                        if (nmax_real > cap || 32 / G < M.n_surfaces) continue;
                        if (!found || cap < best.cap || (cap == best.cap && NL > best.NL)) best = WtGeom{cap, G, NL};
                        found = true;
                    }
            if (found) {
                M.nmax = nmax_real;
                M.variant = TILE_END | TILE_PANE | TILE_WARP;   // one kernel for all such buildings
                M.G = best.G;
                M.NL = best.NL;
                M.walls = 32 / best.G;
                continue;
            }
        }
        const int Wmax = walls_of_group(group_of_nodes(M.nmax));
        if (M.n_surfaces > Wmax || M.n_zones > Wmax)
            fail("model larger than one tile (" + std::to_string(Wmax) + " surfaces / zones at this wall size)");
        // tiles of the face kernel (all-massive / no-mass ends) are as small as the model allows
        const bool face = (M.variant & (TILE_EXT | TILE_DIST)) == 0;
        M.walls = face ? walls_for_model(group_of_nodes(M.nmax), M.n_surfaces, M.n_zones) : Wmax;
        M.NL = slots_of_nodes(M.nmax);
        if (!face && M.NL != kNL)
            fail("surface with no-mass layers between massive ones, cavities or inner solar absorption has more nodes "
                 "than the supported maximum (254)");
        // 16 lanes per wall, few sub-steps (throughput-bound bulk): 16 walls per tile fill the face warp's 32 lanes
        // (the same walls per SM: 2 blocks of 256 threads instead of 4 of 128); long tiles keep the short node phase
        if (face && group_of_nodes(M.nmax) == 16 && M.dt_subdivisions < kBulkSubsteps) M.walls = std::max(M.walls, 16);
    }
    wt_merge_leftovers(B);
    B.model_order.resize(d->n_models);
    std::iota(B.model_order.begin(), B.model_order.end(), 0);
    auto key = [&](int m) {
        const ModelPlan& M = B.models[m];
        return std::make_tuple(M.G, M.NL, M.variant & kTileKey, M.walls, M.dt_subdivisions);
    };
    if (d->n_models > 1)
        std::stable_sort(B.model_order.begin(), B.model_order.end(), [&](int a, int b) { return key(a) < key(b); });

    B.dev_surface_of.assign(n_surf, -1);
    B.dev_zone_of.assign(n_zone, -1);
    B.surface_tile.assign(n_surf, -1);
    B.zone_tile.assign(n_zone, -1);
    B.tiles.clear();
    int64_t node_off = 0, ext_off = 0;
    int32_t surf_slots = 0, zone_slots = 0;
    Tile cur{};
    bool open = false;
    auto close_tile = [&]() {
        if (!open) return;
        cur.node_offset = node_off;
        node_off += (int64_t)cur.NL * cur.G * cur.W;
        cur.ext_offset = -1;
        if (cur.flags & TILE_EXT) {
            cur.ext_offset = ext_off;
            ext_off += (int64_t)cur.nmax * cur.W;
        }
        surf_slots += cur.W;
        zone_slots += cur.W;
        B.tiles.push_back(cur);
        open = false;
    };
    for (int pos = 0; pos < d->n_models; ++pos) {
        const int m = B.model_order[pos];
        const ModelPlan& M = B.models[m];
        const int G = M.G;
        const int W = M.walls;
        const int variant = M.variant & kTileKey;
        if (open && (cur.G != G || cur.NL != M.NL || cur.W != W || (cur.flags & kTileKey) != variant || cur.S != M.dt_subdivisions ||
                     cur.n_surfaces + M.n_surfaces > W || cur.n_zones + M.n_zones > W))
            close_tile();
        if (!open) {
            cur = Tile{};
            cur.surface_begin = surf_slots;
            cur.zone_begin = zone_slots;
            cur.G = G;
            cur.NL = M.NL;
            cur.W = W;
            cur.S = M.dt_subdivisions;
            cur.dt = M.dt;
            cur.flags = variant;
            open = true;
        }
        const int32_t ti = (int32_t)B.tiles.size();
        for (int k = 0; k < M.n_surfaces; ++k) {
            B.dev_surface_of[M.surface_begin + k] = cur.surface_begin + cur.n_surfaces + k;
            B.surface_tile[M.surface_begin + k] = ti;
        }
        for (int k = 0; k < M.n_zones; ++k) {
            B.dev_zone_of[M.zone_begin + k] = cur.zone_begin + cur.n_zones + k;
            B.zone_tile[M.zone_begin + k] = ti;
        }
        cur.n_surfaces += M.n_surfaces;
        cur.n_zones += M.n_zones;
        cur.nmax = std::max(cur.nmax, M.nmax);
        cur.flags |= M.variant;
    }
    close_tile();
    B.node_slab_len = node_off;
    B.ext_len = ext_off;
    B.n_dev_surfaces = surf_slots;
    B.n_dev_zones = zone_slots;
}

// ------------------------------------------------------------------------------------------------------------
// Packing: descriptor-order plans → the SoA buffers the kernels read (DevView layout, hb_dev.hpp).
// ------------------------------------------------------------------------------------------------------------
void pack_arrays(const BatchPlan& B, HostArrays& H) {
    const int64_t NS = B.n_dev_surfaces, NZ = B.n_dev_zones;
    H.n_nodes.assign(NS, 0); H.flags.assign(NS, 0); H.zone_f.assign(NS, -1); H.zone_b.assign(NS, -1);
    H.site.assign(NS, 0); H.desc_index.assign(NS, 0); H.nm_off.assign(NS + 1, 0); H.cav_off.assign(NS + 1, 0);
    H.nm_pairs.clear(); H.cav.clear();
    for (auto* v : {&H.area, &H.nx, &H.ny, &H.cos_tilt, &H.forced_c, &H.eps_f, &H.eps_b, &H.amb_f, &H.amb_b, &H.af0,
                    &H.abL, &H.v0, &H.vL, &H.u_f, &H.u_b, &H.u_f2, &H.u_b2, &H.ax_f, &H.ax_b, &H.wa1f, &H.wa1b, &H.wa2f, &H.wa2b,
                    &H.energy_f, &H.energy_b})
        v->assign(NS, 0.);
    H.hs_ov_f.assign(NS, NAN); H.hs_ov_b.assign(NS, NAN);
    H.nodeT.assign(B.node_slab_len, 0.); H.coefA.assign(B.node_slab_len, 0.); H.coefB.assign(B.node_slab_len, 0.);
    for (auto* v : {&H.ext_u, &H.ext_mass, &H.ext_af, &H.ext_ab}) v->assign(std::max<int64_t>(B.ext_len, 1), 0.);
    H.sc.assign(8 * NS, 0.);
    H.nm_iters.assign(NS, 0ull); H.status.assign(NS, 0);
    H.zoneT.assign(NZ, 0.); H.zvol.assign(NZ, 0.); H.zq.assign(NZ, 0.); H.zinf_v.assign(NZ, 0.); H.zinf_t.assign(NZ, NAN);
    H.zvent_v.assign(NZ, 0.); H.zvent_t.assign(NZ, NAN);
    H.z_surf_begin.assign(NZ, 0); H.z_surf_end.assign(NZ, 0);

    std::vector<int64_t> desc_of(NS, -1);
    for (size_t gd = 0; gd < B.dev_surface_of.size(); ++gd) desc_of[B.dev_surface_of[gd]] = (int64_t)gd;

    for (size_t gd = 0; gd < B.plans.size(); ++gd) {
        const int64_t g = B.dev_surface_of[gd];
        const Tile& tile = B.tiles[B.surface_tile[gd]];
        const int wl = (int)(g - tile.surface_begin);
        const hb_surface& s = B.surfaces[gd];
        const SurfacePlan& P = B.plans[gd];
        const ModelPlan& M = B.models[B.surface_model[gd]];
        const int n = P.n_nodes;
        H.n_nodes[g] = n;
        H.flags[g] = (s.front_kind & 15) | ((s.back_kind & 15) << 4) | ((P.distributed_solar ? 1 : 0) << 8) |
                     ((P.cavities.empty() ? 0 : 1) << 9);
        const bool win4 = (tile.flags & TILE_WARP) && is_window4(P);
        if (win4) {   // solved as a whole by the front face lane: no end-node placement (bits 10 / 11 stay clear)
            H.flags[g] |= 1 << 15;
            H.wa1f[g] = P.front_alpha[1]; H.wa1b[g] = P.back_alpha[1];
            H.wa2f[g] = P.front_alpha[2]; H.wa2b[g] = P.back_alpha[2];
        }
        if ((tile.flags & TILE_END) && !win4)
            for (size_t c = 0; c + 1 < P.nomass_chunks.size(); c += 2)
            {
                const int ini = P.nomass_chunks[c], fin = P.nomass_chunks[c + 1];
                H.flags[g] |= ini == 0 ? (1 << 10) : (1 << 11);
                if (fin - ini == 2 && !is_pane(P)) H.flags[g] |= ini == 0 ? (1 << 13) : (1 << 14);
            }
        const bool pane = (tile.flags & TILE_PANE) && is_pane(P);
        if (pane) H.flags[g] |= (1 << 10) | (1 << 11) | (1 << 12);
        H.u_f[g] = n > 1 ? P.u[0] : 0.;
        H.u_b[g] = n > 1 ? P.u[n - 2] : 0.;
        H.u_f2[g] = n > 2 ? P.u[1] : 0.;
        H.u_b2[g] = n > 2 ? P.u[n - 3] : 0.;
        const int zone_local0 = M.n_zones ? B.dev_zone_of[M.zone_begin] - tile.zone_begin : 0;
        H.zone_f[g] = s.front_kind == HB_BOUNDARY_SPACE ? zone_local0 + s.front_zone : -1;
        H.zone_b[g] = s.back_kind == HB_BOUNDARY_SPACE ? zone_local0 + s.back_zone : -1;
        H.desc_index[g] = (int32_t)gd;
        H.area[g] = s.area;
        H.nx[g] = s.normal[0];
        H.ny[g] = s.normal[1];
        H.cos_tilt[g] = P.cos_tilt;
        H.forced_c[g] = 2.537 * 1.67 * std::sqrt(s.perimeter * P.wind_speed_modifier / s.area);
        H.eps_f[g] = P.front_emissivity;
        H.eps_b[g] = P.back_emissivity;
        H.amb_f[g] = s.front_ambient_t;
        H.amb_b[g] = s.back_ambient_t;
        H.hs_ov_f[g] = s.front_hs_override;
        H.hs_ov_b[g] = s.back_hs_override;
        // scaled tridiagonal coefficients: row i of K times dt/C_i (surface.rs:168-187); no-mass rows are zero
        auto vof = [&](int i) { return P.mass[i] >= kMassiveThreshold ? M.dt / P.mass[i] : 0.0; };
        for (int i = 0; i < n; ++i) {
            const int64_t o = B.slab_index((int64_t)gd, i);
            const double vi = vof(i);
            const double ui = P.u[i];                       // link i → i+1 (−1: cavity, filled in per sub-step)
            const double um = i > 0 ? P.u[i - 1] : 0.0;     // link i−1 → i
            H.nodeT[o] = kInitialT;
            H.coefA[o] = (i + 1 < n && ui > 0. ? ui : 0.) * vi;
            H.coefB[o] = (um > 0. ? um : 0.) * vi;
            if (tile.ext_offset >= 0) {
                const int64_t e = tile.ext_offset + (int64_t)i * tile.W + wl;
                H.ext_u[e] = P.u[i];
                H.ext_mass[e] = P.mass[i];
                H.ext_af[e] = P.front_alpha[i];
                H.ext_ab[e] = P.back_alpha[i];
            }
        }
        H.v0[g] = vof(0);
        H.vL[g] = vof(n - 1);
        if (!P.distributed_solar || pane || win4) {  // distributed: every absorptance goes through the per-node source term
            H.af0[g] = P.front_alpha[0];
            H.abL[g] = P.back_alpha[n - 1];
        }
        if (pane || win4) {
            H.ax_f[g] = P.back_alpha[0];
            H.ax_b[g] = P.front_alpha[n - 1];
        }
        H.sc[0 * NS + g] = kInitialH;
        H.sc[1 * NS + g] = kInitialH;
    }
    // CSR lists in device-slot order
    for (int64_t g = 0; g < NS; ++g) {
        H.nm_off[g] = (int32_t)(H.nm_pairs.size() / 2);
        H.cav_off[g] = (int32_t)H.cav.size();
        const int64_t gd = desc_of[g];
        if (gd < 0) continue;
        const SurfacePlan& P = B.plans[gd];
        H.nm_pairs.insert(H.nm_pairs.end(), P.nomass_chunks.begin(), P.nomass_chunks.end());
        for (const auto& c : P.cavities)
            H.cav.push_back(CavityDev{c.node, c.gas, c.thickness, c.height, c.angle, c.eout, c.ein});
    }
    H.nm_off[NS] = (int32_t)(H.nm_pairs.size() / 2);
    H.cav_off[NS] = (int32_t)H.cav.size();
    if (H.nm_pairs.empty()) H.nm_pairs.push_back(0);
    if (H.cav.empty()) H.cav.push_back(CavityDev{0, 0, 1., 1., 0., 1., 1.});
    // zones
    for (size_t m = 0; m < B.models.size(); ++m) {
        const ModelPlan& M = B.models[m];
        if (M.n_zones == 0) continue;
        const Tile& tile = B.tiles[B.zone_tile[M.zone_begin]];
        const int sb = M.n_surfaces ? B.dev_surface_of[M.surface_begin] - tile.surface_begin : 0;
        for (int z = 0; z < M.n_zones; ++z) {
            const int64_t gz = B.dev_zone_of[M.zone_begin + z];
            H.zoneT[gz] = kInitialT;
            H.zvol[gz] = B.zone_volume[M.zone_begin + z];
            H.z_surf_begin[gz] = sb;
            H.z_surf_end[gz] = sb + M.n_surfaces;
        }
    }
}

DevView HostArrays::view(const BatchPlan& B) {
    DevView V{};
    V.n_nodes = n_nodes.data(); V.flags = flags.data(); V.zone_f = zone_f.data(); V.zone_b = zone_b.data();
    V.site = site.data(); V.desc_index = desc_index.data(); V.area = area.data(); V.nx = nx.data(); V.ny = ny.data();
    V.cos_tilt = cos_tilt.data(); V.forced_c = forced_c.data(); V.eps_f = eps_f.data(); V.eps_b = eps_b.data();
    V.amb_f = amb_f.data(); V.amb_b = amb_b.data(); V.hs_ov_f = hs_ov_f.data(); V.hs_ov_b = hs_ov_b.data();
    V.af0 = af0.data(); V.abL = abL.data(); V.v0 = v0.data(); V.vL = vL.data(); V.u_f = u_f.data(); V.u_b = u_b.data(); V.u_f2 = u_f2.data(); V.u_b2 = u_b2.data(); V.ax_f = ax_f.data(); V.ax_b = ax_b.data();
    V.wa1f = wa1f.data(); V.wa1b = wa1b.data(); V.wa2f = wa2f.data(); V.wa2b = wa2b.data();
    V.nm_off = nm_off.data(); V.nm_pairs = nm_pairs.data(); V.cav_off = cav_off.data(); V.cav = cav.data();
    V.nodeT = nodeT.data(); V.coefA = coefA.data(); V.coefB = coefB.data();
    V.ext_u = ext_u.data(); V.ext_mass = ext_mass.data(); V.ext_af = ext_af.data(); V.ext_ab = ext_ab.data();
    V.sc = sc.data(); V.n_dev_surfaces = B.n_dev_surfaces; V.energy_f = energy_f.data(); V.energy_b = energy_b.data();
    V.nm_iters = nm_iters.data(); V.status = status.data();
    V.zoneT = zoneT.data(); V.zvol = zvol.data(); V.zq = zq.data(); V.zinf_v = zinf_v.data(); V.zinf_t = zinf_t.data();
    V.zvent_v = zvent_v.data(); V.zvent_t = zvent_t.data(); V.z_surf_begin = z_surf_begin.data();
    V.z_surf_end = z_surf_end.data(); V.n_dev_zones = B.n_dev_zones; V.tiles = B.tiles.data();
    return V;
}

void scatter_state(const BatchPlan& B, const double* state, double* nodeT, double* sc, double* zoneT) {
    const int64_t NS = B.n_dev_surfaces;
    for (size_t m = 0; m < B.models.size(); ++m) {
        const ModelPlan& M = B.models[m];
        for (int z = 0; z < M.n_zones; ++z) zoneT[B.dev_zone_of[M.zone_begin + z]] = state[M.state_offset + z];
    }
    for (size_t gd = 0; gd < B.plans.size(); ++gd) {
        const int64_t g = B.dev_surface_of[gd];
        const double* p = state + B.surface_state_offset[gd];
        for (int k = 0; k < 8; ++k) sc[k * NS + g] = p[k];
        for (int i = 0; i < B.plans[gd].n_nodes; ++i) nodeT[B.slab_index((int64_t)gd, i)] = p[8 + i];
    }
}

void gather_state(const BatchPlan& B, const double* nodeT, const double* sc, const double* zoneT, double* state) {
    const int64_t NS = B.n_dev_surfaces;
    for (size_t m = 0; m < B.models.size(); ++m) {
        const ModelPlan& M = B.models[m];
        for (int z = 0; z < M.n_zones; ++z) state[M.state_offset + z] = zoneT[B.dev_zone_of[M.zone_begin + z]];
    }
    for (size_t gd = 0; gd < B.plans.size(); ++gd) {
        const int64_t g = B.dev_surface_of[gd];
        double* p = state + B.surface_state_offset[gd];
        for (int k = 0; k < 8; ++k) p[k] = sc[k * NS + g];
        for (int i = 0; i < B.plans[gd].n_nodes; ++i) p[8 + i] = nodeT[B.slab_index((int64_t)gd, i)];
    }
}

// EPW ingestion.  The reference takes weather through the un-vendored `weather` crate
// (`weather.get_weather_data(date)`, model.rs:372); what it reads of a record is dry bulb, wind direction, wind speed
// (+ horizontal infrared for the sky temperature).  An EPW file is 8 header lines, then one comma-separated record per
// hour; 1-based fields: 7 dry bulb [C], 13 horizontal infrared [W/m2], 14 global horizontal [Wh/m2], 21 wind
// direction [deg], 22 wind speed [m/s] (mapping onto the march inputs: BASELINE.md §3, builder-defined).
// Rows come out in the layouts hb_series consumes: weather [row][3], inputs [row][4] = (GHI, 0, IR, 0).
int64_t read_epw(const char* path, double* weather, double* inputs, int64_t capacity) {
    std::FILE* f = std::fopen(path, "rb");
    if (!f) fail(std::string("EPW: cannot open '") + path + "'");
    std::string line;
    int64_t n = 0, lineno = 0;
    int ch;
    bool eof = false;
    while (!eof) {
        line.clear();
        while ((ch = std::fgetc(f)) != EOF && ch != '\n') line.push_back((char)ch);
        if (ch == EOF) eof = true;
        ++lineno;
        while (!line.empty() && (line.back() == '\r' || line.back() == ' ')) line.pop_back();
        if (lineno <= 8 || line.empty()) continue;   // LOCATION … DATA PERIODS
        double field[22];
        int k = 0;
        const char* p = line.c_str();
        while (k < 22) {
            const char* e = std::strchr(p, ',');
            if (k == 6 || k == 12 || k == 13 || k == 20 || k == 21) {
                char* end = nullptr;
                field[k] = std::strtod(p, &end);
                if (end == p) {
                    std::fclose(f);
                    fail("EPW: line " + std::to_string(lineno) + ": field " + std::to_string(k + 1) + " is not a number");
                }
            }
            ++k;
            if (!e) break;
            p = e + 1;
        }
        if (k < 22) {
            std::fclose(f);
            fail("EPW: line " + std::to_string(lineno) + " has " + std::to_string(k) + " fields, expected at least 22");
        }
        if (n < capacity) {
            if (weather) { weather[3 * n] = field[6]; weather[3 * n + 1] = field[20]; weather[3 * n + 2] = field[21]; }
            if (inputs) { inputs[4 * n] = field[13]; inputs[4 * n + 1] = 0.; inputs[4 * n + 2] = field[12]; inputs[4 * n + 3] = 0.; }
        }
        ++n;
    }
    std::fclose(f);
    return n;
}

}  // namespace hb

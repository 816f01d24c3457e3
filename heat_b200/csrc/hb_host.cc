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
#include <numeric>
#include <stdexcept>

#include "hb_internal.hpp"

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

void plan_batch(const hb_batch_desc* d, BatchPlan& B) {
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
        if (M.n_surfaces > kBlock || M.n_zones > kBlock) fail("model larger than one tile (128 surfaces / zones)");
    }
    B.state_len_total = state;

    // ---- device ordering: sort models so that a warp sees similar node counts and sub-step counts ----
    B.model_order.resize(d->n_models);
    std::iota(B.model_order.begin(), B.model_order.end(), 0);
    if (d->n_models > 1) {
        std::stable_sort(B.model_order.begin(), B.model_order.end(), [&](int a, int b) {
            const ModelPlan &A = B.models[a], &C = B.models[b];
            if (A.n_surfaces != C.n_surfaces) return A.n_surfaces < C.n_surfaces;
            if (A.nmax != C.nmax) return A.nmax < C.nmax;
            return A.dt_subdivisions < C.dt_subdivisions;
        });
    }
    B.dev_surface_of.assign(n_surf, -1);
    B.dev_zone_of.assign(n_zone, -1);
    B.tiles.clear();
    Tile cur{0, 0, 0, 0, 0, 0, 0};
    int64_t node_off = 0;
    auto tier_of = [](int n) { return n <= 8 ? 8 : n <= 16 ? 16 : n <= 24 ? 24 : n <= 32 ? 32 : n <= 64 ? 64 : 256; };
    auto close_tile = [&]() {
        if (cur.n_surfaces == 0 && cur.n_zones == 0) return;
        cur.node_offset = node_off;
        node_off += (int64_t)tier_of(cur.nmax) * kBlock;
        B.tiles.push_back(cur);
        const int32_t t = (int32_t)B.tiles.size();
        cur = Tile{t * kBlock, 0, t * kBlock, 0, 0, 0, 0};
    };
    for (int pos = 0; pos < d->n_models; ++pos) {
        const ModelPlan& M = B.models[B.model_order[pos]];
        const bool tier_change = cur.n_surfaces > 0 && tier_of(M.nmax) != tier_of(cur.nmax);
        if (cur.n_surfaces + M.n_surfaces > kBlock || cur.n_zones + M.n_zones > kBlock || tier_change) close_tile();
        for (int k = 0; k < M.n_surfaces; ++k) B.dev_surface_of[M.surface_begin + k] = cur.surface_begin + cur.n_surfaces + k;
        for (int k = 0; k < M.n_zones; ++k) B.dev_zone_of[M.zone_begin + k] = cur.zone_begin + cur.n_zones + k;
        cur.n_surfaces += M.n_surfaces;
        cur.n_zones += M.n_zones;
        cur.nmax = std::max(cur.nmax, M.nmax);
        cur.smax = std::max(cur.smax, M.dt_subdivisions);
    }
    close_tile();
    B.node_slab_len = node_off;
    B.n_dev_surfaces = (int64_t)B.tiles.size() * kBlock;
    B.n_dev_zones = (int64_t)B.tiles.size() * kBlock;
}

}  // namespace hb

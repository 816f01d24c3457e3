/*
 * heat_oracle.cc — CPU ORACLE (test infrastructure, see heat_oracle.h header comment).
 *
 * Operation-for-operation restatement, in C++17 / f64 / no FMA contraction
 * (build with -ffp-contract=off), of the wall-conduction time-stepping path of
 * SIMPLE-BuildingSimulation/heat.  Every function cites the reference file:line it
 * follows (paths relative to /root/reference).  Structure is deliberately faithful —
 * dense K and C matrices re-zeroed per call, per-iteration clones, four evaluations of
 * the border conditions per sub-step — so that the same code doubles as the timed
 * "reference CPU implementation" baseline.
 *
 * (inferred) marks semantics of un-vendored crates (`matrix`, `polynomial`,
 * `simple_model`, `geometry3d`) that are restated from their call sites and pinned
 * only through the reference's own tests.
 */
#include "heat_oracle.h"

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

namespace {

using Float = double;

// src/lib.rs:41-49
constexpr Float PI = 3.14159265358979323846264338327950288;
constexpr Float SIGMA = 5.670374419e-8;
// src/discretization.rs:21
constexpr Float MAX_RS = 0.05;
// src/convection.rs:22
constexpr Float MIN_H = 0.1;
// src/surface.rs:449, src/discretization.rs:265
constexpr Float DEFAULT_EM = 0.84;

thread_local std::string g_error;

struct Panic : std::runtime_error {
    using std::runtime_error::runtime_error;
};

// LLVM `powi`: repeated multiplication. powi(3) = (x*x)*x, powi(4) = (x*x)*(x*x),
// powi(7) = binary exponentiation as compiler-rt __powidf2 does.
inline Float powi(Float a, int b) {
    // compiler-rt/lib/builtins/powidf2.c
    const bool recip = b < 0;
    Float r = 1;
    while (true) {
        if (b & 1) r *= a;
        b /= 2;
        if (b == 0) break;
        a *= a;
    }
    return recip ? 1 / r : r;
}

inline Float to_radians(Float deg) {
    // core::f64::to_radians: self * (PI / 180.0)
    return deg * (PI / 180.0);
}

/* ------------------------------------------------------------------------------------------
 * `matrix` crate (git dep, un-vendored, unpinned; Cargo.toml:16) — minimal dense Matrix.
 * (inferred) from the call sites in src/surface.rs and src/discretization.rs.
 * ---------------------------------------------------------------------------------------- */
struct Matrix {
    size_t nrows = 0, ncols = 0;
    std::vector<Float> data;

    Matrix() = default;
    Matrix(Float v, size_t r, size_t c) : nrows(r), ncols(c), data(r * c, v) {}
    static Matrix from_data(size_t r, size_t c, std::vector<Float> d) {
        Matrix m;
        m.nrows = r;
        m.ncols = c;
        m.data = std::move(d);
        return m;
    }
    bool in_range(size_t r, size_t c) const { return r < nrows && c < ncols; }
    Float get(size_t r, size_t c) const {
        if (!in_range(r, c)) throw Panic("Matrix index out of bounds");
        return data[r * ncols + c];
    }
    void set(size_t r, size_t c, Float v) {
        if (!in_range(r, c)) throw Panic("Matrix index out of bounds");
        data[r * ncols + c] = v;
    }
    void add_to_element(size_t r, size_t c, Float v) { set(r, c, get(r, c) + v); }
    void scale_element(size_t r, size_t c, Float v) { set(r, c, get(r, c) * v); }
    void scale(Float s) {  // `*= s`
        for (auto& x : data) x *= s;
    }
    void div(Float s) {  // `/= s`
        for (auto& x : data) x /= s;
    }
    void add(const Matrix& o) {  // `+= &o`
        for (size_t i = 0; i < data.size(); ++i) data[i] += o.data[i];
    }
    void scale_into(Float s, Matrix& into) const {
        for (size_t i = 0; i < data.size(); ++i) into.data[i] = data[i] * s;
    }
    void copy_from(const Matrix& o) { data = o.data; }

    // (inferred) Matrix::prod_tri_diag_into — row r = Σ_{c=r-1..r+1} K[r][c]·x[c], accumulated from 0.0
    // in ascending column order.  Call sites: src/surface.rs:268,280,286,292.
    void prod_tri_diag_into(const Matrix& x, Matrix& into) const {
        for (size_t r = 0; r < nrows; ++r) {
            const size_t ini = r == 0 ? 0 : r - 1;
            const size_t fin = r + 1 >= ncols ? ncols - 1 : r + 1;
            Float v = 0.0;
            for (size_t c = ini; c <= fin; ++c) v += data[r * ncols + c] * x.data[c];
            into.data[r] = v;
        }
    }

    // (inferred) Matrix::mut_n_diag_gaussian(b, n) — no-pivot Gaussian elimination restricted to the
    // n-diagonal band, then back substitution.  Call site: src/surface.rs:834 (n = 3).
    Matrix mut_n_diag_gaussian(Matrix b, size_t n) {
        const size_t half = n / 2;
        const size_t N = nrows;
        for (size_t c = 0; c < N; ++c) {
            const Float pivot = data[c * ncols + c];
            const size_t rmax = std::min(N, c + 1 + half);
            for (size_t r = c + 1; r < rmax; ++r) {
                const Float factor = data[r * ncols + c] / pivot;
                const size_t cmax = std::min(ncols, c + 1 + half);
                for (size_t cc = c; cc < cmax; ++cc) data[r * ncols + cc] -= factor * data[c * ncols + cc];
                b.data[r] -= factor * b.data[c];
            }
        }
        for (size_t ri = N; ri-- > 0;) {
            Float v = b.data[ri];
            const size_t cmax = std::min(ncols, ri + 1 + half);
            for (size_t cc = ri + 1; cc < cmax; ++cc) v -= data[ri * ncols + cc] * b.data[cc];
            b.data[ri] = v / data[ri * ncols + ri];
        }
        return b;
    }
};

/* ------------------------------------------------------------------------------------------
 * src/gas.rs
 * ---------------------------------------------------------------------------------------- */
struct Gas {
    // `polynomial::Polynomial` (un-vendored): coefficients in ascending powers, eval by
    // (inferred) Σ c_i·x^i.  Only degree ≤ 1 is used: c0 + c1*x.
    Float lambda[2], mu[2], cp[2];
    int cp_terms;
    Float mass;

    Float thermal_conductivity(Float t) const { return lambda[0] + lambda[1] * t; }  // gas.rs:155
    Float dynamic_viscosity(Float t) const { return mu[0] + mu[1] * t; }             // gas.rs:160
    Float heat_capacity(Float t) const { return cp_terms == 1 ? cp[0] : cp[0] + cp[1] * t; }  // gas.rs:165
    Float density(Float t) const {                                                   // gas.rs:175-179
        const Float R = 8314.46261815324;
        return 101325. * mass / (R * t);
    }
};
// src/gas.rs:45-74
const Gas AIR = {{2.873e-3, 7.760e-5}, {3.723e-6, 4.94e-8}, {1002.7370, 1.2324e-2}, 2, 28.97};
const Gas ARGON = {{2.285e-3, 5.149e-5}, {3.379e-6, 6.451e-8}, {521.9285, 0.}, 1, 39.948};
const Gas KRYPTON = {{9.443e-4, 2.826e-5}, {2.213e-6, 7.777e-8}, {248.0907, 0.}, 1, 83.8};
const Gas XENON = {{4.538e-4, 1.723e-5}, {1.069e-6, 7.414e-8}, {158.3397, 0.}, 1, 131.30};

const Gas& gas_by_id(int id) {
    switch (id) {
        case HO_GAS_AIR: return AIR;
        case HO_GAS_ARGON: return ARGON;
        case HO_GAS_KRYPTON: return KRYPTON;
        case HO_GAS_XENON: return XENON;
    }
    throw Panic("Substance does not have a standard gas.");
}

inline Float in_kelvin(Float t) { return t + 273.15; }  // gas.rs:183-185

// gas.rs:285-309
Float nu_90(Float ra, Float a_gi) {
    Float nu1;
    if (ra <= 1e4) {
        nu1 = 1. + 1.7596678 * 1e-10 * std::pow(ra, 2.2984755);
    } else if (ra < 5e4) {
        nu1 = 0.028154 * std::pow(ra, 0.4134);
    } else if (ra > 5e4) {
        nu1 = 0.0673838 * std::pow(ra, 1. / 3.);
    } else {
        throw Panic("unreachable (nu_90)");
    }
    const Float nu2 = 0.242 * std::pow(ra / a_gi, 0.272);
    return nu1 > nu2 ? nu1 : nu2;
}
// gas.rs:249-264
Float nu_60(Float ra, Float a_gi) {
    const Float g = 0.5 / std::pow(1. + std::pow(ra / 3160., 20.6), 0.1);
    const Float nu1 = std::pow(1. + powi(0.0936 * std::pow(ra, 0.314) / (1. + g), 7), 1. / 7.);
    const Float nu2 = (0.104 + 0.175 / a_gi) * std::pow(ra, 0.283);
    return nu1 > nu2 ? nu1 : nu2;
}
// gas.rs:227-243
Float nu_0_60(Float ra, Float gamma, Float /*a_gi*/) {
    auto aux = [](Float x) { return (x + std::fabs(x)) / 2.; };
    const Float cos_gamma = std::cos(gamma);
    const Float a = aux(1. - 1708. / (ra * cos_gamma));
    const Float b = 1. - 1708. * std::pow(std::sin(1.8 * gamma), 1.6) / (ra * cos_gamma);
    const Float c = std::pow(ra * cos_gamma / 5830., 1. / 3.) - 1.;
    return 1. + 1.44 * a * b + aux(c);
}
// gas.rs:269-280
Float nu_60_90(Float ra, Float gamma, Float a_gi) {
    const Float nu60 = nu_60(ra, a_gi);
    const Float nu90 = nu_90(ra, a_gi);
    const Float x = (gamma - PI / 3.) / (PI / 2. - PI / 3.);
    return nu60 + (nu90 - nu60) * x;
}
// gas.rs:312-315
Float nu_90_180(Float ra, Float a_gi, Float gamma) {
    const Float nu_v = nu_90(ra, a_gi);
    return 1. + (nu_v - 1.) * std::sin(gamma);
}
// gas.rs:197-221
Float nusselt(Float ra, Float gamma_in, Float a_gi) {
    const Float THIRTY_RAD = 30. * PI / 180.;
    const Float EPSILON_RAD = 0.5 * PI / 180.;
    const Float gamma = std::fmod(gamma_in, PI);
    if (gamma >= 0.0 && gamma < 2. * THIRTY_RAD - EPSILON_RAD) {
        return nu_0_60(ra, gamma, a_gi);
    } else if (gamma < 2. * THIRTY_RAD + EPSILON_RAD) {
        return nu_60(ra, a_gi);
    } else if (gamma < 3. * THIRTY_RAD - EPSILON_RAD) {
        return nu_60_90(ra, gamma, a_gi);
    } else if (gamma < 3. * THIRTY_RAD + EPSILON_RAD) {
        return nu_90(ra, a_gi);
    } else if (gamma < 6. * THIRTY_RAD) {
        return nu_90_180(ra, a_gi, gamma);
    }
    throw Panic("unreachable (nusselt gamma)");
}
// gas.rs:82-102
Float raleigh(const Gas& g, Float t_front, Float t_back, Float thickness) {
    const Float G = 9.81;
    if (std::fabs(t_front - t_back) < 1e-10) return 0.0000001;
    const Float temp = (in_kelvin(t_front) + in_kelvin(t_back)) / 2.;
    const Float beta = 1. / temp;
    const Float c_p = g.heat_capacity(temp);
    const Float mu = g.dynamic_viscosity(temp);
    const Float lambda = g.thermal_conductivity(temp);
    const Float rho = g.density(temp);
    return powi(rho, 2) * powi(thickness, 3) * G * beta * c_p * std::fabs(t_front - t_back) / (mu * lambda);
}
// gas.rs:126-152
Float cavity_convection(const Gas& g, Float height, Float thickness, Float gamma, Float t_front, Float t_back) {
    if (t_front > t_back) gamma = to_radians(180.) - gamma;
    const Float a_gi = height / thickness;
    const Float ra = raleigh(g, t_front, t_back, thickness);
    const Float nu = nusselt(ra, gamma, a_gi);
    const Float temp = (in_kelvin(t_front) + in_kelvin(t_back)) / 2.;
    const Float lambda = g.thermal_conductivity(temp);
    return nu * lambda / thickness;
}

/* src/cavity.rs:28-69 */
struct Cavity {
    Float thickness, height;
    const Gas* gas;
    Float eout, ein, angle;
    Float u_value(Float t_front, Float t_back) const {
        const Float conv = cavity_convection(*gas, height, thickness, angle, t_front, t_back);
        const Float tm = (t_back + t_front) / 2. + 273.15;
        const Float rad = 4. * powi(tm, 3) * SIGMA * ein * eout / (1. - (1. - ein) * (1. - eout));
        return rad + conv;
    }
};

/* src/convection.rs:27-168 */
struct ConvectionParams {
    Float air_temperature, air_speed, rad_temperature, surface_temperature;
    size_t roughness_index;
    Float cos_surface_tilt;

    Float natural() const {  // :87-110
        const Float delta_t = air_temperature - surface_temperature;
        const Float abs_delta_t = std::fabs(delta_t);
        Float h;
        if (std::fabs(delta_t) < 1e-3 || std::fabs(cos_surface_tilt) < 1e-3) {
            h = 1.31 * std::pow(abs_delta_t, 1. / 3.);
        } else if ((delta_t < 0. && cos_surface_tilt < 0.) || (delta_t > 0. && cos_surface_tilt > 0.)) {
            h = 9.482 * std::pow(abs_delta_t, 1. / 3.) / (7.238 - std::fabs(cos_surface_tilt));
        } else if ((delta_t > 0. && cos_surface_tilt < 0.) || (delta_t < 0. && cos_surface_tilt > 0.)) {
            h = 1.81 * std::pow(abs_delta_t, 1. / 3.) / (1.382 + std::fabs(cos_surface_tilt));
        } else {
            throw Panic("unreachable (TARP natural; NaN input?)");
        }
        return h < MIN_H ? MIN_H : h;
    }
    Float tarp(Float area, Float perimeter, bool windward) const {  // :151-168
        static const Float COEFFICIENTS[6] = {2.17, 1.67, 1.52, 1.13, 1.11, 1.};
        const Float rf = COEFFICIENTS[roughness_index];
        const Float wf = windward ? 1.0 : 0.5;
        const Float forced = 2.537 * wf * rf * std::sqrt(perimeter * air_speed / area);
        const Float nat = natural();
        return forced + nat;
    }
};

/* src/surface.rs:37-46 */
bool is_windward(Float wind_direction, Float cos_tilt, const Float normal[3]) {
    if (std::fabs(cos_tilt) < 0.98) {
        const Float wx = std::sin(wind_direction), wy = std::cos(wind_direction), wz = 0.0;
        // geometry3d Vector3D dot product (inferred): x*x + y*y + z*z
        return normal[0] * wx + normal[1] * wy + normal[2] * wz > 0.0;
    }
    return true;
}

/* src/surface.rs:135-166 */
Float wind_speed_modifier(Float height, int terrain) {
    if (height < 1e-5) return 0.0;
    Float alpha = 0.0, delta = 0.0;
    switch (terrain) {
        case HO_TERRAIN_COUNTRY: alpha = 0.14; delta = 270.; break;
        case HO_TERRAIN_SUBURBS: alpha = 0.22; delta = 370.; break;
        case HO_TERRAIN_CITY: alpha = 0.33; delta = 460.; break;
        case HO_TERRAIN_OCEAN: alpha = 0.10; delta = 210.; break;
        case HO_TERRAIN_URBAN: alpha = 0.22; delta = 370.; break;
        case HO_TERRAIN_UNSET: break;  // site details without terrain: alpha = delta = 0
        default: alpha = 0.22; delta = 370.; break;  // no site details → Urban
    }
    return std::pow(270. / 10., 0.14) * std::pow(height / delta, alpha);
}

/* src/glazing.rs:31-286 */
struct Glazing {
    Float tau, rho_front, rho_back, alpha_front, alpha_back;
    static Glazing make(Float tau, Float rho_front, Float rho_back) {  // :52-65
        if (!(tau >= 0.0 && tau <= 1. && rho_front >= 0.0 && rho_back >= 0.0 && rho_front <= 1. && rho_back <= 1.))
            throw Panic("Glazing::new assertion failed");
        return Glazing{tau, rho_front, rho_back, 1. - tau - rho_front, 1. - tau - rho_back};
    }
    Float combined_tau(const Glazing& o) const { return tau * o.tau / (1. - rho_back * o.rho_front); }  // :180
    Float combined_rho_front(const Glazing& o) const {                                                   // :191
        return rho_front + powi(tau, 2) * o.rho_front / (1. - rho_back * o.rho_front);
    }
    Float combined_rho_back(const Glazing& o) const {                                                    // :204
        return o.rho_back + powi(o.tau, 2) * rho_back / (1. - o.rho_front * rho_back);
    }
    Glazing combine(const Glazing& o) const {  // :213-218
        const Float rb = combined_rho_back(o);
        const Float rf = combined_rho_front(o);
        const Float t = combined_tau(o);
        return make(t, rf, rb);
    }
    static Glazing combine_layers(const Glazing* l, size_t n) {  // :221-230
        if (n == 1) return l[0];
        const Glazing rest = combine_layers(l + 1, n - 1);
        return l[0].combine(rest);
    }
    void combined_alphas(const Glazing& o, Float& a1, Float& a2) const {  // :247-252
        const Float denom = 1. - rho_back * o.rho_front;
        a1 = alpha_front + alpha_back * tau * o.rho_front / denom;
        a2 = o.alpha_front * tau / denom;
    }
    static std::vector<Float> alphas(const std::vector<Glazing>& layers) {  // :259-286
        std::vector<Float> ret;
        if (layers.empty()) return ret;
        if (layers.size() == 1) {
            ret.push_back(layers[0].alpha_front);
            return ret;
        }
        Float acc_alpha = 0.0;
        for (size_t i = 1; i < layers.size(); ++i) {
            const Glazing g0 = combine_layers(layers.data(), i);
            const Glazing g1 = combine_layers(layers.data() + i, layers.size() - i);
            Float a0, a1;
            g0.combined_alphas(g1, a0, a1);
            ret.push_back(a0 - acc_alpha);
            acc_alpha = a0;
        }
        const Glazing g0 = combine_layers(layers.data(), layers.size() - 1);
        Float a0, a1;
        g0.combined_alphas(layers.back(), a0, a1);
        ret.push_back(a1);
        return ret;
    }
};

inline Float or_default(Float v, Float d) { return std::isnan(v) ? d : v; }

// src/glazing.rs:67-112 (get_glazing_from_iter).  `order` lists layer indices front→back or back→front.
std::vector<Glazing> glazing_from_iter(const ho_layer* layers, const std::vector<int>& order) {
    std::vector<Glazing> ret;
    size_t i = 0;
    while (true) {
        if (i >= order.size()) throw Panic("called `Option::unwrap()` on a `None` value (glazing iterator)");
        const ho_layer& s = layers[order[i++]];
        if (s.kind == HO_LAYER_GAS) throw Panic("NOT expecting a gas");
        const Float tau = or_default(s.solar_transmittance, 0.0);
        // NOTE (quirk kept): the *front* absorbtance is used as alpha_front even when iterating from the back.
        const Float alpha_front = or_default(s.front_solar_abs, 0.84);
        const Float alpha_back = or_default(s.back_solar_abs, 0.84);
        const Float rho_front = 1. - tau - alpha_front;
        const Float rho_back = 1. - tau - alpha_back;
        ret.push_back(Glazing::make(tau, rho_front, rho_back));
        if (tau < 1e-9) break;
        if (i < order.size()) {
            const ho_layer& g = layers[order[i++]];
            if (g.kind == HO_LAYER_SOLID) throw Panic("Expecting a Gas");
        } else {
            break;
        }
    }
    return ret;
}

/* ------------------------------------------------------------------------------------------
 * src/discretization.rs
 * ---------------------------------------------------------------------------------------- */
struct UValue {
    enum Kind { None, Solid, CavityK, Back } kind = None;
    Float u = 0;
    Cavity cavity{};
    Float u_value(Float t_before, Float t_after) const {  // :48-55
        switch (kind) {
            case Solid: return u;
            case CavityK: return cavity.u_value(t_before, t_after);
            case Back: return 0.;
            default: throw Panic("Attempting to get the u-value of None");
        }
    }
};

struct ChunkMemory {  // src/surface.rs:51-87
    Matrix temps, aux, k, c, q, k1, k2, k3, k4;
    ChunkMemory(size_t ini, size_t fin) {
        const size_t n = fin - ini - 1;
        aux = Matrix(0.0, n + 1, 1);
        k = Matrix(0.0, n + 1, n + 1);
        c = Matrix(0.0, n + 1, n + 1);
        q = Matrix(0.0, n + 1, 1);
        temps = Matrix(0.0, n + 1, 1);
        k1 = Matrix(0.0, n + 1, 1);
        k2 = Matrix(0.0, n + 1, 1);
        k3 = Matrix(0.0, n + 1, 1);
        k4 = Matrix(0.0, n + 1, 1);
    }
    ChunkMemory() = default;
};

struct Discretization {
    std::vector<std::pair<Float, UValue>> segments;
    size_t tstep_subdivision = 1;
    std::vector<size_t> n_elements;

    // :410-544 — the recursion `aux(n) → aux(n+1)` restarts the whole construction; restated as a loop.
    // PARITY UNPINNED: no reference test asserts these integers.
    static void discretize_construction(const ho_layer* layers, size_t n_layers, Float main_dt, Float max_dx,
                                        Float min_dt, size_t& out_n, std::vector<size_t>& out_elements) {
        size_t n = 1;
        while (true) {
            const Float dt = main_dt / (Float)n;
            std::vector<size_t> n_elements;
            bool restart = false;
            for (size_t n_layer = 0; n_layer < n_layers; ++n_layer) {
                const ho_layer& L = layers[n_layer];
                if (L.kind == HO_LAYER_GAS) {
                    n_elements.push_back(0);
                    continue;
                }
                const Float thickness = L.thickness;
                const Float k = L.conductivity, rho = L.density, cp = L.specific_heat;
                const Float a_coef = 1.;
                const Float b_coef = -dt / (rho * cp * MAX_RS);
                const Float c_coef = -2. * dt * k / (rho * cp);
                const Float disc = b_coef * b_coef - 4. * a_coef * c_coef;
                const Float min_dx = (-b_coef + std::sqrt(disc)) / (2. * a_coef);
                if (min_dx > thickness) {
                    const Float next_dt = main_dt / (Float)(n + 1);
                    if (next_dt > min_dt) {
                        restart = true;
                        break;
                    } else {
                        n_elements.push_back(0);
                    }
                } else {
                    const Float m = std::floor(thickness / min_dx);
                    const Float dx = thickness / m;
                    if (dx > max_dx) {
                        const Float next_dt = main_dt / (Float)(n + 1);
                        if (next_dt > min_dt) {
                            restart = true;
                            break;
                        } else {
                            n_elements.push_back(0);
                        }
                    } else {
                        n_elements.push_back((size_t)m);
                    }
                }
            }
            if (restart) {
                ++n;
                continue;
            }
            out_n = n;
            out_elements = n_elements;
            return;
        }
    }

    // :163-298
    static Discretization build(const ho_layer* layers, size_t n_layers, size_t tstep_subdivision,
                                std::vector<size_t> n_elements, Float height, Float angle) {
        size_t n_nodes = 0;
        for (auto x : n_elements) n_nodes += x;
        n_nodes += (size_t)std::count(n_elements.begin(), n_elements.end(), (size_t)0) + 1;
        std::vector<std::pair<Float, UValue>> segments(n_nodes);

        size_t n_segment = 0;
        for (size_t n_layer = 0; n_layer < n_elements.size(); ++n_layer) {
            size_t n = n_elements[n_layer];
            const ho_layer& L = layers[n_layer];
            Float mass;
            if (n == 0) {
                mass = 0.0;
            } else if (L.kind == HO_LAYER_SOLID) {
                const Float dx = L.thickness / (Float)n;
                mass = L.density * L.specific_heat * dx;
            } else {
                mass = 0.0;
            }
            if (n == 0) n = 1;
            for (size_t e = 0; e < n; ++e) {
                if (L.kind == HO_LAYER_SOLID) {
                    segments[n_segment].first += mass / 2.;
                    segments[n_segment + 1].first += mass / 2.;
                    const Float dx = L.thickness / (Float)n;
                    segments[n_segment].second.kind = UValue::Solid;
                    segments[n_segment].second.u = L.conductivity / dx;
                } else {
                    const Gas& gas = gas_by_id(L.gas);
                    if (n_layer == 0) throw Panic("Construction has a Gas as its first layer");
                    if (n_layer + 1 >= n_layers) throw Panic("Construction has a Gas as its last layer");
                    const ho_layer& next = layers[n_layer + 1];
                    const ho_layer& prev = layers[n_layer - 1];
                    if (next.kind == HO_LAYER_GAS || prev.kind == HO_LAYER_GAS)
                        throw Panic("Construction has two gases without a solid layer between them");
                    const Float ein = or_default(next.front_emissivity, DEFAULT_EM);
                    const Float eout = or_default(prev.back_emissivity, DEFAULT_EM);
                    Cavity c{L.thickness, height, &gas, eout, ein, angle};
                    segments[n_segment].second.kind = UValue::CavityK;
                    segments[n_segment].second.cavity = c;
                }
                n_segment += 1;
            }
            segments[n_nodes - 1].second = UValue{};
            segments[n_nodes - 1].second.kind = UValue::Back;
        }
        Discretization d;
        d.segments = std::move(segments);
        d.tstep_subdivision = tstep_subdivision;
        d.n_elements = std::move(n_elements);
        return d;
    }

    // :95-114
    static Discretization make(const ho_layer* layers, size_t n_layers, Float model_dt, Float max_dx, Float min_dt,
                               Float height, Float angle) {
        size_t n;
        std::vector<size_t> el;
        discretize_construction(layers, n_layers, model_dt, max_dx, min_dt, n, el);
        return build(layers, n_layers, n, el, height, angle);
    }

    // :117-137
    static std::vector<std::pair<size_t, size_t>> chunk_segments(const std::vector<size_t>& indexes) {
        std::vector<std::pair<size_t, size_t>> ret;
        if (indexes.empty()) return ret;
        size_t start = indexes[0], prev = start;
        for (size_t j = 1; j < indexes.size(); ++j) {
            const size_t i = indexes[j];
            if (i - prev == 1) {
                prev = i;
            } else {
                ret.emplace_back(start, prev + 1);
                start = i;
                prev = start;
            }
        }
        ret.emplace_back(start, prev + 1);
        return ret;
    }
    // :144-160
    void get_chunks(std::vector<std::pair<size_t, size_t>>& mass, std::vector<std::pair<size_t, size_t>>& nomass) const {
        std::vector<size_t> mass_nodes, nomass_nodes;
        for (size_t i = 0; i < segments.size(); ++i) {
            if (segments[i].first >= 1e-5) mass_nodes.push_back(i);
            if (segments[i].first < 1e-5) nomass_nodes.push_back(i);
        }
        mass = chunk_segments(mass_nodes);
        nomass = chunk_segments(nomass_nodes);
    }

    // :305-319
    Float r_value() const {
        Float r = 0.0;
        for (auto& s : segments) {
            switch (s.second.kind) {
                case UValue::Solid: r += 1. / s.second.u; break;
                case UValue::Back: r += 0.0; break;
                default: throw Panic("r_value: todo!/unreachable");
            }
        }
        return r;
    }

    // :596-700
    void get_k_q(size_t ini, size_t fin, const Matrix& temperatures, const ConvectionParams& front_env, Float front_hs,
                 Float front_rad_hs, const ConvectionParams& back_env, Float back_hs, Float back_rad_hs,
                 ChunkMemory& memory) const {
        const size_t nrows = temperatures.nrows;
        if (temperatures.ncols != 1 || nrows != segments.size()) throw Panic("get_k_q: bad temperatures shape");
        memory.k.scale(0.0);
        memory.q.scale(0.0);
        const size_t nnodes = fin - ini;

        auto get_t_after = [&](size_t i) -> Float {
            if (i + 1 < nrows) return temperatures.data[i + 1];
            return back_env.air_temperature;
        };

        for (size_t local_i = 0; local_i + 1 < nnodes; ++local_i) {
            const size_t global_i = ini + local_i;
            const Float t_this = temperatures.get(global_i, 0);
            const Float t_next = get_t_after(global_i);
            const Float u = segments[global_i].second.u_value(t_this, t_next);
            memory.k.add_to_element(local_i, local_i, -u);
            if (memory.k.in_range(local_i + 1, local_i + 1)) {
                const Float old = memory.k.get(local_i + 1, local_i + 1);
                memory.k.set(local_i + 1, local_i + 1, old - u);
            }
            if (memory.k.in_range(local_i, local_i + 1)) {
                const Float old = memory.k.get(local_i, local_i + 1);
                memory.k.set(local_i, local_i + 1, old + u);
            }
            if (memory.k.in_range(local_i + 1, local_i)) {
                const Float old = memory.k.get(local_i + 1, local_i);
                memory.k.set(local_i + 1, local_i, old + u);
            }
        }

        Float hs_front, front_q;
        if (ini == 0) {
            const Float ts = temperatures.get(0, 0);
            front_q = front_env.air_temperature * front_hs + front_rad_hs * (front_env.rad_temperature - ts);
            hs_front = front_hs;
        } else {
            const Float t_before = temperatures.get(ini - 1, 0);
            const Float t_after = temperatures.get(ini, 0);
            const Float u = segments[ini - 1].second.u_value(t_before, t_after);
            hs_front = u;
            front_q = u * t_before;
        }
        memory.q.add_to_element(0, 0, front_q);
        memory.k.add_to_element(0, 0, -hs_front);

        Float hs_back, back_q;
        if (fin == nrows) {
            const Float ts = temperatures.get(fin - 1, 0);
            back_q = back_env.air_temperature * back_hs + back_rad_hs * (back_env.rad_temperature - ts);
            hs_back = back_hs;
        } else {
            const Float t_before = temperatures.get(fin - 1, 0);
            const Float t_after = get_t_after(fin - 1);
            const Float u = segments[fin - 1].second.u_value(t_before, t_after);
            hs_back = u;
            back_q = u * t_after;
        }
        memory.q.add_to_element(nnodes - 1, 0, back_q);
        memory.k.add_to_element(nnodes - 1, nnodes - 1, -hs_back);
    }
};

/* ------------------------------------------------------------------------------------------
 * src/surface.rs
 * ---------------------------------------------------------------------------------------- */
// :168-187
void rearrange_k(Float dt, ChunkMemory& memory) {
    const size_t crows = memory.c.nrows;
    for (size_t nrow = 0; nrow < crows; ++nrow) {
        const Float v = dt / memory.c.get(nrow, nrow);
        const size_t ini = nrow == 0 ? 0 : nrow - 1;
        const size_t fin = nrow == crows - 1 ? crows - 1 : nrow + 1;
        for (size_t ncol = ini; ncol <= fin; ++ncol) memory.k.scale_element(nrow, ncol, v);
        memory.q.scale_element(nrow, 0, v);
    }
}

// :228-308
void rk4(ChunkMemory& memory) {
    if (memory.k.nrows != memory.k.ncols) throw Panic("Expecting 'K' to be a squared matrix");
    if (memory.c.nrows != memory.c.ncols) throw Panic("Expecting 'C' to be a squared matrix");
    if (memory.q.nrows != memory.k.nrows || memory.q.ncols != 1) throw Panic("bad 'q' shape");
    memory.k1.scale(0.0);
    memory.k2.scale(0.0);
    memory.k3.scale(0.0);
    memory.k4.scale(0.0);
    memory.aux.scale(0.0);

    memory.k.prod_tri_diag_into(memory.temps, memory.k1);
    memory.k1.add(memory.q);

    memory.k1.scale_into(0.5, memory.aux);
    memory.aux.add(memory.temps);

    memory.k.prod_tri_diag_into(memory.aux, memory.k2);
    memory.k2.add(memory.q);

    memory.k2.scale_into(0.5, memory.aux);
    memory.aux.add(memory.temps);
    memory.k.prod_tri_diag_into(memory.aux, memory.k3);
    memory.k3.add(memory.q);

    memory.aux.copy_from(memory.k3);
    memory.aux.add(memory.temps);
    memory.k.prod_tri_diag_into(memory.aux, memory.k4);
    memory.k4.add(memory.q);

    memory.k1.div(6.);
    memory.k2.div(3.);
    memory.k3.div(3.);
    memory.k4.div(6.);

    memory.temps.add(memory.k1);
    memory.temps.add(memory.k2);
    memory.temps.add(memory.k3);
    memory.temps.add(memory.k4);
}

struct SurfaceMemory {  // :92-102
    std::vector<ChunkMemory> massive_chunks, nomass_chunks;
    Matrix temperatures, q;
};

// State slots registered by ThermalSurfaceData::new in this order (src/surface.rs:428-442)
enum { SLOT_FRONT_H = 0, SLOT_BACK_H, SLOT_FRONT_FLOW, SLOT_BACK_FLOW, SLOT_FRONT_SOLAR, SLOT_BACK_SOLAR,
       SLOT_FRONT_IR, SLOT_BACK_IR, SLOT_NODES };

struct ThermalSurface {  // :315-381
    size_t state_offset = 0;  // index of SLOT_FRONT_H in the state vector
    Discretization discretization;
    int front_kind = HO_BOUNDARY_OUTDOOR, back_kind = HO_BOUNDARY_OUTDOOR;
    int front_zone = -1, back_zone = -1;
    Float front_ambient_t = 0, back_ambient_t = 0;
    Float front_emissivity = 0, back_emissivity = 0;
    Float area = 0, perimeter = 0;
    Float normal[3] = {0, 0, 0};
    Float wind_speed_modifier_ = 0;
    Float cos_tilt = 0;
    std::vector<std::pair<size_t, size_t>> massive_chunks, nomass_chunks;
    Matrix front_alphas, back_alphas;
    Float front_hs = NAN, back_hs = NAN;  // Option<Float>, NaN = None
    bool is_fenestration = false;
    mutable int64_t nomass_iterations = 0;

    size_t n_nodes() const { return discretization.segments.size(); }
    Float front_temperature(const Float* state) const { return state[state_offset + SLOT_NODES]; }
    Float back_temperature(const Float* state) const { return state[state_offset + SLOT_NODES + n_nodes() - 1]; }

    SurfaceMemory allocate_memory() const {  // :385-410
        SurfaceMemory m;
        for (auto& c : massive_chunks) m.massive_chunks.emplace_back(c.first, c.second);
        for (auto& c : nomass_chunks) m.nomass_chunks.emplace_back(c.first, c.second);
        m.q = Matrix(0.0, n_nodes(), 1);
        m.temperatures = Matrix(0.0, n_nodes(), 1);
        return m;
    }

    // :596-717
    void calc_border_conditions(const Float* state, Float t_front, Float t_back, Float wind_direction,
                                Float wind_speed, ConvectionParams& front_env, ConvectionParams& back_env,
                                Float& front_hs_out, Float& back_hs_out) const {
        const Float ir_front = state[state_offset + SLOT_FRONT_IR];
        const Float ir_back = state[state_offset + SLOT_BACK_IR];
        const bool windward = is_windward(wind_direction, cos_tilt, normal);
        Float fh, bh;
        switch (front_kind) {
            case HO_BOUNDARY_SPACE:
                front_env = ConvectionParams{t_front, 0.0, t_front, front_temperature(state), 1, cos_tilt};
                fh = front_env.natural();
                break;
            case HO_BOUNDARY_AMBIENT:
                front_env = ConvectionParams{front_ambient_t, 0.0, t_front, front_temperature(state), 1, cos_tilt};
                fh = front_env.natural();
                break;
            case HO_BOUNDARY_OUTDOOR:
                front_env = ConvectionParams{t_front, wind_speed * wind_speed_modifier_,
                                             std::pow(ir_front / SIGMA, 0.25) - 273.15, front_temperature(state), 1,
                                             cos_tilt};
                front_env.cos_surface_tilt = -cos_tilt;
                fh = front_env.tarp(area, perimeter, windward);
                break;
            default: throw Panic("unreachable (Boundary::Ground)");
        }
        switch (back_kind) {
            case HO_BOUNDARY_SPACE:
                back_env = ConvectionParams{t_back, 0.0, t_back, back_temperature(state), 1, cos_tilt};
                bh = back_env.natural();
                break;
            case HO_BOUNDARY_AMBIENT:
                // quirk kept (:672-685): the back AmbientTemperature branch is built from *front* values
                back_env = ConvectionParams{back_ambient_t, 0.0, t_front, front_temperature(state), 1, cos_tilt};
                bh = back_env.natural();
                break;
            case HO_BOUNDARY_OUTDOOR:
                back_env = ConvectionParams{t_back, wind_speed * wind_speed_modifier_,
                                            std::pow(ir_back / SIGMA, 0.25) - 273.15, back_temperature(state), 1,
                                            cos_tilt};
                bh = back_env.tarp(area, perimeter, windward);
                break;
            default: throw Panic("unreachable (Boundary::Ground)");
        }
        if (std::isnan(fh) || std::isnan(bh)) throw Panic("Found NaN convection coefficients");
        front_hs_out = std::isnan(front_hs) ? fh : front_hs;  // debug_assertions override :708-714
        back_hs_out = std::isnan(back_hs) ? bh : back_hs;
    }

    // :720-787
    void march_mass(Matrix& global_temperatures, const Matrix& solar_radiation, Float dt, Float t_front, Float t_back,
                    Float front_rad_hs, Float back_rad_hs, Float wind_direction, Float wind_speed, size_t ini,
                    size_t fin, ChunkMemory& memory, const Float* state) const {
        ConvectionParams front_env, back_env;
        Float fhs, bhs;
        calc_border_conditions(state, t_front, t_back, wind_direction, wind_speed, front_env, back_env, fhs, bhs);
        discretization.get_k_q(ini, fin, global_temperatures, front_env, fhs, front_rad_hs, back_env, bhs,
                               back_rad_hs, memory);
        memory.c.scale(0.0);
        for (size_t i = 0; i < fin - ini; ++i) memory.c.set(i, i, discretization.segments[ini + i].first);
        for (size_t local_i = 0, global_i = ini; global_i < fin; ++local_i, ++global_i)
            memory.q.add_to_element(local_i, 0, solar_radiation.get(global_i, 0));
        rearrange_k(dt, memory);
        for (size_t local_i = 0, global_i = ini; global_i < fin; ++local_i, ++global_i)
            memory.temps.set(local_i, 0, global_temperatures.get(global_i, 0));
        rk4(memory);
        for (size_t local_i = 0, global_i = ini; global_i < fin; ++local_i, ++global_i)
            global_temperatures.set(global_i, 0, memory.temps.get(local_i, 0));
    }

    // :790-898
    void march_nomass(Matrix& global_temperatures, const Matrix& solar_radiation, Float t_front, Float t_back,
                      Float front_rad_hs, Float back_rad_hs, Float wind_direction, Float wind_speed, size_t ini,
                      size_t fin, ChunkMemory& memory, const Float* state) const {
        Float old_err = 99999.;
        size_t count = 0;
        while (true) {
            ++nomass_iterations;
            ConvectionParams front_env, back_env;
            Float fhs, bhs;
            calc_border_conditions(state, t_front, t_back, wind_direction, wind_speed, front_env, back_env, fhs, bhs);
            discretization.get_k_q(ini, fin, global_temperatures, front_env, fhs, front_rad_hs, back_env, bhs,
                                   back_rad_hs, memory);
            for (size_t local_i = 0, i = ini; i < fin; ++local_i, ++i)
                memory.q.add_to_element(local_i, 0, solar_radiation.get(i, 0));
            memory.q.scale(-1.);

            Matrix kclone = memory.k;
            const Matrix temps = kclone.mut_n_diag_gaussian(memory.q, 3);

            Float err = 0.0;
            for (size_t local_i = 0, i = ini; i < fin; ++local_i, ++i)
                err += std::fabs(temps.get(local_i, 0) - global_temperatures.get(i, 0));
            if (err > old_err) break;
            if (std::isnan(err)) throw Panic("assertion failed: !err.is_nan()");
            if (!(count < 99199000)) throw Panic("Excessive number of iterations");
            for (size_t local_i = 0, i = ini; i < fin; ++local_i, ++i) {
                global_temperatures.add_to_element(i, 0, temps.get(local_i, 0));
                global_temperatures.scale_element(i, 0, 0.5);
            }
            const Float max_allowed_error = count < 100 ? 0.01 : 0.5;
            if (err / (Float)(fin - ini) < max_allowed_error) break;
            old_err = err;
            count += 1;
        }
    }

    // :902-1024
    void march(const Float* state, Float t_front, Float t_back, Float wind_direction, Float wind_speed, Float dt,
               SurfaceMemory& memory) const {
        const size_t N = n_nodes();
        for (size_t i = 0; i < N; ++i) memory.temperatures.data[i] = state[state_offset + SLOT_NODES + i];

        Float solar_front = state[state_offset + SLOT_FRONT_SOLAR];
        if (std::isnan(solar_front) || solar_front < 0.0) solar_front = 0.0;
        Float solar_back = state[state_offset + SLOT_BACK_SOLAR];
        if (std::isnan(solar_back) || solar_front < 0.0) solar_back = 0.0;  // quirk kept (:921)

        // `&self.front_alphas * solar_front` allocates; `+= &(&self.back_alphas * solar_back)` allocates again
        Matrix solar_radiation(0.0, N, 1);
        front_alphas.scale_into(solar_front, solar_radiation);
        {
            Matrix tmp(0.0, N, 1);
            back_alphas.scale_into(solar_back, tmp);
            solar_radiation.add(tmp);
        }

        ConvectionParams front_env, back_env;
        Float fhs, bhs;
        calc_border_conditions(state, t_front, t_back, wind_direction, wind_speed, front_env, back_env, fhs, bhs);
        Float front_rad_hs = 4. * front_emissivity * SIGMA *
                             powi(273.15 + (front_env.rad_temperature + front_env.surface_temperature) / 2., 3);
        Float back_rad_hs = 4. * back_emissivity * SIGMA *
                            powi(273.15 + (back_env.rad_temperature + back_env.surface_temperature) / 2., 3);

        for (size_t ci = 0; ci < nomass_chunks.size(); ++ci)
            march_nomass(memory.temperatures, solar_radiation, t_front, t_back, front_rad_hs, back_rad_hs,
                         wind_direction, wind_speed, nomass_chunks[ci].first, nomass_chunks[ci].second,
                         memory.nomass_chunks[ci], state);

        calc_border_conditions(state, t_front, t_back, wind_direction, wind_speed, front_env, back_env, fhs, bhs);
        front_rad_hs = 4. * front_emissivity * SIGMA *
                       powi(273.15 + (front_env.rad_temperature + front_env.surface_temperature) / 2., 3);
        back_rad_hs = 4. * back_emissivity * SIGMA *
                      powi(273.15 + (back_env.rad_temperature + back_env.surface_temperature) / 2., 3);

        for (size_t ci = 0; ci < massive_chunks.size(); ++ci)
            march_mass(memory.temperatures, solar_radiation, dt, t_front, t_back, front_rad_hs, back_rad_hs,
                       wind_direction, wind_speed, massive_chunks[ci].first, massive_chunks[ci].second,
                       memory.massive_chunks[ci], state);
    }
};

}  // namespace

/* ------------------------------------------------------------------------------------------
 * src/model.rs + src/zone.rs
 * ---------------------------------------------------------------------------------------- */
struct ho_model {
    std::vector<Float> zone_volume;         // ThermalZone::volume (zone.rs:26-32); state slot z = zone z
    std::vector<ThermalSurface> surfaces;   // surfaces then fenestrations (model.rs:59-62)
    size_t n_plain_surfaces = 0;
    size_t dt_subdivisions = 1;
    Float dt = 0;
    size_t state_len = 0;
    std::vector<SurfaceMemory> alloc;       // ThermalModelMemory (model.rs:46-49)

    // zone.rs:59-65
    Float mcp(size_t z, Float temp) const {
        const Float air_density = AIR.density(temp + 273.15);
        const Float air_specific_heat = AIR.heat_capacity(temp + 273.15);
        return zone_volume[z] * air_density * air_specific_heat / 1.;
    }

    // model.rs:79-96
    Float boundary_temperature(int kind, int zone, Float ambient, Float t_out, const Float* state) const {
        switch (kind) {
            case HO_BOUNDARY_SPACE: return state[zone];
            case HO_BOUNDARY_AMBIENT: return ambient;
            case HO_BOUNDARY_OUTDOOR: return t_out;
            default: throw Panic("not implemented (Boundary::Ground)");
        }
    }

    // model.rs:102-180
    void iterate_surfaces(size_t begin, size_t end, Float wind_direction, Float wind_speed, Float t_out, Float dt_,
                          Float* state) {
        for (size_t si = begin; si < end; ++si) {
            ThermalSurface& s = surfaces[si];
            SurfaceMemory& memory = alloc[si];
            const Float t_front = boundary_temperature(s.front_kind, s.front_zone, s.front_ambient_t, t_out, state);
            const Float t_back = boundary_temperature(s.back_kind, s.back_zone, s.back_ambient_t, t_out, state);
            s.march(state, t_front, t_back, wind_direction, wind_speed, dt_, memory);
            const size_t rows = memory.temperatures.nrows;
            for (size_t i = 0; i < rows; ++i) state[s.state_offset + SLOT_NODES + i] = memory.temperatures.data[i];
            const Float ts_front = memory.temperatures.get(0, 0);
            const Float ts_back = memory.temperatures.get(rows - 1, 0);
            ConvectionParams fe, be;
            Float front_hs, back_hs;
            s.calc_border_conditions(state, t_front, t_back, wind_direction, wind_speed, fe, be, front_hs, back_hs);
            state[s.state_offset + SLOT_FRONT_H] = front_hs;
            state[s.state_offset + SLOT_BACK_H] = back_hs;
            const Float flow_front = (ts_front - t_front) * front_hs;
            const Float flow_back = (ts_back - t_back) * back_hs;
            state[s.state_offset + SLOT_FRONT_FLOW] = flow_front;
            state[s.state_offset + SLOT_BACK_FLOW] = flow_back;
        }
    }

    // model.rs:489-597
    void calculate_zones_abc(const Float* state, const ho_zone_inputs* zin, std::vector<Float>& a,
                             std::vector<Float>& b, std::vector<Float>& c) const {
        const size_t nzones = zone_volume.size();
        a.assign(nzones, 0.0);
        b.assign(nzones, 0.0);
        c.assign(nzones, 0.0);
        if (zin && zin->q_internal)
            for (size_t i = 0; i < nzones; ++i) a[i] += zin->q_internal[i];  // HVAC + luminaires :502-516
        for (size_t i = 0; i < nzones; ++i) {
            if (zin && zin->infiltration_temperature && !std::isnan(zin->infiltration_temperature[i])) {  // :524-533
                const Float t_inf = zin->infiltration_temperature[i];
                const Float v_inf = zin->infiltration_volume[i];
                const Float cp_inf = AIR.heat_capacity(t_inf + 273.15);
                const Float rho_inf = AIR.density(t_inf + 273.15);
                a[i] += rho_inf * v_inf * cp_inf * t_inf;
                b[i] += rho_inf * v_inf * cp_inf;
            }
            if (zin && zin->ventilation_temperature && !std::isnan(zin->ventilation_temperature[i])) {  // :536-544
                const Float t_v = zin->ventilation_temperature[i];
                const Float v_v = zin->ventilation_volume[i];
                const Float cp_v = AIR.heat_capacity(t_v + 273.15);
                const Float rho_v = AIR.density(t_v + 273.15);
                a[i] += rho_v * v_v * cp_v * t_v;
                b[i] += rho_v * v_v * cp_v;
            }
            c[i] = mcp(i, state[i]);  // :549-552
        }
        for (const ThermalSurface& s : surfaces) {  // surfaces, then fenestrations :556-590
            const Float h_front = state[s.state_offset + SLOT_FRONT_H];
            const Float h_back = state[s.state_offset + SLOT_BACK_H];
            const Float ai = s.area;
            if (s.front_kind == HO_BOUNDARY_SPACE) {
                const Float temp = s.front_temperature(state);
                a[s.front_zone] += h_front * ai * temp;
                b[s.front_zone] += h_front * ai;
            }
            if (s.back_kind == HO_BOUNDARY_SPACE) {
                const Float temp = s.back_temperature(state);
                a[s.back_zone] += h_back * ai * temp;
                b[s.back_zone] += h_back * ai;
            }
        }
    }

    // model.rs:359-427
    void march(Float* state, const Float* weather, const ho_zone_inputs* zin) {
        const size_t nzones = zone_volume.size();
        for (size_t sub = 0; sub < dt_subdivisions; ++sub) {
            const Float wind_direction = to_radians(weather[3 * sub + 1]);
            const Float wind_speed = weather[3 * sub + 2];
            const Float t_out = weather[3 * sub + 0];
            std::vector<Float> t_current(state, state + nzones);  // :385, :601-610
            iterate_surfaces(0, n_plain_surfaces, wind_direction, wind_speed, t_out, dt, state);
            iterate_surfaces(n_plain_surfaces, surfaces.size(), wind_direction, wind_speed, t_out, dt, state);
            std::vector<Float> a, b, c;
            calculate_zones_abc(state, zin, a, b, c);
            // :650-674
            std::vector<Float> future;
            future.reserve(nzones);
            for (size_t i = 0; i < nzones; ++i) {
                if (std::fabs(b[i]) > 1e-9) {
                    future.push_back(a[i] / b[i] + (t_current[i] - a[i] / b[i]) * std::exp(-b[i] * dt / c[i]));
                } else {
                    future.push_back(t_current[i]);
                }
            }
            for (size_t i = 0; i < nzones; ++i) {
                if (std::isnan(future[i])) throw Panic("Future temperatures is NaN");
                state[i] = future[i];
            }
        }
    }
};

namespace {

// ThermalSurfaceData::new (src/surface.rs:414-567) + the per-surface part of ThermalModel::new (model.rs:245-282)
ThermalSurface make_surface(const ho_model_desc* d, const ho_surface& hs, Float main_dt, Float max_dx, Float min_dt,
                            size_t& state_len) {
    if (hs.construction < 0 || hs.construction >= d->n_constructions) throw Panic("Construction not found");
    const int l0 = d->construction_offset[hs.construction], l1 = d->construction_offset[hs.construction + 1];
    const ho_layer* layers = d->layers + l0;
    const size_t n_layers = (size_t)(l1 - l0);
    if (n_layers == 0) throw Panic("empty construction");

    ThermalSurface ts;
    const Float cos_tilt = hs.normal[0] * 0. + hs.normal[1] * 0. + hs.normal[2] * 1.;  // model.rs:249
    const Float height = 1.;                                                            // model.rs:252
    const Float angle = std::acos(cos_tilt);                                            // model.rs:253
    ts.discretization = Discretization::make(layers, n_layers, main_dt, max_dx, min_dt, height, angle);

    ts.state_offset = state_len;  // surface.rs:428-442
    state_len += SLOT_NODES + ts.discretization.segments.size();

    const ho_layer& front_sub = layers[0];
    const ho_layer& back_sub = layers[n_layers - 1];
    if (front_sub.kind != HO_LAYER_SOLID || back_sub.kind != HO_LAYER_SOLID)
        throw Panic("Front Emissivity not available for this particular kind of Substance");
    ts.front_emissivity = or_default(front_sub.front_emissivity, DEFAULT_EM);  // :450-455
    ts.back_emissivity = or_default(back_sub.back_emissivity, DEFAULT_EM);     // :456-461
    ts.discretization.get_chunks(ts.massive_chunks, ts.nomass_chunks);         // :463

    // :466-537 solar absorption per node
    const size_t n_nodes = ts.discretization.segments.size();
    std::vector<int> fwd(n_layers), rev(n_layers);
    for (size_t i = 0; i < n_layers; ++i) {
        fwd[i] = (int)i;
        rev[i] = (int)(n_layers - 1 - i);
    }
    const std::vector<Float> front_alphas_prev = Glazing::alphas(glazing_from_iter(layers, fwd));
    // DEVIATION (documented in DESIGN.md): the reference panics unless len ∈ {1, n_layers} (:470-472), which makes
    // every glass–gas–… stack unusable; we also accept len == (n_layers+1)/2, the case its indexing
    // `layer_index = 2*alpha_i` (:479) is written for.
    auto len_ok = [&](size_t len) { return len == 1 || len == n_layers || len == (n_layers + 1) / 2; };
    if (!len_ok(front_alphas_prev.size()))
        throw Panic("Construction seems to have a mixture of transparent and opaque layers");
    ts.front_alphas = Matrix(0.0, n_nodes, 1);
    {
        size_t global_i = 0;
        for (size_t alpha_i = 0; alpha_i < front_alphas_prev.size(); ++alpha_i) {
            const size_t layer_index = 2 * alpha_i;
            const size_t n = ts.discretization.n_elements[layer_index] == 0 ? 1 : ts.discretization.n_elements[layer_index];
            const ho_layer& sub = layers[layer_index];
            if (sub.kind != HO_LAYER_SOLID) throw Panic("unreachable (alpha on gas)");
            const Float tau = or_default(sub.solar_transmittance, 0.0);
            const Float alpha = front_alphas_prev[alpha_i];
            if (tau > 0.0) {
                for (size_t local_i = 0; local_i <= n; ++local_i)
                    ts.front_alphas.add_to_element(global_i + local_i, 0, alpha / (Float)(n + 1));
            } else {
                ts.front_alphas.add_to_element(global_i, 0, alpha);
            }
            global_i += n + 1;
        }
    }
    const std::vector<Float> back_alphas_prev = Glazing::alphas(glazing_from_iter(layers, rev));
    if (!len_ok(back_alphas_prev.size()))
        throw Panic("Construction seems to have a mixture of transparent and opaque layers");
    ts.back_alphas = Matrix(0.0, n_nodes, 1);
    {
        size_t global_i = n_nodes;
        for (size_t alpha_i = 0; alpha_i < back_alphas_prev.size(); ++alpha_i) {
            const size_t layer_index = n_layers - 2 * alpha_i - 1;
            const size_t n = ts.discretization.n_elements[layer_index] == 0 ? 1 : ts.discretization.n_elements[layer_index];
            const ho_layer& sub = layers[layer_index];
            if (sub.kind != HO_LAYER_SOLID) throw Panic("unreachable (alpha on gas)");
            const Float tau = or_default(sub.solar_transmittance, 0.0);
            const Float alpha = back_alphas_prev[alpha_i];
            if (tau > 0.0) {
                for (size_t local_i = 0; local_i <= n; ++local_i) {
                    const size_t row = global_i - local_i - 1;
                    ts.back_alphas.add_to_element(row, 0, alpha / (Float)(n + 1));
                }
            } else {
                ts.back_alphas.add_to_element(global_i - 1, 0, alpha);
            }
            global_i -= n + 1;
        }
    }

    ts.cos_tilt = cos_tilt;  // :539
    ts.wind_speed_modifier_ = wind_speed_modifier(hs.centroid_z, d->terrain);  // :540
    ts.area = hs.area;
    ts.perimeter = hs.perimeter;
    for (int i = 0; i < 3; ++i) ts.normal[i] = hs.normal[i];
    ts.front_kind = hs.front_kind;
    ts.back_kind = hs.back_kind;
    ts.front_zone = hs.front_zone;
    ts.back_zone = hs.back_zone;
    ts.front_ambient_t = hs.front_ambient_t;
    ts.back_ambient_t = hs.back_ambient_t;
    ts.front_hs = hs.front_hs_override;
    ts.back_hs = hs.back_hs_override;
    if (!std::isnan(hs.front_emissivity_override)) ts.front_emissivity = hs.front_emissivity_override;
    if (!std::isnan(hs.back_emissivity_override)) ts.back_emissivity = hs.back_emissivity_override;
    ts.is_fenestration = hs.is_fenestration != 0;
    for (int kind_zone = 0; kind_zone < 2; ++kind_zone) {
        const int kind = kind_zone ? ts.back_kind : ts.front_kind;
        const int zone = kind_zone ? ts.back_zone : ts.front_zone;
        if (kind == HO_BOUNDARY_SPACE && (zone < 0 || zone >= d->n_zones)) throw Panic("Space not found");
    }
    return ts;
}

template <class F>
auto guarded(F&& f, decltype(f()) on_error) -> decltype(f()) {
    try {
        return f();
    } catch (const std::exception& e) {
        g_error = e.what();
        return on_error;
    }
}

}  // namespace

extern "C" {

const char* ho_last_error(void) { return g_error.c_str(); }

// ThermalModel::new — src/model.rs:215-354
ho_model* ho_model_new(const ho_model_desc* d) {
    return ho_model_new_ex(d, 0., 0.04 /* model.rs:236 */, 60. /* model.rs:237 */);
}

// Same, with the discretization parameters the reference's unit tests pass to Discretization::new directly
// (src/surface.rs:1124-1127,1264-1267,1377-1380,1485-1488).  main_dt_override <= 0: main_dt = 3600/n (model.rs:240).
ho_model* ho_model_new_ex(const ho_model_desc* d, double main_dt_override, double max_dx, double min_dt) {
    return guarded(
        [&]() -> ho_model* {
            auto m = std::make_unique<ho_model>();
            m->zone_volume.assign(d->zone_volume, d->zone_volume + d->n_zones);
            size_t state_len = (size_t)d->n_zones;  // one SpaceDryBulbTemperature per space, zone.rs:43-50
            size_t dt_subdivisions = 1;
            const size_t n = (size_t)d->steps_per_hour;
            const Float main_dt = main_dt_override > 0. ? main_dt_override : 60. * 60. / (Float)n;  // :240
            // surfaces first, then fenestrations (:244-323); the descriptor must already be in that order
            bool seen_fen = false;
            for (int i = 0; i < d->n_surfaces; ++i) {
                const ho_surface& hs = d->surfaces[i];
                if (hs.is_fenestration) seen_fen = true;
                else if (seen_fen) throw Panic("descriptor must list surfaces before fenestrations");
                ThermalSurface ts = make_surface(d, hs, main_dt, max_dx, min_dt, state_len);
                if (ts.discretization.tstep_subdivision > dt_subdivisions)
                    dt_subdivisions = ts.discretization.tstep_subdivision;
                if (!hs.is_fenestration) m->n_plain_surfaces += 1;
                m->surfaces.push_back(std::move(ts));
            }
            Float dt = 60. * 60. / ((Float)n * (Float)dt_subdivisions);  // :326
            const size_t SAFETY = 2;                                      // :329
            dt /= (Float)SAFETY;
            dt_subdivisions *= SAFETY;
            m->dt = dt;
            m->dt_subdivisions = dt_subdivisions;
            m->state_len = state_len;
            for (auto& s : m->surfaces) m->alloc.push_back(s.allocate_memory());  // allocate_memory :193-207
            return m.release();
        },
        nullptr);
}

void ho_model_free(ho_model* m) { delete m; }

int32_t ho_dt_subdivisions(const ho_model* m) { return (int32_t)m->dt_subdivisions; }
double ho_dt(const ho_model* m) { return m->dt; }
int32_t ho_state_len(const ho_model* m) { return (int32_t)m->state_len; }
int32_t ho_surface_n_nodes(const ho_model* m, int32_t s) { return (int32_t)m->surfaces[s].n_nodes(); }
int32_t ho_surface_state_offset(const ho_model* m, int32_t s) { return (int32_t)m->surfaces[s].state_offset; }
int32_t ho_surface_tstep_subdivision(const ho_model* m, int32_t s) {
    return (int32_t)m->surfaces[s].discretization.tstep_subdivision;
}
int32_t ho_surface_n_elements(const ho_model* m, int32_t s, int32_t* out, int32_t cap) {
    const auto& v = m->surfaces[s].discretization.n_elements;
    for (size_t i = 0; i < v.size() && (int32_t)i < cap; ++i) out[i] = (int32_t)v[i];
    return (int32_t)v.size();
}
int32_t ho_surface_segments(const ho_model* m, int32_t s, double* mass, double* u) {
    const auto& seg = m->surfaces[s].discretization.segments;
    for (size_t i = 0; i < seg.size(); ++i) {
        mass[i] = seg[i].first;
        u[i] = seg[i].second.kind == UValue::Solid ? seg[i].second.u : (seg[i].second.kind == UValue::CavityK ? -1. : 0.);
    }
    return (int32_t)seg.size();
}
int32_t ho_surface_chunks(const ho_model* m, int32_t s, int32_t massive, int32_t* out, int32_t cap) {
    const auto& v = massive ? m->surfaces[s].massive_chunks : m->surfaces[s].nomass_chunks;
    for (size_t i = 0; i < v.size() && (int32_t)(2 * i + 1) < cap; ++i) {
        out[2 * i] = (int32_t)v[i].first;
        out[2 * i + 1] = (int32_t)v[i].second;
    }
    return (int32_t)v.size();
}
int32_t ho_surface_alphas(const ho_model* m, int32_t s, double* front, double* back) {
    const auto& S = m->surfaces[s];
    for (size_t i = 0; i < S.n_nodes(); ++i) {
        front[i] = S.front_alphas.data[i];
        back[i] = S.back_alphas.data[i];
    }
    return (int32_t)S.n_nodes();
}
double ho_surface_wind_modifier(const ho_model* m, int32_t s) { return m->surfaces[s].wind_speed_modifier_; }
double ho_surface_emissivity(const ho_model* m, int32_t s, int32_t back) {
    return back ? m->surfaces[s].back_emissivity : m->surfaces[s].front_emissivity;
}

void ho_initial_state(const ho_model* m, double* state) {
    for (size_t z = 0; z < m->zone_volume.size(); ++z) state[z] = 22.0;  // zone.rs:45-50
    for (const auto& s : m->surfaces) {
        double* p = state + s.state_offset;
        p[SLOT_FRONT_H] = 1.739658084820765;  // surface_trait.rs:231
        p[SLOT_BACK_H] = 1.739658084820765;   // surface_trait.rs:248
        p[SLOT_FRONT_FLOW] = p[SLOT_BACK_FLOW] = 0.0;
        p[SLOT_FRONT_SOLAR] = p[SLOT_BACK_SOLAR] = 0.0;
        p[SLOT_FRONT_IR] = p[SLOT_BACK_IR] = 0.0;
        for (size_t i = 0; i < s.n_nodes(); ++i) p[SLOT_NODES + i] = 22.0;  // surface_trait.rs:365-370
    }
}

int32_t ho_march(const ho_model* m, double* state, const double* weather, const ho_zone_inputs* zin) {
    return guarded(
        [&]() -> int32_t {
            const_cast<ho_model*>(m)->march(state, weather, zin);
            return 0;
        },
        1);
}

int32_t ho_iterate_surfaces(const ho_model* m, double* state, double wind_direction_rad, double wind_speed,
                            double t_out, double dt) {
    return guarded(
        [&]() -> int32_t {
            ho_model* mm = const_cast<ho_model*>(m);
            mm->iterate_surfaces(0, mm->n_plain_surfaces, wind_direction_rad, wind_speed, t_out, dt, state);
            mm->iterate_surfaces(mm->n_plain_surfaces, mm->surfaces.size(), wind_direction_rad, wind_speed, t_out, dt,
                                 state);
            return 0;
        },
        1);
}

int32_t ho_zones_abc(const ho_model* m, const double* state, const ho_zone_inputs* zin, double* a, double* b,
                     double* c) {
    return guarded(
        [&]() -> int32_t {
            std::vector<Float> va, vb, vc;
            m->calculate_zones_abc(state, zin, va, vb, vc);
            std::copy(va.begin(), va.end(), a);
            std::copy(vb.begin(), vb.end(), b);
            std::copy(vc.begin(), vc.end(), c);
            return 0;
        },
        1);
}

int64_t ho_nomass_iterations(const ho_model* m) {
    int64_t n = 0;
    for (auto& s : m->surfaces) n += s.nomass_iterations;
    return n;
}

double ho_nusselt(double ra, double gamma, double a_gi) {
    return guarded([&]() -> double { return nusselt(ra, gamma, a_gi); }, NAN);
}
double ho_gas_property(int32_t gas, int32_t which, double t) {
    return guarded(
        [&]() -> double {
            const Gas& g = gas_by_id(gas);
            switch (which) {
                case 0: return g.thermal_conductivity(t);
                case 1: return g.dynamic_viscosity(t);
                case 2: return g.heat_capacity(t);
                case 3: return g.mass;
                default: return g.density(t);
            }
        },
        NAN);
}
double ho_cavity_u_value(int32_t gas, double thickness, double height, double angle, double eout, double ein,
                         double t_front, double t_back) {
    return guarded(
        [&]() -> double {
            Cavity c{thickness, height, &gas_by_id(gas), eout, ein, angle};
            return c.u_value(t_front, t_back);
        },
        NAN);
}
double ho_tarp_natural(double t_air, double t_surf, double cos_tilt) {
    return guarded([&]() -> double { return ConvectionParams{t_air, 0., t_air, t_surf, 1, cos_tilt}.natural(); }, NAN);
}
double ho_tarp_forced_plus_natural(double t_air, double t_surf, double cos_tilt, double air_speed, double area,
                                   double perimeter, int32_t windward) {
    return guarded(
        [&]() -> double {
            return ConvectionParams{t_air, air_speed, t_air, t_surf, 1, cos_tilt}.tarp(area, perimeter, windward != 0);
        },
        NAN);
}
int32_t ho_is_windward(double wd, double cos_tilt, const double* normal) { return is_windward(wd, cos_tilt, normal); }
double ho_wind_speed_modifier(double height, int32_t terrain) { return wind_speed_modifier(height, terrain); }

void ho_rk4_steps(int32_t n, double* k, const double* c, double* q, double* temps, double dt, int32_t nsteps,
                  int32_t rearrange) {
    ChunkMemory mem(0, (size_t)n);
    mem.k.data.assign(k, k + (size_t)n * n);
    for (int i = 0; i < n; ++i) mem.c.set(i, i, c[i]);
    mem.q.data.assign(q, q + n);
    mem.temps.data.assign(temps, temps + n);
    if (rearrange) rearrange_k(dt, mem);
    for (int s = 0; s < nsteps; ++s) rk4(mem);
    std::copy(mem.temps.data.begin(), mem.temps.data.end(), temps);
    std::copy(mem.k.data.begin(), mem.k.data.end(), k);
    std::copy(mem.q.data.begin(), mem.q.data.end(), q);
}

int32_t ho_discretize(const ho_layer* layers, int32_t n_layers, double model_dt, double max_dx, double min_dt,
                      double height, double angle, int32_t* tstep_subdivision, int32_t* n_elements, double* mass,
                      double* u, int32_t cap_nodes) {
    return guarded(
        [&]() -> int32_t {
            Discretization d = Discretization::make(layers, (size_t)n_layers, model_dt, max_dx, min_dt, height, angle);
            *tstep_subdivision = (int32_t)d.tstep_subdivision;
            for (size_t i = 0; i < d.n_elements.size(); ++i) n_elements[i] = (int32_t)d.n_elements[i];
            if ((int32_t)d.segments.size() > cap_nodes) throw Panic("cap_nodes too small");
            for (size_t i = 0; i < d.segments.size(); ++i) {
                mass[i] = d.segments[i].first;
                const auto& uv = d.segments[i].second;
                u[i] = uv.kind == UValue::Solid ? uv.u : (uv.kind == UValue::CavityK ? -1. : 0.);
            }
            return (int32_t)d.segments.size();
        },
        -1);
}


/* Discretization::build on given element counts (src/discretization.rs:163-298) */
int32_t ho_build(const ho_layer* layers, int32_t n_layers, int32_t tstep_subdivision, const int32_t* n_elements,
                 double height, double angle, double* mass, double* u, int32_t cap_nodes) {
    return guarded(
        [&]() -> int32_t {
            std::vector<size_t> el(n_elements, n_elements + n_layers);
            Discretization d = Discretization::build(layers, (size_t)n_layers, (size_t)tstep_subdivision, el, height, angle);
            if ((int32_t)d.segments.size() > cap_nodes) throw Panic("cap_nodes too small");
            for (size_t i = 0; i < d.segments.size(); ++i) {
                mass[i] = d.segments[i].first;
                const auto& uv = d.segments[i].second;
                u[i] = uv.kind == UValue::Solid ? uv.u : (uv.kind == UValue::CavityK ? -1. : 0.);
            }
            return (int32_t)d.segments.size();
        },
        -1);
}

/* Discretization::get_chunks on bare node masses (src/discretization.rs:117-160); returns the chunk count */
int32_t ho_chunks(const double* mass, int32_t n, int32_t massive, int32_t* out, int32_t cap) {
    return guarded(
        [&]() -> int32_t {
            Discretization d;
            d.segments.resize((size_t)n);
            for (int i = 0; i < n; ++i) d.segments[i].first = mass[i];
            std::vector<std::pair<size_t, size_t>> ma, nm;
            d.get_chunks(ma, nm);
            const auto& v = massive ? ma : nm;
            if ((int32_t)(2 * v.size()) > cap) throw Panic("cap too small");
            for (size_t i = 0; i < v.size(); ++i) {
                out[2 * i] = (int32_t)v[i].first;
                out[2 * i + 1] = (int32_t)v[i].second;
            }
            return (int32_t)v.size();
        },
        -1);
}

/* Discretization::get_k_q on solid segments (src/discretization.rs:596-700): segments = (mass[i], Solid(u[i])),
 * last one Back.  k_out: dense (fin-ini)x(fin-ini) row-major, q_out: fin-ini. */
int32_t ho_get_k_q(int32_t n, const double* mass, const double* u, int32_t ini, int32_t fin, const double* temps,
                   double front_air, double front_rad_t, double front_hs, double front_rad_hs, double back_air,
                   double back_rad_t, double back_hs, double back_rad_hs, double* k_out, double* q_out) {
    return guarded(
        [&]() -> int32_t {
            Discretization d;
            d.segments.resize((size_t)n);
            for (int i = 0; i < n; ++i) {
                d.segments[i].first = mass[i];
                d.segments[i].second.kind = i + 1 < n ? UValue::Solid : UValue::Back;
                d.segments[i].second.u = u[i];
            }
            Matrix t = Matrix::from_data((size_t)n, 1, std::vector<Float>(temps, temps + n));
            ConvectionParams fe{}, be{};
            fe.air_temperature = front_air; fe.rad_temperature = front_rad_t; fe.surface_temperature = 22.;
            fe.roughness_index = 1;
            be.air_temperature = back_air; be.rad_temperature = back_rad_t; be.surface_temperature = 22.;
            be.roughness_index = 1;
            ChunkMemory mem((size_t)ini, (size_t)fin);
            d.get_k_q((size_t)ini, (size_t)fin, t, fe, front_hs, front_rad_hs, be, back_hs, back_rad_hs, mem);
            const size_t m = (size_t)(fin - ini);
            for (size_t r = 0; r < m; ++r) {
                q_out[r] = mem.q.get(r, 0);
                for (size_t c = 0; c < m; ++c) k_out[r * m + c] = mem.k.get(r, c);
            }
            return 0;
        },
        1);
}

/* ISO 9050 pair combination (src/glazing.rs:180-218): out = tau, rho_front, rho_back, alpha_front, alpha_back of
 * the n panes combined front to back */
int32_t ho_glazing_combine(const double* tau, const double* rho_front, const double* rho_back, int32_t n, double* out) {
    return guarded(
        [&]() -> int32_t {
            std::vector<Glazing> g;
            for (int i = 0; i < n; ++i) g.push_back(Glazing::make(tau[i], rho_front[i], rho_back[i]));
            Glazing c = Glazing::combine_layers(g.data(), g.size());
            out[0] = c.tau; out[1] = c.rho_front; out[2] = c.rho_back; out[3] = c.alpha_front; out[4] = c.alpha_back;
            return 0;
        },
        1);
}

int32_t ho_glazing_alphas(const double* tau, const double* rho_front, const double* rho_back, int32_t n,
                          double* out) {
    return guarded(
        [&]() -> int32_t {
            std::vector<Glazing> g;
            for (int i = 0; i < n; ++i) g.push_back(Glazing::make(tau[i], rho_front[i], rho_back[i]));
            auto a = Glazing::alphas(g);
            std::copy(a.begin(), a.end(), out);
            return (int32_t)a.size();
        },
        -1);
}

int64_t ho_run_ensemble_q(ho_model* const* models, int32_t n_models, int32_t n_steps, const double* weather,
                          const double* surf_inputs, int32_t weather_linear, int32_t threads,
                          double* final_state_concat, const double* q_internal_concat);

int64_t ho_run_ensemble(ho_model* const* models, int32_t n_models, int32_t n_steps, const double* weather,
                        const double* surf_inputs, int32_t weather_linear, int32_t threads,
                        double* final_state_concat) {
    return ho_run_ensemble_q(models, n_models, n_steps, weather, surf_inputs, weather_linear, threads,
                             final_state_concat, nullptr);
}

// Same, with the zones' HVAC + luminaire power (model.rs:502-516) given per zone, models concatenated (or NULL).
int64_t ho_run_ensemble_q(ho_model* const* models, int32_t n_models, int32_t n_steps, const double* weather,
                          const double* surf_inputs, int32_t weather_linear, int32_t threads,
                          double* final_state_concat, const double* q_internal_concat) {
    std::vector<size_t> offsets(n_models + 1, 0), zoffsets(n_models + 1, 0);
    for (int i = 0; i < n_models; ++i) offsets[i + 1] = offsets[i] + models[i]->state_len;
    for (int i = 0; i < n_models; ++i) zoffsets[i + 1] = zoffsets[i] + models[i]->zone_volume.size();
    std::atomic<int> next{0};
    std::atomic<int64_t> updates{0};
    std::atomic<int> failed{0};
    auto worker = [&]() {
        int64_t local = 0;
        std::vector<Float> state, wrow;
        while (true) {
            const int i = next.fetch_add(1);
            if (i >= n_models) break;
            ho_model* m = models[i];
            state.assign(m->state_len, 0.0);
            ho_initial_state(m, state.data());
            const size_t S = m->dt_subdivisions;
            wrow.assign(3 * S, 0.0);
            int64_t nodes = 0;
            for (auto& s : m->surfaces) nodes += (int64_t)s.n_nodes();
            try {
                for (int step = 0; step < n_steps; ++step) {
                    const double* w0 = weather + 3 * (size_t)step;
                    const double* w1 = weather + 3 * (size_t)(step + 1);  // needs n_steps+1 rows when linear
                    for (size_t s = 0; s < S; ++s) {
                        if (weather_linear) {
                            const Float f = (Float)(s + 1) / (Float)S;
                            for (int k = 0; k < 3; ++k) wrow[3 * s + k] = w0[k] + (w1[k] - w0[k]) * f;
                        } else {
                            for (int k = 0; k < 3; ++k) wrow[3 * s + k] = w0[k];
                        }
                    }
                    if (surf_inputs) {
                        const double* in = surf_inputs + 4 * (size_t)step;
                        for (auto& s : m->surfaces) {
                            state[s.state_offset + SLOT_FRONT_SOLAR] = in[0];
                            state[s.state_offset + SLOT_BACK_SOLAR] = in[1];
                            state[s.state_offset + SLOT_FRONT_IR] = in[2];
                            state[s.state_offset + SLOT_BACK_IR] = in[3];
                        }
                    }
                    ho_zone_inputs zin{};
                    if (q_internal_concat) zin.q_internal = q_internal_concat + zoffsets[i];
                    m->march(state.data(), wrow.data(), q_internal_concat ? &zin : nullptr);
                    local += nodes * (int64_t)S;
                }
            } catch (const std::exception&) {
                failed.fetch_add(1);
            }
            if (final_state_concat) std::copy(state.begin(), state.end(), final_state_concat + offsets[i]);
        }
        updates.fetch_add(local);
    };
    std::vector<std::thread> pool;
    const int nt = std::max(1, threads);
    for (int t = 1; t < nt; ++t) pool.emplace_back(worker);
    worker();
    for (auto& t : pool) t.join();
    if (failed.load()) {
        g_error = "ho_run_ensemble: a model panicked";
        return -1;
    }
    return updates.load();
}

}  // extern "C"

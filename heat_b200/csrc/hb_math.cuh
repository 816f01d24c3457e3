// hb_math.cuh — scalar physics of the wall march, callable from the CUDA kernels (and, for the test emulator in
// tests/emu, from host code: every function is HB_HD).  All f64.
//
//   tarp_natural      convection.rs:87-110
//   cavity_u/nusselt  cavity.rs:59-69, gas.rs:45-74,82-315
//   border            surface.rs:596-717 (calc_border_conditions)
//   solve_nomass      surface.rs:790-898 around discretization.rs:596-700 (get_k_q) for one no-mass chunk
//   zone_update       model.rs:489-597 (calculate_zones_abc), zone.rs:59-65 (mcp), model.rs:650-674
//
// pow(x,1/3) is evaluated as cbrt(x) and pow(x,1/4) as sqrt(sqrt(x)) (≤ 1 ulp apart from libm pow).
#pragma once
#include <cmath>
#include <cstdint>

#include "hb_dev.hpp"

#if defined(__CUDACC__)
#define HB_HD __host__ __device__ __forceinline__
#define HB_HD_NOINLINE static __host__ __device__ __noinline__
#else
#define HB_HD inline
#define HB_HD_NOINLINE static inline
#endif

namespace hb {

constexpr double kPi = 3.14159265358979323846;
// passes the no-mass fixed point may take before the reference gives up ("Excessive number of iterations",
// surface.rs:869-873).  A solve that gets anywhere near it is a lane spinning for seconds — as the reference's would be:
// some all-no-mass shoeboxes need > 10^5 passes in a sub-step and the reference carries on (tests: check_slow_nomass_…).
constexpr int kNoMassIterCap = 99199000;

HB_HD double hb_nan() {
#if defined(__CUDA_ARCH__)
    return __longlong_as_double(0x7ff8000000000000LL);
#else
    return NAN;
#endif
}

// Per-wall constants (shared memory in the kernel), derived once per launch by load_wallc().
enum : int32_t {
    WF_F_OUT = 1, WF_F_SPACE = 2, WF_B_OUT = 4, WF_B_SPACE = 8, WF_B_AMB = 16, WF_TILTED = 32, WF_ANY_OUT = 64,
    WF_DIST = 128, WF_OV_F = 256, WF_OV_B = 512,
};
struct WallC {
    double area, cos_tilt, forced_c, eps_f, eps4s_f, eps4s_b, amb_f, amb_b, ov_f, ov_b, nx, ny, af0, abL, v0, vL;
    // TARP natural-convection factor by sign of (T_air − T_surface), per face (convection.rs:91-100 with the
    // face's signed cos(tilt) folded in; 1.31 for a vertical face)
    double kpos_f, kneg_f, kpos_b, kneg_b;
    int32_t n, fl, zf, zb, site, nm_begin, nm_end, cav_begin, cav_end, desc_index;
};
static_assert(sizeof(WallC) % 8 == 0, "WallC is copied as doubles");

HB_HD void load_wallc(const DevView& V, int64_t sg, bool single_site, WallC& c) {
    c.area = V.area[sg];
    c.cos_tilt = V.cos_tilt[sg];
    c.forced_c = V.forced_c[sg];
    c.eps_f = V.eps_f[sg];
    c.eps4s_f = 4. * V.eps_f[sg] * kSigma;   // surface.rs:941-948: 4·ε·σ·T³ evaluated left to right
    c.eps4s_b = 4. * V.eps_b[sg] * kSigma;
    c.amb_f = V.amb_f[sg]; c.amb_b = V.amb_b[sg];
    c.ov_f = V.hs_ov_f[sg]; c.ov_b = V.hs_ov_b[sg];
    c.nx = V.nx[sg]; c.ny = V.ny[sg];
    c.af0 = V.af0[sg]; c.abL = V.abL[sg]; c.v0 = V.v0[sg]; c.vL = V.vL[sg];
    c.n = V.n_nodes[sg];
    const int flags = V.flags[sg];
    const int kind_f = flags & 15, kind_b = (flags >> 4) & 15;
    int fl = 0;
    if (kind_f == HB_BOUNDARY_OUTDOOR) fl |= WF_F_OUT | WF_ANY_OUT;
    if (kind_f == HB_BOUNDARY_SPACE) fl |= WF_F_SPACE;
    if (kind_b == HB_BOUNDARY_OUTDOOR) fl |= WF_B_OUT | WF_ANY_OUT;
    if (kind_b == HB_BOUNDARY_SPACE) fl |= WF_B_SPACE;
    if (kind_b == HB_BOUNDARY_AMBIENT) fl |= WF_B_AMB;
    const double act = fabs(c.cos_tilt);
    if (act < 0.98) fl |= WF_TILTED;       // surface.rs:38
    if ((flags >> 8) & 1) fl |= WF_DIST;
    if (c.ov_f == c.ov_f) fl |= WF_OV_F;
    if (c.ov_b == c.ov_b) fl |= WF_OV_B;
    c.fl = fl;
    const double c_same = 9.482 / (7.238 - act), c_opp = 1.81 / (1.382 + act);
    const double tilt_f = (fl & WF_F_OUT) ? -c.cos_tilt : c.cos_tilt;   // front Outdoor flips the tilt, surface.rs:652
    const double tilt_b = c.cos_tilt;
    const bool flat = act < 1e-3;
    c.kpos_f = flat ? 1.31 : tilt_f > 0. ? c_same : c_opp;
    c.kneg_f = flat ? 1.31 : tilt_f < 0. ? c_same : c_opp;
    c.kpos_b = flat ? 1.31 : tilt_b > 0. ? c_same : c_opp;
    c.kneg_b = flat ? 1.31 : tilt_b < 0. ? c_same : c_opp;
    c.zf = V.zone_f[sg]; c.zb = V.zone_b[sg];
    c.site = single_site ? 0 : V.site[sg];
    c.nm_begin = V.nm_off[sg]; c.nm_end = V.nm_off[sg + 1];
    c.cav_begin = V.cav_off[sg]; c.cav_end = V.cav_off[sg + 1];
    c.desc_index = V.desc_index[sg];
}

// Per-wall dynamic scalar state (registers of the scalar lane).
struct WallS {
    double h_f, h_b, flow_f, flow_b, sol_f, sol_b, ir_f, ir_b;   // the 8 SimulationState slots (surface.rs:428-442)
    double e_f, e_b;                 // ∫ flow·A dt
    double ts_f, ts_b;               // temperatures of the two end nodes
    double solar_front, solar_back;  // clamped irradiances of this main step (surface.rs:916-923)
    double trad_out_f, trad_out_b;   // outdoor radiant temperatures of this main step (surface.rs:647,692)
    double t_front, t_back, forced;  // carried from the pre- to the post-march phase of a sub-step
    unsigned long long nm_iters;
    int32_t status;
};

// What the scalar lane hands to the lanes that march the nodes.
struct BoundaryLinks {
    double ghost_f, bf;   // front ghost temperature and its link coefficient h_f·dt/C_0
    double ghost_b, ab;   // back ghost temperature and h_b·dt/C_{n-1}
};

struct SubWeather {
    double t_out, sqrt_ws, sin_wd, cos_wd;
};

// Products and sums that feed a DISCONTINUOUS decision (the windward test) are rounded separately, as the reference
// does (Rust never contracts a*b+c): with an FMA, 300° + (0° − 300°)·0.7 is 90.00000000000001° instead of 90°, which
// flips cos(wind direction) from +6e-17 to −1e-14 and with it the forced-convection factor of a south wall.
HB_HD double mul_rn(double a, double b) {
#if defined(__CUDA_ARCH__)
    return __dmul_rn(a, b);
#else
    return a * b;
#endif
}
HB_HD double add_rn(double a, double b) {
#if defined(__CUDA_ARCH__)
    return __dadd_rn(a, b);
#else
    return a + b;
#endif
}
// row0 + (row1 − row0)·f with every operation rounded (model.rs:371-382 samples the weather at the sub-step's end)
HB_HD double lerp_rn(double a, double b, double f) { return add_rn(a, mul_rn(add_rn(b, -a), f)); }

// ---- short-latency device arithmetic for the scalar lane (it is the block's critical path) -------------------
// 1/x: MUFU.RCP64H seed (2^-23) + two Newton steps, no special-case branch (x is a positive, normal number here).
HB_HD double fast_rcp(double x) {
#if defined(__CUDA_ARCH__)
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
    r = fma(fma(-x, r, 1.0), r, r);
    r = fma(fma(-x, r, 1.0), r, r);
    return r;
#else
    return 1.0 / x;
#endif
}
// x^(1/3) for x >= 0: single-precision seed 2^(log2(x)/3) + two Newton steps with a frozen derivative (error
// contracts by ~4e-6 per step -> < 1 ulp).  Below 1e-24 the result only ever feeds MIN_H-clamped coefficients.
HB_HD double fast_cbrt(double x) {
#if defined(__CUDA_ARCH__)
    if (!(x >= 1e-24)) return x != x ? x : 0.0;
    const float xf = (float)x;
    const float yf = exp2f(__log2f(xf) * 0.33333334f);
    const double y0 = (double)yf;
    double w;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(w) : "d"(3.0 * y0 * y0));
    double y = fma(-fma(y0 * y0, y0, -x), w, y0);
    y = fma(-fma(y * y, y, -x), w, y);
    return y;
#else
    return cbrt(x);
#endif
}

// ---- branch-free variants for the lean face path (hb_face_kernel<.., END = false>) ---------------------------------
// They trade the last bits of the Newton polish (results stay within ~1e-12 relative, six orders below the parity
// tolerance of 1e-6 degC on temperatures) for a shorter dependent chain and no slow-path branches.
#if defined(__CUDACC__)
// exp(r) on |r| <= ln2/2 as sum r^i / i!, i <= 12 (truncation 4e-16), + range-reduction constants
static __constant__ double hb_expc[18] = {
    1.0, 1.0, 1.0 / 2, 1.0 / 6, 1.0 / 24, 1.0 / 120, 1.0 / 720, 1.0 / 5040, 1.0 / 40320, 1.0 / 362880, 1.0 / 3628800,
    1.0 / 39916800, 1.0 / 479001600, 0.0,
    1.4426950408889634074,      // [14] log2(e)
    6755399441055744.0,         // [15] 1.5 * 2^52: round-to-nearest-integer shifter
    6.93147180369123816490e-01, // [16] ln2 high part (low 32 bits of the mantissa are zero: k*hi is exact)
    1.90821492927058770002e-10, // [17] ln2 low part
};
#endif
// 1/x, one Newton step on the MUFU seed: relative error ~1e-12.
HB_HD double rcp_lean(double x) {
#if defined(__CUDA_ARCH__)
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
    const double e = fma(-x, r, 1.0);
    return fma(fma(e, e, e), r, r);   // r(1 + e + e^2): third-order step for the price of one more FMA
#else
    return 1.0 / x;
#endif
}
// x^(1/3), x >= 0 (NaN propagates, x < 1e-24 gives 0 as in fast_cbrt): no branch.
HB_HD double cbrt_lean(double x) {
#if defined(__CUDA_ARCH__)
    const float xf = (float)x;
    float l, yf;
    asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(l) : "f"(xf));
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(yf) : "f"(l * 0.33333334f));
    const double y0 = (double)yf;
    const double y2 = y0 * y0;
    double w;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(w) : "d"(3.0 * y2));
    // one Newton step: seed error 3e-7 (lg2/ex2.approx), derivative error 1e-6 (rcp.approx) -> 4e-13 relative
    const double y = fma(-fma(y2, y0, -x), w, y0);
    return x >= 1e-24 ? y : (x != x ? x : 0.0);
#else
    return x >= 1e-24 ? cbrt(x) : (x != x ? x : 0.0);
#endif
}
// exp(x) for any x (clamped to +-700; NaN propagates): Cody-Waite reduction, degree-12 Estrin, exponent add.
HB_HD double exp_lean(double x) {
#if defined(__CUDA_ARCH__)
    const double xc = x < -700. ? -700. : x > 700. ? 700. : x;
    const double t = fma(xc, hb_expc[14], hb_expc[15]);
    const double k = t - hb_expc[15];
    double r = fma(-k, hb_expc[16], xc);
    r = fma(-k, hb_expc[17], r);
    const double r2 = r * r, r4 = r2 * r2, r8 = r4 * r4;
    const double q0 = fma(hb_expc[1], r, hb_expc[0]), q1 = fma(hb_expc[3], r, hb_expc[2]);
    const double q2 = fma(hb_expc[5], r, hb_expc[4]), q3 = fma(hb_expc[7], r, hb_expc[6]);
    const double q4 = fma(hb_expc[9], r, hb_expc[8]), q5 = fma(hb_expc[11], r, hb_expc[10]);
    const double s0 = fma(q1, r2, q0), s1 = fma(q3, r2, q2), s2 = fma(q5, r2, q4);
    const double lo = fma(s1, r4, s0), hi = fma(hb_expc[12], r4, s2);
    const double p = fma(hi, r8, lo);
    const int ki = __double2loint(t);   // k as an integer (low word of the shifted sum)
    const double res = __hiloint2double(__double2hiint(p) + (ki << 20), __double2loint(p));
    return x == x ? res : x;
#else
    return exp(x);
#endif
}

HB_HD void sub_weather(double t_out, double wdir_deg, double wspeed, SubWeather& w) {
    w.t_out = t_out;
    w.sqrt_ws = sqrt(wspeed);
    const double a = wdir_deg * (kPi / 180.0);
#if defined(__CUDA_ARCH__)
    sincos(a, &w.sin_wd, &w.cos_wd);
#else
    w.sin_wd = sin(a);
    w.cos_wd = cos(a);
#endif
}

// convection.rs:87-110: h = k·|ΔT|^(1/3), k = 1.31 | 9.482/(7.238−|cosΣ|) | 1.81/(1.382+|cosΣ|) by the signs of ΔT and
// of the face's cos(tilt) (kpos/kneg, load_wallc), clamped at MIN_H = 0.1 (convection.rs:22).  NaN propagates.
HB_HD double tarp_natural(double t_air, double t_surf, double kpos, double kneg) {
    const double d = t_air - t_surf;
    const double ad = fabs(d);
    const double k = ad < 1e-3 ? 1.31 : d > 0. ? kpos : kneg;
    const double h = k * fast_cbrt(ad);
    return h < 0.1 ? 0.1 : h;
}

// the same for d = t_air − t_surf, branch-free (lean face path)
HB_HD double tarp_lean(double d, double kpos, double kneg) {
    const double ad = fabs(d);
    const double k = ad < 1e-3 ? 1.31 : d > 0. ? kpos : kneg;
    const double h = k * cbrt_lean(ad);
    return h < 0.1 ? 0.1 : h;
}

// gas.rs:45-74 — ascending-power polynomials in T[K]
struct GasDev {
    double l0, l1, m0, m1, c0, c1, mass;
};
HB_HD GasDev gas_of(int id) {
    switch (id) {
        case HB_GAS_ARGON: return GasDev{2.285e-3, 5.149e-5, 3.379e-6, 6.451e-8, 521.9285, 0., 39.948};
        case HB_GAS_KRYPTON: return GasDev{9.443e-4, 2.826e-5, 2.213e-6, 7.777e-8, 248.0907, 0., 83.8};
        case HB_GAS_XENON: return GasDev{4.538e-4, 1.723e-5, 1.069e-6, 7.414e-8, 158.3397, 0., 131.30};
        default: return GasDev{2.873e-3, 7.760e-5, 3.723e-6, 4.94e-8, 1002.7370, 1.2324e-2, 28.97};
    }
}

// gas.rs:285-309
HB_HD_NOINLINE double nu_90(double ra, double a_gi) {
    double nu1;
    if (ra <= 1e4)
        nu1 = 1. + 1.7596678 * 1e-10 * pow(ra, 2.2984755);
    else if (ra < 5e4)
        nu1 = 0.028154 * pow(ra, 0.4134);
    else
        nu1 = 0.0673838 * cbrt(ra);
    const double nu2 = 0.242 * pow(ra / a_gi, 0.272);
    return nu1 > nu2 ? nu1 : nu2;
}
// gas.rs:249-264
HB_HD_NOINLINE double nu_60(double ra, double a_gi) {
    const double g = 0.5 / pow(1. + pow(ra / 3160., 20.6), 0.1);
    const double base = 0.0936 * pow(ra, 0.314) / (1. + g);
    const double b2 = base * base, b4 = b2 * b2;
    const double nu1 = pow(1. + b4 * b2 * base, 1. / 7.);
    const double nu2 = (0.104 + 0.175 / a_gi) * pow(ra, 0.283);
    return nu1 > nu2 ? nu1 : nu2;
}
// gas.rs:197-315
HB_HD_NOINLINE double nusselt(double ra, double gamma_in, double a_gi, int& status) {
    const double THIRTY = 30. * kPi / 180.;
    const double EPS = 0.5 * kPi / 180.;
    const double gamma = fmod(gamma_in, kPi);
    if (gamma >= 0.0 && gamma < 2. * THIRTY - EPS) {
        const double cg = cos(gamma);
        const double x1 = 1. - 1708. / (ra * cg);
        const double a = (x1 + fabs(x1)) / 2.;
        const double b = 1. - 1708. * pow(sin(1.8 * gamma), 1.6) / (ra * cg);
        const double c = cbrt(ra * cg / 5830.) - 1.;
        return 1. + 1.44 * a * b + (c + fabs(c)) / 2.;
    } else if (gamma < 2. * THIRTY + EPS) {
        return nu_60(ra, a_gi);
    } else if (gamma < 3. * THIRTY - EPS) {
        const double n60 = nu_60(ra, a_gi), n90 = nu_90(ra, a_gi);
        const double x = (gamma - kPi / 3.) / (kPi / 2. - kPi / 3.);
        return n60 + (n90 - n60) * x;
    } else if (gamma < 3. * THIRTY + EPS) {
        return nu_90(ra, a_gi);
    } else if (gamma < 6. * THIRTY) {
        return 1. + (nu_90(ra, a_gi) - 1.) * sin(gamma);
    }
    status |= ST_BAD_GAMMA;
    return hb_nan();
}
// cavity.rs:59-69 + gas.rs:82-152
HB_HD_NOINLINE double cavity_u(const CavityDev& c, double t_front, double t_back, int& status) {
    const GasDev g = gas_of(c.gas);
    double gamma = c.angle;
    if (t_front > t_back) gamma = 180. * (kPi / 180.0) - gamma;
    const double a_gi = c.height / c.thickness;
    const double temp = ((t_front + 273.15) + (t_back + 273.15)) / 2.;
    const double lambda = g.l0 + g.l1 * temp;
    double ra;
    if (fabs(t_front - t_back) < 1e-10) {
        ra = 0.0000001;
    } else {
        const double beta = 1. / temp;
        const double cp = g.c1 == 0. ? g.c0 : g.c0 + g.c1 * temp;
        const double mu = g.m0 + g.m1 * temp;
        const double rho = 101325. * g.mass / (8314.46261815324 * temp);
        ra = (rho * rho) * ((c.thickness * c.thickness) * c.thickness) * 9.81 * beta * cp * fabs(t_front - t_back) /
             (mu * lambda);
    }
    const double nu = nusselt(ra, gamma, a_gi, status);
    const double conv = nu * lambda / c.thickness;
    const double tm = (t_back + t_front) / 2. + 273.15;
    const double rad = 4. * ((tm * tm) * tm) * kSigma * c.ein * c.eout / (1. - (1. - c.ein) * (1. - c.eout));
    return rad + conv;
}

struct Face {
    double air, trad, ts_env;  // ConvectionParams.air_temperature / rad_temperature / surface_temperature
    double h;
};

// surface.rs:596-717: environment and film coefficient of both faces for given end-node temperatures.
//   front: air = t_front; Outdoor → T_rad from the IR irradiance, forced + natural; Space/Ambient → T_rad = t_front.
//   back : same with t_back, except the AmbientTemperature arm, which the reference builds from FRONT values
//          (surface.rs:672-685): T_rad = t_front, surface temperature = front node.
HB_HD void border(const WallC& c, const WallS& w, double ts_f, double ts_b, Face& F, Face& Bk, int& status) {
    const int fl = c.fl;
    F.air = w.t_front;
    F.trad = (fl & WF_F_OUT) ? w.trad_out_f : w.t_front;
    F.ts_env = ts_f;
    double hf = tarp_natural(w.t_front, ts_f, c.kpos_f, c.kneg_f);
    if (fl & WF_F_OUT) hf += w.forced;
    Bk.air = w.t_back;
    Bk.trad = (fl & WF_B_OUT) ? w.trad_out_b : (fl & WF_B_AMB) ? w.t_front : w.t_back;
    Bk.ts_env = (fl & WF_B_AMB) ? ts_f : ts_b;
    double hb = tarp_natural(w.t_back, Bk.ts_env, c.kpos_b, c.kneg_b);
    if (fl & WF_B_OUT) hb += w.forced;
    if (hf != hf || hb != hb) status |= ST_NAN_H;   // surface.rs:704-707 (panic in the reference)
    F.h = (fl & WF_OV_F) ? c.ov_f : hf;             // surface.rs:708-714 (debug-only overrides)
    Bk.h = (fl & WF_OV_B) ? c.ov_b : hb;
}

// Access to one wall's node temperatures and wall-major constant arrays from the scalar lane (TILE_EXT tiles).
struct ExtWall {
    double* t;           // this wall's row of the shared-memory temperature mirror: slot s at t[(s % kNL) * G + s / kNL]
    double* a;           // same layout: coefficient A mirror (tiles with cavities)
    double* b;
    const double* u;     // wall-major: element of node i at [i * W]
    const double* mass;
    const double* af;
    const double* ab;
    int G, W;
    HB_HD int at(int node) const { const int s = node + 1; return (s % kNL) * G + s / kNL; }
    HB_HD double T(int node) const { return t[at(node)]; }
    HB_HD void setT(int node, double v) const { t[at(node)] = v; }
    HB_HD double link(int node) const { return u[(int64_t)node * W]; }
    HB_HD double m(int node) const { return mass[(int64_t)node * W]; }
};

// One no-mass chunk [ini, fin): damped fixed point around a tridiagonal solve (surface.rs:790-898).
// F/Bk are the border conditions evaluated with the previous sub-step's surface temperatures, hf/hb/rad_f/rad_b
// the film and radiative coefficients derived from them.
HB_HD_NOINLINE void solve_nomass(const WallC& c, WallS& w, const ExtWall& X, const CavityDev* cavs, int ini, int fin,
                                 const Face& F, const Face& Bk, double hf, double hb, double rad_f, double rad_b,
                                 double dt) {
    (void)dt;
    const int n = c.n;
    const int m = fin - ini;
    double tl[kMaxNomassChunk + 2];   // T[ini-1 .. fin]
    double ul[kMaxNomassChunk + 1];   // links ini-1 .. fin-1
    double sl[kMaxNomassChunk];       // absorbed solar per node
    int cavl[kMaxNomassChunk + 1];
    for (int j = 0; j < m + 2; ++j) {
        const int node = ini - 1 + j;
        tl[j] = (node >= 0 && node < n) ? X.T(node) : 0.;
    }
    for (int j = 0; j < m + 1; ++j) {
        const int link = ini - 1 + j;
        ul[j] = (link >= 0 && link < n - 1) ? X.link(link) : 0.;
        cavl[j] = -1;
        if (ul[j] < 0.)
            for (int q = c.cav_begin; q < c.cav_end; ++q)
                if (cavs[q].node == link) cavl[j] = q;
    }
    const bool dist = (c.fl & WF_DIST) != 0;
    for (int j = 0; j < m; ++j) {
        const int node = ini + j;
        if (dist)
            sl[j] = X.af[(int64_t)node * X.W] * w.solar_front + X.ab[(int64_t)node * X.W] * w.solar_back;
        else
            sl[j] = (node == 0 ? c.af0 * w.solar_front : 0.) + (node == n - 1 ? c.abL * w.solar_back : 0.);
    }
    double old_err = 99999.;
    int count = 0;
    while (true) {
        ++w.nm_iters;
        // assemble (discretization.rs:622-697)
        double dl[kMaxNomassChunk], dd[kMaxNomassChunk], du[kMaxNomassChunk], rhs[kMaxNomassChunk];
        for (int j = 0; j < m; ++j) { dl[j] = 0.; dd[j] = 0.; du[j] = 0.; rhs[j] = 0.; }
        for (int j = 0; j + 1 < m; ++j) {
            double uu = ul[j + 1];
            if (uu < 0.) uu = cavity_u(cavs[cavl[j + 1]], tl[j + 1], tl[j + 2], w.status);
            dd[j] += -uu;
            dd[j + 1] = dd[j + 1] - uu;
            du[j] = du[j] + uu;
            dl[j + 1] = dl[j + 1] + uu;
        }
        double hs_front, front_q;
        if (ini == 0) {
            front_q = F.air * hf + rad_f * (F.trad - tl[1]);
            hs_front = hf;
        } else {
            double uu = ul[0];
            if (uu < 0.) uu = cavity_u(cavs[cavl[0]], tl[0], tl[1], w.status);
            hs_front = uu;
            front_q = uu * tl[0];
        }
        rhs[0] += front_q;
        dd[0] += -hs_front;
        double hs_back, back_q;
        if (fin == n) {
            back_q = Bk.air * hb + rad_b * (Bk.trad - tl[m]);
            hs_back = hb;
        } else {
            double uu = ul[m];
            if (uu < 0.) uu = cavity_u(cavs[cavl[m]], tl[m], tl[m + 1], w.status);
            hs_back = uu;
            back_q = uu * tl[m + 1];
        }
        rhs[m - 1] += back_q;
        dd[m - 1] += -hs_back;
        for (int j = 0; j < m; ++j) rhs[j] = -(rhs[j] + sl[j]);  // surface.rs:828-832
        // 3-diagonal Gaussian elimination without pivoting (surface.rs:834)
        for (int j = 0; j + 1 < m; ++j) {
            const double f = dl[j + 1] / dd[j];
            dd[j + 1] -= f * du[j];
            rhs[j + 1] -= f * rhs[j];
        }
        for (int j = m - 1; j >= 0; --j) {
            double vv = rhs[j];
            if (j + 1 < m) vv -= du[j] * rhs[j + 1];
            rhs[j] = vv / dd[j];
        }
        double err = 0.0;
        for (int j = 0; j < m; ++j) err += fabs(rhs[j] - tl[j + 1]);
        if (err > old_err) break;                           // surface.rs:842-848
        // reference: panics on a NaN error (:850), then on the pass limit (surface.rs:869-873) — one cold branch here
        if (err != err || !(count < kNoMassIterCap)) { w.status |= err != err ? ST_NAN_ERR : ST_ITER_CAP; break; }
        for (int j = 0; j < m; ++j) tl[j + 1] = (tl[j + 1] + rhs[j]) * 0.5;  // :878-883
        const double tol = count < 100 ? 0.01 : 0.5;       // :885
        if (err / (double)m < tol) break;
        old_err = err;
        ++count;
    }
    for (int j = 0; j < m; ++j) X.setT(ini + j, tl[j + 1]);
}

// Main-step inputs → per-main-step derived quantities (surface.rs:916-931, 647, 692).
HB_HD void begin_main_step(const WallC& c, WallS& w) {
    double solar_front = w.sol_f;
    if (solar_front != solar_front || solar_front < 0.0) solar_front = 0.0;  // surface.rs:916-919
    double solar_back = w.sol_b;
    // surface.rs:920-923 tests `solar_front < 0.0` AFTER the front clamp above made it >= 0: only a NaN zeroes it
    if (solar_back != solar_back || solar_front < 0.0) solar_back = 0.0;
    w.solar_front = solar_front;
    w.solar_back = solar_back;
    w.trad_out_f = sqrt(sqrt(w.ir_f / kSigma)) - 273.15;
    w.trad_out_b = sqrt(sqrt(w.ir_b / kSigma)) - 273.15;
    (void)c;
}

// Before the nodes march: weather → boundary temperatures → border conditions with the OLD surface temperatures
// (surface.rs:939-948) → no-mass chunks / cavity links (TILE_EXT) → boundary links for the massive march.
template <bool EXT>
HB_HD void scalar_pre(const WallC& c, WallS& w, const SubWeather& sw, double zone_t_front, double zone_t_back,
                      double dt, const ExtWall& X, const CavityDev* cavs, const int32_t* nm_pairs, BoundaryLinks& out) {
    const int fl = c.fl;
    // surface.rs:37-46 (windward unless the tilted surface's normal points away from the wind) and the forced term
    // 2.537·Wf·Rf·sqrt(P·v/A) (convection.rs:157-163); only Outdoor faces use it
    const bool leeward = (fl & WF_TILTED) && !(add_rn(mul_rn(c.nx, sw.sin_wd), mul_rn(c.ny, sw.cos_wd)) > 0.0);
    w.forced = c.forced_c * (leeward ? 0.5 : 1.0) * sw.sqrt_ws;
    // boundary temperatures (model.rs:79-96)
    w.t_front = (fl & WF_F_OUT) ? sw.t_out : (fl & WF_F_SPACE) ? zone_t_front : c.amb_f;
    w.t_back = (fl & WF_B_OUT) ? sw.t_out : (fl & WF_B_SPACE) ? zone_t_back : c.amb_b;

    Face F, Bk;
    border(c, w, w.ts_f, w.ts_b, F, Bk, w.status);
    const double hf = F.h, hb = Bk.h;
    const double xr_f = 273.15 + (F.trad + F.ts_env) / 2.;
    const double xr_b = 273.15 + (Bk.trad + Bk.ts_env) / 2.;
    const double rad_f = c.eps4s_f * ((xr_f * xr_f) * xr_f);
    const double rad_b = c.eps4s_b * ((xr_b * xr_b) * xr_b);

    if (EXT) {
        // no-mass chunks first (surface.rs:950-965)
        for (int k = c.nm_begin; k < c.nm_end; ++k)
            solve_nomass(c, w, X, cavs, nm_pairs[2 * k], nm_pairs[2 * k + 1], F, Bk, hf, hb, rad_f, rad_b, dt);
        if (c.nm_end > c.nm_begin) {
            w.ts_f = X.T(0);
            w.ts_b = X.T(c.n - 1);
        }
        // cavity links: conductance at the assembly temperatures, scaled by dt/C of the two nodes
        for (int q = c.cav_begin; q < c.cav_end; ++q) {
            const CavityDev cv = cavs[q];
            const int i = cv.node;
            const double uu = cavity_u(cv, X.T(i), X.T(i + 1), w.status);
            const double mi = X.m(i), mj = X.m(i + 1);
            X.a[X.at(i)] = uu * (mi >= kMassiveThreshold ? dt / mi : 0.);
            X.b[X.at(i + 1)] = uu * (mj >= kMassiveThreshold ? dt / mj : 0.);
        }
    }
    // film coefficients as links to ghost nodes at the equivalent air temperature:
    //   h·(T_air − T_s) + rad·(T_rad − T_s⁰) + q_solar  =  h·(T_ghost − T_s)        (discretization.rs:658-697)
    // T_s⁰ is the surface temperature at assembly time: the radiative term is explicit in the reference.
    const double hf_s = hf > 1e-12 ? hf : 1e-12, hb_s = hb > 1e-12 ? hb : 1e-12;
    out.ghost_f = F.air + (rad_f * (F.trad - w.ts_f) + c.af0 * w.solar_front) * fast_rcp(hf_s);
    out.ghost_b = Bk.air + (rad_b * (Bk.trad - w.ts_b) + c.abL * w.solar_back) * fast_rcp(hb_s);
    out.bf = hf_s * c.v0;
    out.ab = hb_s * c.vL;
}

// After the nodes marched: film coefficients with the NEW surface temperatures, convective flows
// (model.rs:145-169) and this wall's contributions to calculate_zones_abc (model.rs:556-590).
HB_HD void scalar_post(const WallC& c, WallS& w, double ts_f, double ts_b, double dt, double (&pub)[4]) {
    w.ts_f = ts_f;
    w.ts_b = ts_b;
    Face F, Bk;
    border(c, w, ts_f, ts_b, F, Bk, w.status);
    w.h_f = F.h;
    w.h_b = Bk.h;
    w.flow_f = (ts_f - w.t_front) * w.h_f;
    w.flow_b = (ts_b - w.t_back) * w.h_b;
    w.e_f += w.flow_f * c.area * dt;
    w.e_b += w.flow_b * c.area * dt;
    pub[0] = w.h_f * c.area;
    pub[1] = w.h_f * c.area * ts_f;
    pub[2] = w.h_b * c.area;
    pub[3] = w.h_b * c.area * ts_b;
}

struct ZoneC {
    double vol, q, inf_v, inf_t, vent_v, vent_t;
    int32_t sb, se;
};

// model.rs:489-597 + 650-674, in three pieces so that a zone with a single wall can be advanced by that wall's lane.
// zone_sources: a, b of everything that is not a surface (model.rs:502-544).
HB_HD void zone_sources(const ZoneC& z, double& a, double& b) {
    a = 0.0;
    b = 0.0;
    a += z.q;
    if (z.inf_t == z.inf_t) {
        const double tk = z.inf_t + 273.15;
        const double cp = 1002.7370 + 1.2324e-2 * tk;
        const double rho = 101325. * 28.97 / (8314.46261815324 * tk);
        a += rho * z.inf_v * cp * z.inf_t;
        b += rho * z.inf_v * cp;
    }
    if (z.vent_t == z.vent_t) {
        const double tk = z.vent_t + 273.15;
        const double cp = 1002.7370 + 1.2324e-2 * tk;
        const double rho = 101325. * 28.97 / (8314.46261815324 * tk);
        a += rho * z.vent_v * cp * z.vent_t;
        b += rho * z.vent_v * cp;
    }
}
// Scheduled zone inputs of main step `step` (hb_series.zone_gain / zone_infiltration_* / zone_ventilation_*): what the
// reference's caller writes into SimulationState before a march call (model.rs:502-544).  Returns whether anything is
// scheduled (the caller then re-sums the non-surface sources in the reference's order, zone_sources).
HB_HD bool load_zone_schedule(const SeriesDev& P, int step, int64_t gz, ZoneC& z) {
    const int64_t o = (int64_t)step * P.n_dev_zones + gz;
    if (P.zone_gain) z.q = P.zone_gain[o];
    if (P.zone_inf_v) z.inf_v = P.zone_inf_v[o];
    if (P.zone_inf_t) z.inf_t = P.zone_inf_t[o];
    if (P.zone_vent_v) z.vent_v = P.zone_vent_v[o];
    if (P.zone_vent_t) z.vent_t = P.zone_vent_t[o];
    return P.zone_gain || P.zone_inf_v || P.zone_inf_t || P.zone_vent_v || P.zone_vent_t;
}

// 1/c with c = V·ρ(T)·cp(T) (zone.rs:59-65), ρ = 101325·28.97/(8314.46261815324·T_K)
HB_HD double zone_inv_c(const ZoneC& z, double tz) {
    const double tk = tz + 273.15;
    return (8314.46261815324 * tk) * fast_rcp(z.vol * (101325. * 28.97) * (1002.7370 + 1.2324e-2 * tk));
}
// model.rs:650-674: T' = a/b + (T − a/b)·exp(−b·dt/c) when |b| > 1e-9.  LEAN: branch-free, short-chain arithmetic.
template <bool LEAN = false>
HB_HD double zone_advance(double tz, double dt, double inv_cc, double a, double b) {
    if constexpr (LEAN) {
        const double ab = a * rcp_lean(b);
        const double tn = ab + (tz - ab) * exp_lean(-b * dt * inv_cc);
        return fabs(b) > 1e-9 ? tn : tz;
    } else {
        double tn = tz;
        if (fabs(b) > 1e-9) {
            const double ab = a * fast_rcp(b);   // b > 0: sums of h·A and ρ·V̇·cp
            tn = ab + (tz - ab) * exp(-b * dt * inv_cc);
        }
        return tn;
    }
}
template <bool LEAN = false>
HB_HD double zone_inv_c_t(double vol, double tz) {
    const double tk = tz + 273.15;
    const double den = vol * (101325. * 28.97) * (1002.7370 + 1.2324e-2 * tk);
    return (8314.46261815324 * tk) * (LEAN ? rcp_lean(den) : fast_rcp(den));
}
// The zone update with the non-surface sources (a0, b0 of zone_sources) already known.
template <bool LEAN = false>
HB_HD double zone_update_ab(const ZoneC& z, int zi, double tz, double dt, double a, double b, const double* pub_hA_f,
                            const double* pub_hAT_f, const double* pub_hA_b, const double* pub_hAT_b,
                            const int32_t* zf, const int32_t* zb) {
    const double inv_cc = zone_inv_c_t<LEAN>(z.vol, tz);
    if constexpr (LEAN) {
        // four independent partial sums (front/back x a/b), loads issued ahead of the adds: the sums differ from the
        // reference's strictly sequential order by rounding only
        double af = 0., bf = 0., ab = 0., bb = 0.;
#pragma unroll 4
        for (int j = z.sb; j < z.se; ++j) {
            const bool f = zf[j] == zi, k = zb[j] == zi;
            const double haf = pub_hA_f[j], hatf = pub_hAT_f[j], hab = pub_hA_b[j], hatb = pub_hAT_b[j];
            af += f ? hatf : 0.;
            bf += f ? haf : 0.;
            ab += k ? hatb : 0.;
            bb += k ? hab : 0.;
        }
        a += af + ab;
        b += bf + bb;
    } else {
#pragma unroll 1
        for (int j = z.sb; j < z.se; ++j) {   // surfaces then fenestrations, front face before back face
            if (zf[j] == zi) { a += pub_hAT_f[j]; b += pub_hA_f[j]; }
            if (zb[j] == zi) { a += pub_hAT_b[j]; b += pub_hA_b[j]; }
        }
    }
    return zone_advance<LEAN>(tz, dt, inv_cc, a, b);
}
// pub_* are the per-wall contributions of the tile, zf/zb the walls' zone indices.
template <bool LEAN = false>
HB_HD double zone_update(const ZoneC& z, int zi, double tz, double dt, const double* pub_hA_f, const double* pub_hAT_f,
                         const double* pub_hA_b, const double* pub_hAT_b, const int32_t* zf, const int32_t* zb) {
    double a, b;
    zone_sources(z, a, b);
    return zone_update_ab<LEAN>(z, zi, tz, dt, a, b, pub_hA_f, pub_hAT_f, pub_hA_b, pub_hAT_b, zf, zb);
}

// ------------------------------------------------------------------------------------------------------------
// Per-FACE formulation of the same scalar physics (hb_face_kernel): the front and the back face of a wall are
// evaluated by two different lanes (two warps of the block), which halves the dependent chain that separates two
// node marches.  Arithmetic and operation order are those of border()/scalar_pre()/scalar_post() above.
// ------------------------------------------------------------------------------------------------------------
enum : int32_t {
    FC_OUT = 1, FC_SPACE = 2, FC_BACK = 4, FC_TILTED = 8, FC_OV = 16, FC_NM = 32,
    FC_BACK_AMB = 64,        // back face, Boundary::AmbientTemperature: built from FRONT values (surface.rs:672-685)
    FC_FRONT_OUT = 128, FC_FRONT_SPACE = 256,   // kind of the wall's front boundary (needed by FC_BACK_AMB)
    FC_NM2 = 1024,           // FC_NM with a chunk of TWO no-mass nodes at this end (end node + its neighbour)
    FC_PANE = 512,           // face of a two-node all-no-mass wall (single glass pane): both faces solve one 2x2 system
    FC_WIN4 = 2048,          // face of a double-glazed window (pane | gas | pane, four no-mass nodes; warp tiles)
};
// Constants of one face of one wall.
struct FaceC {
    double kpos, kneg;     // TARP natural-convection factors of this face (load_wallc)
    double eps4s;          // 4·ε·σ of this face
    double area;
    double v_end;          // dt/C of the face's end node (0: no mass)
    double alpha_end;      // solar absorptance landing on the end node from this side
    double alpha_x;        // … and from the OTHER side's irradiance (FC_PANE: distributed absorption, else 0)
    double amb;            // AmbientTemperature of this boundary
    double ov;             // film-coefficient override (debug-only field of the reference) when FC_OV
    double forced_c, nx, ny;
    double eps;            // emissivity of this face (front: IR net-gain conversion)
    double u_nb;           // conductance of the link between the end node and its neighbour (FC_NM)
    double u_nb2;          // … between that neighbour and the next node (FC_NM2)
    double amb_front;      // FC_BACK_AMB: the front boundary's ambient temperature
    int32_t fl, zone, zone_front, site, desc_index, n;
};
HB_HD void load_facec(const WallC& c, bool back, double u_nb, bool nm, FaceC& f, bool pane = false,
                      double alpha_own = 0., double alpha_x = 0., bool nm2 = false, double u_nb2 = 0., bool win4 = false) {
    const int wf = c.fl;
    int fl = back ? FC_BACK : 0;
    if (wf & (back ? WF_B_OUT : WF_F_OUT)) fl |= FC_OUT;
    if (wf & (back ? WF_B_SPACE : WF_F_SPACE)) fl |= FC_SPACE;
    if (wf & WF_TILTED) fl |= FC_TILTED;
    if (wf & (back ? WF_OV_B : WF_OV_F)) fl |= FC_OV;
    if (back && (wf & WF_B_AMB)) fl |= FC_BACK_AMB;
    if (wf & WF_F_OUT) fl |= FC_FRONT_OUT;
    if (wf & WF_F_SPACE) fl |= FC_FRONT_SPACE;
    if (nm) fl |= FC_NM;
    if (pane) fl |= FC_PANE;
    if (win4) fl |= FC_WIN4;
    if (nm && nm2) fl |= FC_NM2;
    f.fl = fl;
    f.kpos = back ? c.kpos_b : c.kpos_f;
    f.kneg = back ? c.kneg_b : c.kneg_f;
    f.eps4s = back ? c.eps4s_b : c.eps4s_f;
    f.area = c.area;
    f.v_end = back ? c.vL : c.v0;
    f.alpha_end = pane || win4 ? alpha_own : back ? c.abL : c.af0;
    f.alpha_x = pane || win4 ? alpha_x : 0.;
    f.amb = back ? c.amb_b : c.amb_f;
    f.ov = back ? c.ov_b : c.ov_f;
    f.forced_c = c.forced_c; f.nx = c.nx; f.ny = c.ny;
    f.eps = c.eps_f;
    f.u_nb = u_nb;
    f.u_nb2 = u_nb2;
    f.amb_front = c.amb_f;
    f.zone = back ? c.zb : c.zf;
    f.zone_front = c.zf;
    f.site = c.site; f.desc_index = c.desc_index; f.n = c.n;
}
// Dynamic state of one face (registers of the face lane).
struct FaceS {
    double h, flow, sol, ir;   // this face's 4 SimulationState slots (surface.rs:428-442)
    double e;                  // ∫ flow·A dt
    double ts;                 // temperature of the face's end node
    double solar;              // clamped irradiance of this main step (surface.rs:916-923)
    double trad_out;           // outdoor radiant temperature of this main step (surface.rs:647,692)
    double air, forced;        // carried from the pre- to the post-march phase of a sub-step
    unsigned long long nm_iters;
    int32_t status;
};
// surface.rs:916-931,647,692 for one face.  The reference's back clamp tests `solar_front < 0.0` (surface.rs:921) after
// the front clamp has already made solar_front >= 0, so the back irradiance is zeroed by a NaN only.
HB_HD void face_begin_main_step(const FaceC& c, FaceS& w) {
    double s = w.sol;
    if (c.fl & FC_BACK) {
        if (s != s) s = 0.0;
    } else {
        if (s != s || s < 0.0) s = 0.0;
    }
    w.solar = s;
    w.trad_out = sqrt(sqrt(w.ir / kSigma)) - 273.15;
}
// Film coefficient of a face for end-node temperature ts_env (border() for one face).
HB_HD double face_h(const FaceC& c, const FaceS& w, double ts_env, int& status) {
    double h = tarp_natural(w.air, ts_env, c.kpos, c.kneg);
    if (c.fl & FC_OUT) h += w.forced;
    if (h != h) status |= ST_NAN_H;
    return (c.fl & FC_OV) ? c.ov : h;
}
// One no-mass end node (chunk (0,1) or (n-1,n)): damped fixed point of surface.rs:790-898 on a 1x1 system.
//   K = -(h + u),  q = air·h + rad·(trad - T) + u·T_neighbour (+ absorbed solar)
template <bool LEAN = false>
HB_HD double solve_end_node(double air, double h, double rad, double trad, double u, double t_nb, double sol, double t,
                            unsigned long long& iters, int& status) {
    const double dd = -h - u;
    const double qn = u * t_nb;
    double old_err = 99999.;
    int count = 0;
    if constexpr (LEAN) {
        // the update x(t) = K⁻¹(−q(t)) is affine in t: x = alpha + beta·t.  Same iteration and the same stopping rules.
        // The first pass (all there is in ~87 % of the solves: sub-steps are short) evaluates x as above; from the second
        // pass on the residual e = x − t is carried instead of recomputed — it contracts by rho = (1 + beta)/2 per
        // pass, t moves by e/2 — so that a pass is one multiplication on the dependent chain (≈ 35 cycles) instead of
        // five operations (≈ 90): the walls that need 5–10 passes bound the longest tiles of a series.  Identical in
        // exact arithmetic; in floating point the iterates differ in the last bits (far below 1e-6 °C).
        const double r = -rcp_lean(h + u);
        const double beta = rad * r;
        const double alpha = -((((air * h + rad * trad) + qn) + sol) * r);
        ++iters;
        double e = fma(beta, t, alpha) - t;
        double err = fabs(e);
        if (err > old_err) return t;
        if (err != err) { status |= ST_NAN_ERR; return t; }
        t = fma(0.5, e, t);
        if (err < 0.01) return t;
        const double rho = fma(0.5, beta, 0.5);
        old_err = err;
        count = 1;
        while (true) {
            ++iters;
            e *= rho;
            err = fabs(e);
            if (err > old_err) break;
            // reference: panics on a NaN error (:850), then on the pass limit (surface.rs:869-873) — one cold branch here
            if (err != err || !(count < kNoMassIterCap)) { status |= err != err ? ST_NAN_ERR : ST_ITER_CAP; break; }
            t = fma(0.5, e, t);
            const double tol = count < 100 ? 0.01 : 0.5;
            if (err < tol) break;
            old_err = err;
            ++count;
        }
        return t;
    }
    const double r = 1.0 / dd;
    while (true) {
        ++iters;
        const double rhs = -(((air * h + rad * (trad - t)) + qn) + sol);
        const double x = rhs * r;
        const double err = fabs(x - t);
        if (err > old_err) break;
        // reference: panics on a NaN error (:850), then on the pass limit (surface.rs:869-873) — one cold branch here
        if (err != err || !(count < kNoMassIterCap)) { status |= err != err ? ST_NAN_ERR : ST_ITER_CAP; break; }
        t = (t + x) * 0.5;
        const double tol = count < 100 ? 0.01 : 0.5;
        if (err < tol) break;
        old_err = err;
        ++count;
    }
    return t;
}
// Before the nodes march (scalar_pre for one face).  zone_t: temperature of the zone this face looks at (FC_SPACE);
// t_front/ts_front: air temperature and end-node temperature of the wall's FRONT face (FC_BACK_AMB only);
// t_nb: neighbour node of a no-mass end node (FC_NM).  Returns ghost temperature and link coefficient for the
// march, and (FC_NM) the end node's new temperature in w.ts.
struct FaceEnv {
    double h, rad, trad;   // film coefficient, radiative coefficient, radiant temperature (old surface temperatures)
};
HB_HD void face_env(const FaceC& c, FaceS& w, const SubWeather& sw, double zone_t, double t_front, double ts_front,
                    FaceEnv& e) {
    const int fl = c.fl;
    const bool leeward = (fl & FC_TILTED) && !(add_rn(mul_rn(c.nx, sw.sin_wd), mul_rn(c.ny, sw.cos_wd)) > 0.0);
    w.forced = c.forced_c * (leeward ? 0.5 : 1.0) * sw.sqrt_ws;
    w.air = (fl & FC_OUT) ? sw.t_out : (fl & FC_SPACE) ? zone_t : c.amb;
    e.trad = (fl & FC_OUT) ? w.trad_out : (fl & FC_BACK_AMB) ? t_front : w.air;
    const double ts_env = (fl & FC_BACK_AMB) ? ts_front : w.ts;
    e.h = face_h(c, w, ts_env, w.status);
    const double xr = 273.15 + (e.trad + ts_env) / 2.;
    e.rad = c.eps4s * ((xr * xr) * xr);
}
// A single pane (two no-mass nodes, chunk (0, 2)): solve_nomass for m = 2 written out — tridiagonal elimination of
//   [-(u+h_f)  u; u  -(u+h_b)] x = -(q + absorbed solar)  inside the damped fixed point (surface.rs:790-898).
// sol0/sol1: absorbed solar of node 0 / node 1 (front_alpha·I_front + back_alpha·I_back, discretization.rs).
template <bool LEAN = false>
HB_HD void solve_pane(double air_f, double h_f, double rad_f, double trad_f, double sol0, double air_b, double h_b,
                      double rad_b, double trad_b, double sol1, double u, double& t0, double& t1,
                      unsigned long long& iters, int& status) {
    const double dd0 = -u - h_f, dd1 = -u - h_b;
    const double i0 = LEAN ? -fast_rcp(-dd0) : 1.0 / dd0;
    const double f = LEAN ? u * i0 : u / dd0;
    const double dd1e = dd1 - f * u;
    const double i1 = LEAN ? -fast_rcp(-dd1e) : 1.0 / dd1e;
    if constexpr (LEAN) {
        // the elimination is affine in (t0, t1): x1 = c1 + m10 t0 + m11 t1, x0 = c0 + n00 t0 + n01 x1 — the same
        // iteration and stopping rules on a chain of three dependent operations per pass
        const double k0 = -((air_f * h_f + rad_f * trad_f) + sol0), k1 = -((air_b * h_b + rad_b * trad_b) + sol1);
        const double c1 = (k1 - f * k0) * i1, m10 = -(f * rad_f) * i1, m11 = rad_b * i1;
        const double c0 = k0 * i0, n00 = rad_f * i0, n01 = -(u * i0);
        double old_err = 99999.;
        int count = 0;
        while (true) {
            ++iters;
            const double x1 = fma(m11, t1, fma(m10, t0, c1));
            const double x0 = fma(n01, x1, fma(n00, t0, c0));
            const double err = fabs(x0 - t0) + fabs(x1 - t1);
            if (err > old_err) break;
            // reference: panics on a NaN error (:850), then on the pass limit (surface.rs:869-873) — one cold branch here
            if (err != err || !(count < kNoMassIterCap)) { status |= err != err ? ST_NAN_ERR : ST_ITER_CAP; break; }
            t0 = (t0 + x0) * 0.5;
            t1 = (t1 + x1) * 0.5;
            const double tol = count < 100 ? 0.01 : 0.5;
            if (err * 0.5 < tol) break;
            old_err = err;
            ++count;
        }
        return;
    }
    double old_err = 99999.;
    int count = 0;
    while (true) {
        ++iters;
        const double r0 = -((air_f * h_f + rad_f * (trad_f - t0)) + sol0);
        double r1 = -((air_b * h_b + rad_b * (trad_b - t1)) + sol1);
        r1 -= f * r0;
        const double x1 = LEAN ? r1 * i1 : r1 / dd1e;
        const double x0 = LEAN ? (r0 - u * x1) * i0 : (r0 - u * x1) / dd0;
        const double err = fabs(x0 - t0) + fabs(x1 - t1);
        if (err > old_err) break;
        // reference: panics on a NaN error (:850), then on the pass limit (surface.rs:869-873) — one cold branch here
        if (err != err || !(count < kNoMassIterCap)) { status |= err != err ? ST_NAN_ERR : ST_ITER_CAP; break; }
        t0 = (t0 + x0) * 0.5;
        t1 = (t1 + x1) * 0.5;
        const double tol = count < 100 ? 0.01 : 0.5;
        if (err / 2.0 < tol) break;
        old_err = err;
        ++count;
    }
}
// A double-glazed window: pane | gas | pane, four no-mass nodes, one chunk (0, 4) — solve_nomass written out for m = 4,
// ini = 0, fin = n, with the cavity's temperature-dependent conductance (cavity.rs:59-69, gas.rs:82-315) re-evaluated
// in every pass of the damped fixed point (surface.rs:790-898 around discretization.rs:596-700).
// u01 / u23: conductance of the panes; sol[j]: absorbed solar of node j; t[0..3] in/out.
HB_HD_NOINLINE void solve_window4(double air_f, double h_f, double rad_f, double trad_f, double air_b, double h_b,
                                  double rad_b, double trad_b, const double (&sol)[4], double u01, double u23,
                                  const CavityDev& cav, double (&t)[4], unsigned long long& iters, int& status) {
    double old_err = 99999.;
    int count = 0;
    while (true) {
        ++iters;
        const double u12 = cavity_u(cav, t[1], t[2], status);
        // assemble (discretization.rs:622-697): interior links, then the film edges
        double dd[4] = {0., 0., 0., 0.}, du[3], rhs[4] = {0., 0., 0., 0.};
        const double uu[3] = {u01, u12, u23};
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            dd[j] += -uu[j];
            dd[j + 1] = dd[j + 1] - uu[j];
            du[j] = uu[j];
        }
        rhs[0] += air_f * h_f + rad_f * (trad_f - t[0]);
        dd[0] += -h_f;
        rhs[3] += air_b * h_b + rad_b * (trad_b - t[3]);
        dd[3] += -h_b;
#pragma unroll
        for (int j = 0; j < 4; ++j) rhs[j] = -(rhs[j] + sol[j]);   // surface.rs:828-832
        // 3-diagonal Gaussian elimination without pivoting (surface.rs:834); the lower diagonal equals the upper one
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            const double f = du[j] / dd[j];
            dd[j + 1] -= f * du[j];
            rhs[j + 1] -= f * rhs[j];
        }
        rhs[3] = rhs[3] / dd[3];
#pragma unroll
        for (int j = 2; j >= 0; --j) rhs[j] = (rhs[j] - du[j] * rhs[j + 1]) / dd[j];
        double err = 0.0;
#pragma unroll
        for (int j = 0; j < 4; ++j) err += fabs(rhs[j] - t[j]);
        if (err > old_err) break;                         // surface.rs:842-848
        // reference: panics on a NaN error (:850), then on the pass limit (surface.rs:869-873) — one cold branch here
        if (err != err || !(count < kNoMassIterCap)) { status |= err != err ? ST_NAN_ERR : ST_ITER_CAP; break; }
#pragma unroll
        for (int j = 0; j < 4; ++j) t[j] = (t[j] + rhs[j]) * 0.5;   // :878-883
        const double tol = count < 100 ? 0.01 : 0.5;     // :885
        if (err / 4.0 < tol) break;
        old_err = err;
        ++count;
    }
}

// Two no-mass nodes at an end of a wall (chunk (0,2) or (n-2,n)): the chunk's 2x2 system is solve_pane's with the
// inner side replaced by the link to the massive neighbour (h := u2, air := T_neighbour, no radiation, no solar);
// the chunk's first node is the end node at the front, the inner node at the back (solve_nomass's ordering).
template <bool LEAN = false>
HB_HD void solve_end_pair(bool back, double air, double h, double rad, double trad, double sol, double u1, double u2,
                          double t_nb2, double& t_end, double& t_inner, unsigned long long& iters, int& status) {
    if (back)
        solve_pane<LEAN>(t_nb2, u2, 0., 0., 0., air, h, rad, trad, sol, u1, t_inner, t_end, iters, status);
    else
        solve_pane<LEAN>(air, h, rad, trad, sol, t_nb2, u2, 0., 0., 0., u1, t_end, t_inner, iters, status);
}
// t_nb: the end node's neighbour (FC_NM; for FC_NM2 in/out: the chunk's inner node); t_nb2: the massive node behind it.
HB_HD void face_links(const FaceC& c, FaceS& w, const FaceEnv& e, double& t_nb, double t_nb2, double& ghost,
                      double& link) {
    if (c.fl & FC_NM2)
        solve_end_pair(c.fl & FC_BACK, w.air, e.h, e.rad, e.trad, c.alpha_end * w.solar, c.u_nb, c.u_nb2, t_nb2, w.ts,
                       t_nb, w.nm_iters, w.status);
    else if (c.fl & FC_NM)
        w.ts = solve_end_node(w.air, e.h, e.rad, e.trad, c.u_nb, t_nb, c.alpha_end * w.solar, w.ts, w.nm_iters,
                              w.status);
    const double hs = e.h > 1e-12 ? e.h : 1e-12;
    ghost = w.air + (e.rad * (e.trad - w.ts) + c.alpha_end * w.solar) * fast_rcp(hs);
    link = hs * c.v_end;
}
HB_HD void face_pre(const FaceC& c, FaceS& w, const SubWeather& sw, double zone_t, double t_front, double ts_front,
                    double t_nb, double& ghost, double& link) {
    FaceEnv e;
    face_env(c, w, sw, zone_t, t_front, ts_front, e);
    face_links(c, w, e, t_nb, 0., ghost, link);
}
// After the nodes marched (scalar_post for one face): new film coefficient, convective flow, energy integral and
// this face's contribution (h·A, h·A·T_s) to calculate_zones_abc.
HB_HD void face_post(const FaceC& c, FaceS& w, double ts, double ts_front, double dt, double& hA, double& hAT) {
    w.ts = ts;
    const double ts_env = (c.fl & FC_BACK_AMB) ? ts_front : ts;
    w.h = face_h(c, w, ts_env, w.status);
    w.flow = (ts - w.air) * w.h;
    w.e += w.flow * c.area * dt;
    hA = w.h * c.area;
    hAT = w.h * c.area * ts;
}

// ------------------------------------------------------------------------------------------------------------
// Massive nodes: classic RK4 on  Ṫ = C⁻¹(K·T + q)  (surface.rs:228-308) for one lane's kNL slots.
// K and q are frozen during a sub-step, so the four stages collapse to
//     T⁺ = T + p + M p/2 + M² p/6 + M³ p/24,   p = M·T (ghost nodes carry q),   M = diag(dt/C)·K,
// evaluated as four tridiagonal sweeps in flux form  (M x)_i = A_i (x_{i+1} − x_i) − B_i (x_i − x_{i−1}).
// `left`/`right` are the neighbouring lanes' edge values of x (halo); ghost and padding slots have A = B = 0.
// ------------------------------------------------------------------------------------------------------------
template <int NL>
HB_HD void sweep_first(double (&T)[NL], double (&V)[NL], const double (&A)[NL], const double (&B)[NL], double left,
                       double right) {
    double dprev = T[0] - left;
#pragma unroll
    for (int i = 0; i < NL; ++i) {
        const double dnext = (i + 1 < NL ? T[i + 1 < NL ? i + 1 : i] : right) - T[i];
        V[i] = A[i] * dnext - B[i] * dprev;
        dprev = dnext;
    }
}
template <int NL>
HB_HD void sweep_next(double (&T)[NL], double (&V)[NL], const double (&A)[NL], const double (&B)[NL], double left,
                      double right, double c) {
    double dprev = V[0] - left;
#pragma unroll
    for (int i = 0; i < NL; ++i) {
        const double dnext = (i + 1 < NL ? V[i + 1 < NL ? i + 1 : i] : right) - V[i];
        V[i] = A[i] * dnext - B[i] * dprev;
        dprev = dnext;
        T[i] = fma(c, V[i], T[i]);
    }
}

}  // namespace hb

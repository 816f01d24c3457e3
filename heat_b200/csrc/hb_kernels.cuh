// hb_kernels.cuh — hand-written CUDA (sm_100a) for the wall-conduction time-stepping path.
//
// K1/K2  hb_march_kernel<NMAX, DIST, LOCAL>: persistent over all main steps and sub-steps of a series.
//        One CUDA block = one tile of independent models; one thread = one surface (wall, roof, window…),
//        whose node temperatures stay in registers across sub-steps; the first n_zones threads also own the
//        tile's zones (shared memory).  Per sub-step (reference call stack):
//          boundary temperatures (model.rs:79-96) → border conditions / TARP film coefficients
//          (surface.rs:596-717, convection.rs:87-168) → radiative coefficients (surface.rs:941-948) →
//          no-mass chunks: damped fixed point around a tridiagonal solve (surface.rs:790-898) →
//          massive nodes: RK4 on  Ṫ = C⁻¹(K·T + q)  (surface.rs:168-308, discretization.rs:596-700) →
//          film coefficients with the new surface temperatures + convective flows (model.rs:145-169) →
//          per-zone a,b,c sums over the model's surfaces and analytic update (model.rs:489-597,650-674).
// K3     hb_reduce_kernel: ensemble sums / extrema feeding hb_reduce() (and NCCL in the host layer).
//
// Arithmetic notes (all f64).  The tridiagonal row is evaluated in flux form,
//   k_i = a_i·(x_{i+1} − x_i) − b_i·(x_i − x_{i−1}),  a_i = u_i·dt/C_i,  b_i = u_{i−1}·dt/C_i,
// with the film coefficients as links to virtual end nodes holding the equivalent air temperature
//   T_eq = T_air + (rad_hs·(T_rad − T_s) + solar)/h.
// a_i and b_i are bitwise the reference's scaled off-diagonals (surface.rs:168-187); the diagonal is never
// formed, which changes rounding by O(1e-16) relative — far inside the 1e-6 °C parity tolerance.
// pow(x,1/3) → cbrt(x), pow(x,1/4) → sqrt(sqrt(x)) (≤ 1 ulp apart); a no-mass node is a node with dt/C = 0.
#pragma once
#include <cuda_runtime.h>
#include <math_constants.h>

#include "hb_dev.hpp"

namespace hb {

// ----------------------------------------------------------------------------------------------------
// small device helpers
// ----------------------------------------------------------------------------------------------------
// Register-resident arrays may only be indexed with compile-time constants: dynamic reads/writes go through a
// fully unrolled select chain.  LOCAL tiers (large walls) keep their arrays in local memory and index directly.
template <int NMAX, bool LOCAL>
__device__ __forceinline__ double reg_get(const double (&A)[NMAX], int idx) {
    if constexpr (LOCAL) return A[idx];
    double v = 0.0;
#pragma unroll
    for (int i = 0; i < NMAX; ++i)
        if (i == idx) v = A[i];
    return v;
}
template <int NMAX, bool LOCAL>
__device__ __forceinline__ void reg_set(double (&A)[NMAX], int idx, double v) {
    if constexpr (LOCAL) {
        A[idx] = v;
        return;
    }
#pragma unroll
    for (int i = 0; i < NMAX; ++i)
        if (i == idx) A[i] = v;
}

// convection.rs:87-110 with the per-surface constants 9.482/(7.238-|cosΣ|) and 1.81/(1.382+|cosΣ|) hoisted.
__device__ __forceinline__ double tarp_natural(double t_air, double t_surf, double tilt, double c_same,
                                               double c_opp) {
    const double d = t_air - t_surf;
    const double ad = fabs(d);
    const double r = cbrt(ad);
    double h;
    if (ad < 1e-3 || fabs(tilt) < 1e-3)
        h = 1.31 * r;
    else if ((d < 0. && tilt < 0.) || (d > 0. && tilt > 0.))
        h = c_same * r;
    else
        h = c_opp * r;
    return h < 0.1 ? 0.1 : h;  // MIN_H, convection.rs:22; NaN propagates and is flagged by the caller
}

// gas.rs:45-74 — ascending-power polynomials in T[K]
struct GasDev {
    double l0, l1, m0, m1, c0, c1, mass;
};
__device__ __forceinline__ GasDev gas_of(int id) {
    switch (id) {
        case HB_GAS_ARGON: return GasDev{2.285e-3, 5.149e-5, 3.379e-6, 6.451e-8, 521.9285, 0., 39.948};
        case HB_GAS_KRYPTON: return GasDev{9.443e-4, 2.826e-5, 2.213e-6, 7.777e-8, 248.0907, 0., 83.8};
        case HB_GAS_XENON: return GasDev{4.538e-4, 1.723e-5, 1.069e-6, 7.414e-8, 158.3397, 0., 131.30};
        default: return GasDev{2.873e-3, 7.760e-5, 3.723e-6, 4.94e-8, 1002.7370, 1.2324e-2, 28.97};
    }
}

// gas.rs:285-309
static __device__ double nu_90(double ra, double a_gi) {
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
static __device__ double nu_60(double ra, double a_gi) {
    const double g = 0.5 / pow(1. + pow(ra / 3160., 20.6), 0.1);
    const double base = 0.0936 * pow(ra, 0.314) / (1. + g);
    const double b2 = base * base, b4 = b2 * b2;
    const double nu1 = pow(1. + b4 * b2 * base, 1. / 7.);
    const double nu2 = (0.104 + 0.175 / a_gi) * pow(ra, 0.283);
    return nu1 > nu2 ? nu1 : nu2;
}
// gas.rs:197-315
static __device__ double nusselt(double ra, double gamma_in, double a_gi, int& status) {
    const double PI = 3.14159265358979323846;
    const double THIRTY = 30. * PI / 180.;
    const double EPS = 0.5 * PI / 180.;
    const double gamma = fmod(gamma_in, PI);
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
        const double x = (gamma - PI / 3.) / (PI / 2. - PI / 3.);
        return n60 + (n90 - n60) * x;
    } else if (gamma < 3. * THIRTY + EPS) {
        return nu_90(ra, a_gi);
    } else if (gamma < 6. * THIRTY) {
        return 1. + (nu_90(ra, a_gi) - 1.) * sin(gamma);
    }
    status |= ST_BAD_GAMMA;
    return CUDART_NAN;
}
// cavity.rs:59-69 + gas.rs:82-152
static __device__ double cavity_u(const CavityDev& c, double t_front, double t_back, int& status) {
    const double PI = 3.14159265358979323846;
    const GasDev g = gas_of(c.gas);
    double gamma = c.angle;
    if (t_front > t_back) gamma = 180. * (PI / 180.0) - gamma;
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

// ----------------------------------------------------------------------------------------------------
// K1/K2
// ----------------------------------------------------------------------------------------------------
template <int NMAX, bool DIST, bool LOCAL>
__global__ void __launch_bounds__(kBlock) hb_march_kernel(DevView V, SeriesDev P, const int32_t* tile_ids) {
    const double PI = 3.14159265358979323846;
    const Tile tile = V.tiles[tile_ids[blockIdx.x]];
    const int t = threadIdx.x;
    const int64_t g = (int64_t)tile.surface_begin + t;
    const int n = V.n_nodes[g];              // 0 on padding threads
    const bool surf = n > 0;
    const bool zthread = t < tile.n_zones;
    const int64_t gz = (int64_t)tile.zone_begin + t;

    __shared__ double s_zT[kBlock];
    __shared__ double s_pub[4][kBlock];     // hA_f, hA·T_f, hA_b, hA·T_b
    __shared__ int32_t s_zf[kBlock], s_zb[kBlock];

    // ---- per-surface constants ----
    const int flags = surf ? V.flags[g] : 0;
    const int kind_f = flags & 15, kind_b = (flags >> 4) & 15;
    const bool has_cav = (flags >> 9) & 1;
    const int zf = surf ? V.zone_f[g] : -1, zb = surf ? V.zone_b[g] : -1;
    const int S = surf ? V.S[g] : 0;
    const double dt = surf ? V.dt[g] : 0.;
    const double area = surf ? V.area[g] : 0.;
    const double cos_tilt = surf ? V.cos_tilt[g] : 0.;
    const double act = fabs(cos_tilt);
    const double c_same = 9.482 / (7.238 - act), c_opp = 1.81 / (1.382 + act);
    const double forced_c = surf ? V.forced_c[g] : 0.;
    const double eps_f = surf ? V.eps_f[g] : 0., eps_b = surf ? V.eps_b[g] : 0.;
    const double amb_f = surf ? V.amb_f[g] : 0., amb_b = surf ? V.amb_b[g] : 0.;
    const double ov_f = surf ? V.hs_ov_f[g] : CUDART_NAN, ov_b = surf ? V.hs_ov_b[g] : CUDART_NAN;
    const bool any_outdoor = kind_f == HB_BOUNDARY_OUTDOOR || kind_b == HB_BOUNDARY_OUTDOOR;
    const bool tilted = act < 0.98;  // surface.rs:38
    const double nx = surf ? V.nx[g] : 0., ny = surf ? V.ny[g] : 0.;
    const int site = surf ? V.site[g] : 0;
    const double* wtab = P.weather + (int64_t)site * P.weather_rows * 3;

    const double* slab_mass = V.mass + tile.node_offset + t;
    const double* slab_u = V.u + tile.node_offset + t;
    double* slab_T = V.nodeT + tile.node_offset + t;

    // ---- node state in registers ----
    double T[NMAX], A[NMAX], B[NMAX];
    double af0 = 0., abL = 0.;      // solar absorptance of the two end nodes
    double v0 = 0., vL = 0.;        // dt/C of the two end nodes (0 when no-mass)
#pragma unroll(LOCAL ? 1 : NMAX)
    for (int i = 0; i < (LOCAL ? n : NMAX); ++i) {
        T[i] = 0.;
        A[i] = 0.;
        B[i] = 0.;
        if (i < n) {
            T[i] = slab_T[(int64_t)i * kBlock];
            const double m = slab_mass[(int64_t)i * kBlock];
            const double vi = m >= kMassiveThreshold ? dt / m : 0.0;
            const double ui = slab_u[(int64_t)i * kBlock];                          // link i → i+1 (−1: cavity)
            const double um = i > 0 ? slab_u[(int64_t)(i - 1) * kBlock] : 0.0;      // link i−1 → i
            A[i] = (ui > 0. ? ui : 0.) * vi;
            B[i] = (um > 0. ? um : 0.) * vi;
            if (i == 0) v0 = vi;
            if (i == n - 1) vL = vi;
        }
    }
    if (surf) {
        af0 = V.alpha_f[tile.node_offset + t];
        abL = V.alpha_b[tile.node_offset + (int64_t)(n - 1) * kBlock + t];
    }
    const int nm_begin = surf ? V.nm_off[g] : 0, nm_end = surf ? V.nm_off[g + 1] : 0;
    const int cav_begin = surf ? V.cav_off[g] : 0, cav_end = surf ? V.cav_off[g + 1] : 0;

    const int64_t NS = V.n_dev_surfaces;
    double h_f = surf ? V.sc[0 * NS + g] : 0., h_b = surf ? V.sc[1 * NS + g] : 0.;
    double flow_f = surf ? V.sc[2 * NS + g] : 0., flow_b = surf ? V.sc[3 * NS + g] : 0.;
    double sol_f = surf ? V.sc[4 * NS + g] : 0., sol_b = surf ? V.sc[5 * NS + g] : 0.;
    double ir_f = surf ? V.sc[6 * NS + g] : 0., ir_b = surf ? V.sc[7 * NS + g] : 0.;
    double e_f = 0., e_b = 0.;
    unsigned long long nm_iters = 0;
    int status = 0;

    // ---- zones of the tile ----
    s_zf[t] = zf;
    s_zb[t] = zb;
    double z_vol = 0., z_q = 0., z_inf_v = 0., z_inf_t = CUDART_NAN, z_vent_v = 0., z_vent_t = CUDART_NAN, z_dt = 0.;
    int z_S = 0, z_sb = 0, z_se = 0;
    if (zthread) {
        s_zT[t] = V.zoneT[gz];
        z_vol = V.zvol[gz];
        z_q = V.zq[gz];
        z_inf_v = V.zinf_v[gz];
        z_inf_t = V.zinf_t[gz];
        z_vent_v = V.zvent_v[gz];
        z_vent_t = V.zvent_t[gz];
        z_dt = V.zdt[gz];
        z_S = V.zS[gz];
        z_sb = V.z_surf_begin[gz];
        z_se = V.z_surf_end[gz];
    }
    __syncthreads();

    int rec = 0;
    for (int step = 0; step < P.n_steps; ++step) {
        // ---- per-main-step inputs (what the caller writes into SimulationState before `march`) ----
        if (surf) {
            if (P.input_mode == HB_INPUT_SITE) {
                const double* in = P.inputs + ((int64_t)site * P.n_steps + step) * 4;
                sol_f = in[0]; sol_b = in[1]; ir_f = in[2]; ir_b = in[3];
            } else if (P.input_mode == HB_INPUT_PER_SURFACE) {
                const double* in = P.inputs + ((int64_t)step * P.n_desc_surfaces + V.desc_index[g]) * 4;
                sol_f = in[0]; sol_b = in[1]; ir_f = in[2]; ir_b = in[3];
            }
            if (P.input_mode != HB_INPUT_STATE && P.ir_mode == HB_IR_NET_GAIN) {
                // validate_wall_heat_transfer.rs:689-699 (front face only, as in the reference driver)
                if (eps_f > 1e-3) {
                    const double tk = T[0] + 273.15;
                    const double tk2 = tk * tk;
                    ir_f = ir_f / area / eps_f + kSigma * (tk2 * tk2);
                } else {
                    ir_f = V.sc[6 * NS + g];
                }
            }
        }
        // quantities that are constant over the main step
        double solar_front = sol_f;
        if (isnan(solar_front) || solar_front < 0.0) solar_front = 0.0;  // surface.rs:916-919
        double solar_back = sol_b;
        if (isnan(solar_back) || sol_f < 0.0) solar_back = 0.0;          // surface.rs:920-923 (tests solar_front)
        const double trad_out_f = sqrt(sqrt(ir_f / kSigma)) - 273.15;   // surface.rs:647
        const double trad_out_b = sqrt(sqrt(ir_b / kSigma)) - 273.15;   // surface.rs:692
        const double q_sol0 = af0 * solar_front;   // absorbed solar on the two end nodes (surface.rs:930-931)
        const double q_solL = abL * solar_back;

        const double* w0 = wtab + (int64_t)step * 3;
        if (P.weather_mode == HB_WEATHER_PER_SUBSTEP) w0 = wtab + (int64_t)step * S * 3;

        for (int s = 0; s < tile.smax; ++s) {
            if (surf && s < S) {
                // ---- weather sample at the END of the sub-step (model.rs:371-382) ----
                double t_out, wdir_deg, wspeed;
                if (P.weather_mode == HB_WEATHER_LINEAR) {
                    const double f = (double)(s + 1) / (double)S;
                    t_out = w0[0] + (w0[3] - w0[0]) * f;
                    wdir_deg = w0[1] + (w0[4] - w0[1]) * f;
                    wspeed = w0[2] + (w0[5] - w0[2]) * f;
                } else if (P.weather_mode == HB_WEATHER_PER_SUBSTEP) {
                    t_out = w0[3 * s + 0]; wdir_deg = w0[3 * s + 1]; wspeed = w0[3 * s + 2];
                } else {
                    t_out = w0[0]; wdir_deg = w0[1]; wspeed = w0[2];
                }
                double forced = 0.;
                if (any_outdoor) {
                    bool windward = true;  // surface.rs:37-46
                    if (tilted) {
                        double sw, cw;
                        sincos(wdir_deg * (PI / 180.0), &sw, &cw);
                        windward = nx * sw + ny * cw > 0.0;
                    }
                    forced = forced_c * (windward ? 1.0 : 0.5) * sqrt(wspeed);
                }
                // ---- boundary temperatures (model.rs:79-96) ----
                const double t_front = kind_f == HB_BOUNDARY_SPACE ? s_zT[zf] : kind_f == HB_BOUNDARY_AMBIENT ? amb_f : t_out;
                const double t_back = kind_b == HB_BOUNDARY_SPACE ? s_zT[zb] : kind_b == HB_BOUNDARY_AMBIENT ? amb_b : t_out;

                // ---- border conditions with the OLD surface temperatures (surface.rs:939-948) ----
                const double ts_f_old = T[0];
                const double ts_b_old = reg_get<NMAX, LOCAL>(T, n - 1);
                Face F, Bk;
                auto border = [&](double ts_f, double ts_b) {
                    if (kind_f == HB_BOUNDARY_OUTDOOR) {
                        F.air = t_front; F.trad = trad_out_f; F.ts_env = ts_f;
                        F.h = forced + tarp_natural(t_front, ts_f, -cos_tilt, c_same, c_opp);  // tilt flipped, :652
                    } else {
                        F.air = kind_f == HB_BOUNDARY_SPACE ? t_front : amb_f; F.trad = t_front; F.ts_env = ts_f;
                        F.h = tarp_natural(F.air, ts_f, cos_tilt, c_same, c_opp);
                    }
                    if (kind_b == HB_BOUNDARY_OUTDOOR) {
                        Bk.air = t_back; Bk.trad = trad_out_b; Bk.ts_env = ts_b;
                        Bk.h = forced + tarp_natural(t_back, ts_b, cos_tilt, c_same, c_opp);
                    } else if (kind_b == HB_BOUNDARY_SPACE) {
                        Bk.air = t_back; Bk.trad = t_back; Bk.ts_env = ts_b;
                        Bk.h = tarp_natural(t_back, ts_b, cos_tilt, c_same, c_opp);
                    } else {
                        // surface.rs:672-685: the back AmbientTemperature arm is built from FRONT values
                        Bk.air = amb_b; Bk.trad = t_front; Bk.ts_env = ts_f;
                        Bk.h = tarp_natural(amb_b, ts_f, cos_tilt, c_same, c_opp);
                    }
                    if (isnan(F.h) || isnan(Bk.h)) status |= ST_NAN_H;  // surface.rs:704-707 (panic in the reference)
                    if (!isnan(ov_f)) F.h = ov_f;                        // surface.rs:708-714
                    if (!isnan(ov_b)) Bk.h = ov_b;
                };
                border(ts_f_old, ts_b_old);
                const double hf = F.h, hb_ = Bk.h;
                double xr_f = 273.15 + (F.trad + F.ts_env) / 2.;
                double xr_b = 273.15 + (Bk.trad + Bk.ts_env) / 2.;
                const double rad_f = 4. * eps_f * kSigma * ((xr_f * xr_f) * xr_f);
                const double rad_b = 4. * eps_b * kSigma * ((xr_b * xr_b) * xr_b);

                // ---- no-mass chunks (surface.rs:950-965 → 790-898) ----
                for (int c = nm_begin; c < nm_end; ++c) {
                    const int ini = V.nm_pairs[2 * c], fin = V.nm_pairs[2 * c + 1];
                    const int m = fin - ini;
                    double tl[kMaxNomassChunk + 2];   // T[ini-1 .. fin]
                    double ul[kMaxNomassChunk + 1];   // links ini-1 .. fin-1
                    double sl[kMaxNomassChunk];       // absorbed solar per node
                    int cavl[kMaxNomassChunk + 1];
                    for (int j = 0; j < m + 2; ++j) {
                        const int node = ini - 1 + j;
                        tl[j] = (node >= 0 && node < n) ? reg_get<NMAX, LOCAL>(T, node) : 0.;
                    }
                    for (int j = 0; j < m + 1; ++j) {
                        const int link = ini - 1 + j;
                        ul[j] = (link >= 0 && link < n - 1) ? slab_u[(int64_t)link * kBlock] : 0.;
                        cavl[j] = -1;
                        if (ul[j] < 0.)
                            for (int q = cav_begin; q < cav_end; ++q)
                                if (V.cav[q].node == link) cavl[j] = q;
                    }
                    for (int j = 0; j < m; ++j) {
                        const int node = ini + j;
                        if constexpr (DIST) {
                            sl[j] = V.alpha_f[tile.node_offset + (int64_t)node * kBlock + t] * solar_front +
                                    V.alpha_b[tile.node_offset + (int64_t)node * kBlock + t] * solar_back;
                        } else {
                            sl[j] = (node == 0 ? af0 * solar_front : 0.) + (node == n - 1 ? abL * solar_back : 0.);
                        }
                    }
                    double old_err = 99999.;
                    int count = 0;
                    while (true) {
                        ++nm_iters;
                        // assemble (discretization.rs:622-697)
                        double dl[kMaxNomassChunk], dd[kMaxNomassChunk], du[kMaxNomassChunk], rhs[kMaxNomassChunk];
                        for (int j = 0; j < m; ++j) { dl[j] = 0.; dd[j] = 0.; du[j] = 0.; rhs[j] = 0.; }
                        for (int j = 0; j + 1 < m; ++j) {
                            double uu = ul[j + 1];
                            if (uu < 0.) uu = cavity_u(V.cav[cavl[j + 1]], tl[j + 1], tl[j + 2], status);
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
                            if (uu < 0.) uu = cavity_u(V.cav[cavl[0]], tl[0], tl[1], status);
                            hs_front = uu;
                            front_q = uu * tl[0];
                        }
                        rhs[0] += front_q;
                        dd[0] += -hs_front;
                        double hs_back, back_q;
                        if (fin == n) {
                            back_q = Bk.air * hb_ + rad_b * (Bk.trad - tl[m]);
                            hs_back = hb_;
                        } else {
                            double uu = ul[m];
                            if (uu < 0.) uu = cavity_u(V.cav[cavl[m]], tl[m], tl[m + 1], status);
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
                        if (err > old_err) break;                       // surface.rs:842-848
                        if (isnan(err)) { status |= ST_NAN_ERR; break; }  // reference panics, :850
                        for (int j = 0; j < m; ++j) tl[j + 1] = (tl[j + 1] + rhs[j]) * 0.5;  // :878-883
                        const double tol = count < 100 ? 0.01 : 0.5;   // :885
                        if (err / (double)m < tol) break;
                        old_err = err;
                        ++count;
                        if (count > 100000) { status |= ST_ITER_CAP; break; }
                    }
                    for (int j = 0; j < m; ++j) reg_set<NMAX, LOCAL>(T, ini + j, tl[j + 1]);
                }

                // ---- cavity links between massive nodes: conductance at the assembly temperatures ----
                if (has_cav) {
                    for (int q = cav_begin; q < cav_end; ++q) {
                        const CavityDev cv = V.cav[q];
                        const int i = cv.node;
                        const double uu = cavity_u(cv, reg_get<NMAX, LOCAL>(T, i), reg_get<NMAX, LOCAL>(T, i + 1), status);
                        const double mi = slab_mass[(int64_t)i * kBlock], mj = slab_mass[(int64_t)(i + 1) * kBlock];
                        reg_set<NMAX, LOCAL>(A, i, uu * (mi >= kMassiveThreshold ? dt / mi : 0.));
                        reg_set<NMAX, LOCAL>(B, i + 1, uu * (mj >= kMassiveThreshold ? dt / mj : 0.));
                    }
                }

                // ---- massive nodes: RK4 (surface.rs:984-1000 → 720-787, 228-308) ----
                {
                    // film coefficients as links to virtual nodes at the equivalent air temperature
                    const double t0_asm = T[0];
                    const double tL_asm = reg_get<NMAX, LOCAL>(T, n - 1);
                    const double hf_s = hf > 1e-300 ? hf : 1e-300, hb_s = hb_ > 1e-300 ? hb_ : 1e-300;
                    const double xf = F.air + (rad_f * (F.trad - t0_asm) + q_sol0) / hf_s;
                    const double xb = Bk.air + (rad_b * (Bk.trad - tL_asm) + q_solL) / hb_s;
                    const double b0 = hf_s * v0;
                    const double aL = hb_s * vL;
                    double X[NMAX], ACC[NMAX];
                    double Q[DIST ? NMAX : 1];
#pragma unroll(LOCAL ? 1 : NMAX)
                    for (int i = 0; i < (LOCAL ? n : NMAX); ++i) {
                        X[i] = T[i];
                        ACC[i] = T[i];
                        if constexpr (DIST) {
                            Q[i] = 0.;
                            if (i > 0 && i < n - 1) {
                                const double m_i = slab_mass[(int64_t)i * kBlock];
                                const double v_i = m_i >= kMassiveThreshold ? dt / m_i : 0.;
                                Q[i] = v_i * (V.alpha_f[tile.node_offset + (int64_t)i * kBlock + t] * solar_front +
                                              V.alpha_b[tile.node_offset + (int64_t)i * kBlock + t] * solar_back);
                            } else if (i < n) {
                                // end nodes: the part of the distributed absorptance not already in q_sol0/q_solL
                                const double extra = (i == 0 ? V.alpha_b[tile.node_offset + t] * solar_back : 0.) +
                                                     (i == n - 1 && n > 1
                                                          ? V.alpha_f[tile.node_offset + (int64_t)i * kBlock + t] * solar_front
                                                          : 0.);
                                Q[i] = (i == 0 ? v0 : vL) * extra;
                            }
                        }
                    }
                    auto stage = [&](const double cx, const double w, const bool last) {
                        double dleft = X[0] - xf;
#pragma unroll(LOCAL ? 1 : NMAX)
                        for (int i = 0; i < (LOCAL ? n : NMAX); ++i) {
                            if (i < n) {
                                const bool end = i + 1 >= n;
                                const double xr = end ? xb : X[i + 1 < NMAX ? i + 1 : i];
                                const double ai = end ? aL : A[i];
                                const double bi = i == 0 ? b0 : B[i];
                                const double dright = xr - X[i];
                                double k = ai * dright - bi * dleft;
                                if constexpr (DIST) k += Q[i];
                                dleft = dright;
                                ACC[i] += w * k;
                                if (!last) X[i] = T[i] + cx * k;
                            }
                        }
                    };
                    stage(0.5, 1. / 6., false);
                    stage(0.5, 1. / 3., false);
                    stage(1.0, 1. / 3., false);
                    stage(0.0, 1. / 6., true);
#pragma unroll(LOCAL ? 1 : NMAX)
                    for (int i = 0; i < (LOCAL ? n : NMAX); ++i) T[i] = ACC[i];
                }

                // ---- new film coefficients and convective flows (model.rs:145-169) ----
                const double ts_f = T[0];
                const double ts_b = reg_get<NMAX, LOCAL>(T, n - 1);
                border(ts_f, ts_b);
                h_f = F.h;
                h_b = Bk.h;
                flow_f = (ts_f - t_front) * h_f;
                flow_b = (ts_b - t_back) * h_b;
                e_f += flow_f * area * dt;
                e_b += flow_b * area * dt;
                // contributions to calculate_zones_abc (model.rs:556-590)
                s_pub[0][t] = h_f * area;
                s_pub[1][t] = h_f * area * ts_f;
                s_pub[2][t] = h_b * area;
                s_pub[3][t] = h_b * area * ts_b;
            }
            __syncthreads();
            // ---- zones: a, b, c and the analytic update (model.rs:412-423, 489-597, 650-674) ----
            if (zthread && s < z_S) {
                const double tz = s_zT[t];
                double a = 0.0, b = 0.0;
                a += z_q;
                if (!isnan(z_inf_t)) {
                    const double tk = z_inf_t + 273.15;
                    const double cp = 1002.7370 + 1.2324e-2 * tk;
                    const double rho = 101325. * 28.97 / (8314.46261815324 * tk);
                    a += rho * z_inf_v * cp * z_inf_t;
                    b += rho * z_inf_v * cp;
                }
                if (!isnan(z_vent_t)) {
                    const double tk = z_vent_t + 273.15;
                    const double cp = 1002.7370 + 1.2324e-2 * tk;
                    const double rho = 101325. * 28.97 / (8314.46261815324 * tk);
                    a += rho * z_vent_v * cp * z_vent_t;
                    b += rho * z_vent_v * cp;
                }
                const double tk = tz + 273.15;
                const double c = z_vol * (101325. * 28.97 / (8314.46261815324 * tk)) * (1002.7370 + 1.2324e-2 * tk) / 1.;
                for (int j = z_sb; j < z_se; ++j) {   // surfaces then fenestrations, front face before back face
                    if (s_zf[j] == t) { a += s_pub[1][j]; b += s_pub[0][j]; }
                    if (s_zb[j] == t) { a += s_pub[3][j]; b += s_pub[2][j]; }
                }
                double tn = tz;
                if (fabs(b) > 1e-9) tn = a / b + (tz - a / b) * exp(-b * z_dt / c);
                if (isnan(tn)) atomicOr(&V.status[(int64_t)tile.surface_begin + z_sb], ST_NAN_ZONE);
                s_zT[t] = tn;
            }
            __syncthreads();
        }

        if (P.record_every > 0 && (step + 1) % P.record_every == 0) {
            if (P.zone_rec && zthread) P.zone_rec[(int64_t)rec * V.n_dev_zones + gz] = s_zT[t];
            if (P.node_rec && surf) {
#pragma unroll(LOCAL ? 1 : NMAX)
                for (int i = 0; i < (LOCAL ? n : NMAX); ++i)
                    if (i < n) P.node_rec[(int64_t)rec * P.slab_len + tile.node_offset + (int64_t)i * kBlock + t] = T[i];
                double* r = P.sc_rec + (int64_t)rec * 8 * NS;
                r[0 * NS + g] = h_f; r[1 * NS + g] = h_b; r[2 * NS + g] = flow_f; r[3 * NS + g] = flow_b;
                r[4 * NS + g] = sol_f; r[5 * NS + g] = sol_b; r[6 * NS + g] = ir_f; r[7 * NS + g] = ir_b;
            }
            ++rec;
        }
    }

    // ---- write the state back ----
    if (surf) {
#pragma unroll(LOCAL ? 1 : NMAX)
        for (int i = 0; i < (LOCAL ? n : NMAX); ++i)
            if (i < n) slab_T[(int64_t)i * kBlock] = T[i];
        V.sc[0 * NS + g] = h_f; V.sc[1 * NS + g] = h_b; V.sc[2 * NS + g] = flow_f; V.sc[3 * NS + g] = flow_b;
        V.sc[4 * NS + g] = sol_f; V.sc[5 * NS + g] = sol_b; V.sc[6 * NS + g] = ir_f; V.sc[7 * NS + g] = ir_b;
        V.energy_f[g] += e_f;
        V.energy_b[g] += e_b;
        V.nm_iters[g] += nm_iters;
        if (status) atomicOr(&V.status[g], status);
    }
    if (zthread) V.zoneT[gz] = s_zT[t];
}

#ifdef HB_AUX_KERNELS
// ----------------------------------------------------------------------------------------------------
// K3: ensemble reduction. out[0..8): Σzone, min zone, max zone, Σnode, min node, max node, E_f, E_b;
// iout[0..4): n_zones, n_nodes, no-mass iterations, status OR.
// ----------------------------------------------------------------------------------------------------
__device__ __forceinline__ double atomic_min_f64(double* addr, double v) {
    unsigned long long* p = (unsigned long long*)addr;
    unsigned long long old = *p, assumed;
    do {
        assumed = old;
        if (__longlong_as_double(assumed) <= v) break;
        old = atomicCAS(p, assumed, __double_as_longlong(v));
    } while (assumed != old);
    return __longlong_as_double(old);
}
__device__ __forceinline__ double atomic_max_f64(double* addr, double v) {
    unsigned long long* p = (unsigned long long*)addr;
    unsigned long long old = *p, assumed;
    do {
        assumed = old;
        if (__longlong_as_double(assumed) >= v) break;
        old = atomicCAS(p, assumed, __double_as_longlong(v));
    } while (assumed != old);
    return __longlong_as_double(old);
}

__global__ void __launch_bounds__(kBlock) hb_reduce_kernel(DevView V, int n_tiles, double* out,
                                                           unsigned long long* iout) {
    double sz = 0., mnz = CUDART_INF, mxz = -CUDART_INF, sn = 0., mnn = CUDART_INF, mxn = -CUDART_INF, ef = 0., eb = 0.;
    unsigned long long nz = 0, nn = 0, it = 0, st = 0;
    const int t = threadIdx.x;
    for (int ti = blockIdx.x; ti < n_tiles; ti += gridDim.x) {
        const Tile tile = V.tiles[ti];
        const int64_t g = (int64_t)tile.surface_begin + t;
        const int n = V.n_nodes[g];
        for (int i = 0; i < n; ++i) {
            const double v = V.nodeT[tile.node_offset + (int64_t)i * kBlock + t];
            sn += v;
            mnn = fmin(mnn, v);
            mxn = fmax(mxn, v);
        }
        if (n > 0) {
            nn += n;
            ef += V.energy_f[g];
            eb += V.energy_b[g];
            it += V.nm_iters[g];
            st |= (unsigned long long)V.status[g];
        }
        if (t < tile.n_zones) {
            const double v = V.zoneT[(int64_t)tile.zone_begin + t];
            sz += v;
            mnz = fmin(mnz, v);
            mxz = fmax(mxz, v);
            nz += 1;
        }
    }
    // warp reduce, then one atomic per warp
    for (int o = 16; o > 0; o >>= 1) {
        sz += __shfl_down_sync(0xffffffffu, sz, o);
        sn += __shfl_down_sync(0xffffffffu, sn, o);
        ef += __shfl_down_sync(0xffffffffu, ef, o);
        eb += __shfl_down_sync(0xffffffffu, eb, o);
        mnz = fmin(mnz, __shfl_down_sync(0xffffffffu, mnz, o));
        mxz = fmax(mxz, __shfl_down_sync(0xffffffffu, mxz, o));
        mnn = fmin(mnn, __shfl_down_sync(0xffffffffu, mnn, o));
        mxn = fmax(mxn, __shfl_down_sync(0xffffffffu, mxn, o));
        nz += __shfl_down_sync(0xffffffffu, nz, o);
        nn += __shfl_down_sync(0xffffffffu, nn, o);
        it += __shfl_down_sync(0xffffffffu, it, o);
        st |= __shfl_down_sync(0xffffffffu, st, o);
    }
    if ((t & 31) == 0) {
        atomicAdd(&out[0], sz);
        atomic_min_f64(&out[1], mnz);
        atomic_max_f64(&out[2], mxz);
        atomicAdd(&out[3], sn);
        atomic_min_f64(&out[4], mnn);
        atomic_max_f64(&out[5], mxn);
        atomicAdd(&out[6], ef);
        atomicAdd(&out[7], eb);
        atomicAdd(&iout[0], nz);
        atomicAdd(&iout[1], nn);
        atomicAdd(&iout[2], it);
        atomicOr(&iout[3], st);
    }
}

// ----------------------------------------------------------------------------------------------------
// probes
// ----------------------------------------------------------------------------------------------------
__global__ void hb_dfma_probe_kernel(double* out, int iters, double seed) {
    double a0 = seed + threadIdx.x, a1 = a0 + 1., a2 = a0 + 2., a3 = a0 + 3., a4 = a0 + 4., a5 = a0 + 5., a6 = a0 + 6.,
           a7 = a0 + 7.;
    const double m = 0.999999, c = 1e-9;
    for (int i = 0; i < iters; ++i) {
        a0 = fma(a0, m, c); a1 = fma(a1, m, c); a2 = fma(a2, m, c); a3 = fma(a3, m, c);
        a4 = fma(a4, m, c); a5 = fma(a5, m, c); a6 = fma(a6, m, c); a7 = fma(a7, m, c);
    }
    out[(int64_t)blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
}

__global__ void hb_fill_kernel(double* p, int64_t n, double v) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) p[i] = v;
}

#endif  // HB_AUX_KERNELS

}  // namespace hb

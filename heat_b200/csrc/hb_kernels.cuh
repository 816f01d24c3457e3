// hb_kernels.cuh — hand-written CUDA (sm_100a) for the wall-conduction time-stepping path.
//
// K1/K2  hb_march_kernel<G, W, VAR>: persistent over all main steps and sub-steps of a series.
//        One CUDA block = one tile of W walls belonging to whole models; the tile is uniform in sub-step count.
//        Two roles share the block:
//          * NODE lanes — every wall is owned by a group of G consecutive lanes, each lane keeping kNL slots
//            (nodes + the two ghost nodes that stand for the air on either side) of temperature and of the scaled
//            tridiagonal coefficients in REGISTERS for the whole series.  A sub-step's RK4 (surface.rs:228-308) is
//            four tridiagonal sweeps with one warp-shuffle halo exchange each (hb_math.cuh).
//          * SCALAR lanes — lane j of one warp of the block evaluates wall j's per-sub-step scalar physics once:
//            boundary temperatures (model.rs:79-96), border conditions / TARP film coefficients
//            (surface.rs:596-717, convection.rs:87-168), radiative coefficients (surface.rs:941-948), no-mass
//            chunks and cavity links (surface.rs:790-898, cavity.rs, gas.rs), then — after the nodes marched —
//            the new film coefficients, convective flows (model.rs:145-169) and the zone a/b/c sums and analytic
//            update (model.rs:489-597, 650-674).
//        The roles meet twice per sub-step at a block barrier and exchange 4 + 2 doubles per wall through shared
//        memory.  Per-sub-step weather (interpolation, sin/cos of the wind direction, sqrt of the wind speed) is
//        computed once per block and main step into a shared table.
// K3     hb_reduce_kernel: ensemble sums / extrema feeding hb_reduce() (and NCCL in the host layer).
#pragma once
#include <cuda_runtime.h>
#include <math_constants.h>

#include "hb_math.cuh"

namespace hb {

constexpr int kVarPlain = 0, kVarExt = 1, kVarDist = 2, kVarEnd = 3, kVarGen = 4, kVarPane = 5, kVarPlainSimple = 6,
              kVarEndSimple = 7;
// hb_face_kernel modes: lean (all ends massive), lean with no-mass end nodes, general (everything), lean with
// no-mass end nodes and single panes
constexpr int kFaceLean = 0, kFaceLeanEnd = 1, kFaceGen = 2, kFacePane = 3;

#ifndef HB_MARCH_OCC
// resident threads per SM the general kernel is compiled for.  384 (168 registers per thread, spills halved) measured
// 5.5 % faster than 512 (128 registers) on double-glazed buildings, 256 (no spills) the same as 384, 768 (80
// registers) 14 % slower (profiles/r1x_occ_ab.log, r1x_occ_ab2.log): the scalar lane's chain sets the time, not occupancy
#define HB_MARCH_OCC 384
#endif
template <int G, int W, int VAR>
__global__ void __launch_bounds__(G* W, HB_MARCH_OCC / (G * W) > 0 ? HB_MARCH_OCC / (G * W) : 1) hb_march_kernel(DevView V, SeriesDev P, const int32_t* tile_ids) {
    constexpr int TPB = G * W;
    constexpr int NL = kNL;
    constexpr bool EXT = VAR >= kVarExt, DIST = VAR == kVarDist;
    constexpr int NWARP = TPB / 32 > 0 ? TPB / 32 : 1;
    constexpr int ROW = G * NL + 1;   // padded row of the shared-memory mirrors
    const Tile tile = V.tiles[tile_ids[blockIdx.x]];
    const int tid = threadIdx.x;
    const int64_t NS = V.n_dev_surfaces;

    __shared__ WallC s_wc[W];
    __shared__ ZoneC s_zc[W];
    __shared__ double s_zT[W];
    __shared__ double s_pub[4][W];
    __shared__ int32_t s_zf[W], s_zb[W];
    __shared__ BoundaryLinks s_bc[W];
    __shared__ double s_ts[2][W];
    __shared__ double s_sol[2][W];
    __shared__ SubWeather s_w[kMaxSubsteps];
    extern __shared__ double s_dyn[];   // EXT: temperature / coefficient mirrors [3][W][ROW]
    double* const sT = s_dyn;
    double* const sA = s_dyn + W * ROW;
    double* const sB = s_dyn + 2 * W * ROW;

    const int S = tile.S;
    const double dt = tile.dt;
    const bool has_cav = (tile.flags & TILE_CAV) != 0;

    // ---- node role ----
    const int wl = tid / G, part = tid % G;
    const int64_t g = (int64_t)tile.surface_begin + wl;
    const int n = V.n_nodes[g];
    const int iN = n - part * NL;        // lane-local index of the last node's slot (slot n)
    const int iG = n + 1 - part * NL;    // … of the back ghost (slot n+1)
    const int row = wl * ROW + part;     // mirror address of lane-local slot i: row + i*G
    double T[NL], A[NL], B[NL], Vv[NL];
    double Q[DIST ? NL : 1];
    {
        const int64_t o = tile.node_offset + tid;
#pragma unroll
        for (int i = 0; i < NL; ++i) {
            T[i] = V.nodeT[o + (int64_t)i * TPB];
            A[i] = V.coefA[o + (int64_t)i * TPB];
            B[i] = V.coefB[o + (int64_t)i * TPB];
            if constexpr (EXT) {
                sT[row + i * G] = T[i];
                sA[row + i * G] = A[i];
                sB[row + i * G] = B[i];
            }
        }
    }
    const bool wall_dist = DIST && ((V.flags[g] >> 8) & 1);

    // ---- scalar role ----
    const int swarp = NWARP > 1 ? (int)(blockIdx.x % NWARP) : 0;
    const int stid = tid - 32 * swarp;
    const bool scalar = stid >= 0 && stid < W && stid < 32;
    if (tid < W) {
        const int64_t sg = (int64_t)tile.surface_begin + tid;
        WallC c;
        load_wallc(V, sg, P.single_site != 0, c);
        s_wc[tid] = c;
        s_zf[tid] = c.zf;
        s_zb[tid] = c.zb;
        s_bc[tid] = BoundaryLinks{0., 0., 0., 0.};
        const int64_t gz = (int64_t)tile.zone_begin + tid;
        ZoneC z{};
        if (tid < tile.n_zones) {
            z.vol = V.zvol[gz]; z.q = V.zq[gz]; z.inf_v = V.zinf_v[gz]; z.inf_t = V.zinf_t[gz];
            z.vent_v = V.zvent_v[gz]; z.vent_t = V.zvent_t[gz]; z.sb = V.z_surf_begin[gz]; z.se = V.z_surf_end[gz];
            s_zT[tid] = V.zoneT[gz];
        }
        s_zc[tid] = z;
    }
    auto publish_ends = [&]() {
        if (part == 0) s_ts[0][wl] = T[1];
#pragma unroll
        for (int i = 0; i < NL; ++i)
            if (i == iN) s_ts[1][wl] = T[i];
    };
    publish_ends();
    __syncthreads();

    const int64_t sg = (int64_t)tile.surface_begin + (scalar ? stid : 0);
    const bool swall = scalar && s_wc[scalar ? stid : 0].n > 0;
    const bool szone = scalar && stid < tile.n_zones;
    const int64_t sgz = (int64_t)tile.zone_begin + (scalar ? stid : 0);
    WallS w{};
    ExtWall X{};
    if (swall) {
        w.h_f = V.sc[0 * NS + sg]; w.h_b = V.sc[1 * NS + sg]; w.flow_f = V.sc[2 * NS + sg]; w.flow_b = V.sc[3 * NS + sg];
        w.sol_f = V.sc[4 * NS + sg]; w.sol_b = V.sc[5 * NS + sg]; w.ir_f = V.sc[6 * NS + sg]; w.ir_b = V.sc[7 * NS + sg];
        w.ts_f = s_ts[0][stid];
        w.ts_b = s_ts[1][stid];
        if constexpr (EXT) {
            X.t = sT + stid * ROW; X.a = sA + stid * ROW; X.b = sB + stid * ROW;
            X.u = V.ext_u + tile.ext_offset + stid; X.mass = V.ext_mass + tile.ext_offset + stid;
            X.af = V.ext_af + tile.ext_offset + stid; X.ab = V.ext_ab + tile.ext_offset + stid;
            X.G = G; X.W = W;
        }
    }
    const double ir_f_state = w.ir_f;

    int rec = 0;
    for (int step = 0; step < P.n_steps; ++step) {
        // ---- main-step prologue: weather table, per-step inputs (what the caller writes into SimulationState) ----
        if (P.zone_gain || P.zone_inf_v || P.zone_inf_t || P.zone_vent_v || P.zone_vent_t) {
            // scheduled zone inputs of this main step (read by the zone lanes after the syncs below)
            __syncthreads();
            if (tid < tile.n_zones) {
                ZoneC z = s_zc[tid];
                load_zone_schedule(P, step, (int64_t)tile.zone_begin + tid, z);
                s_zc[tid] = z;
            }
            __syncthreads();
        }
        if (P.single_site) {
            const double* w0 = P.weather + (int64_t)step * 3;
            if (P.weather_mode == HB_WEATHER_PER_SUBSTEP) w0 = P.weather + (int64_t)step * S * 3;
            for (int s = tid; s < S; s += TPB) {
                double t_out, wdir, wspeed;
                if (P.weather_mode == HB_WEATHER_LINEAR) {
                    const double f = (double)(s + 1) / (double)S;
                    t_out = lerp_rn(w0[0], w0[3], f);
                    wdir = lerp_rn(w0[1], w0[4], f);
                    wspeed = lerp_rn(w0[2], w0[5], f);
                } else if (P.weather_mode == HB_WEATHER_PER_SUBSTEP) {
                    t_out = w0[3 * s + 0]; wdir = w0[3 * s + 1]; wspeed = w0[3 * s + 2];
                } else {
                    t_out = w0[0]; wdir = w0[1]; wspeed = w0[2];
                }
                sub_weather(t_out, wdir, wspeed, s_w[s]);
            }
        }
        if (swall) {
            const WallC& c = s_wc[stid];
            if (P.input_mode == HB_INPUT_SITE) {
                const double* in = P.inputs + ((int64_t)c.site * P.n_steps + step) * 4;
                w.sol_f = in[0]; w.sol_b = in[1]; w.ir_f = in[2]; w.ir_b = in[3];
            } else if (P.input_mode == HB_INPUT_PER_SURFACE) {
                const double* in = P.inputs + ((int64_t)step * P.n_desc_surfaces + c.desc_index) * 4;
                w.sol_f = in[0]; w.sol_b = in[1]; w.ir_f = in[2]; w.ir_b = in[3];
            }
            if (P.input_mode != HB_INPUT_STATE && P.ir_mode == HB_IR_NET_GAIN) {
                // validate_wall_heat_transfer.rs:689-699 (front face only, as in the reference driver)
                if (c.eps_f > 1e-3) {
                    const double tk = w.ts_f + 273.15;
                    const double tk2 = tk * tk;
                    w.ir_f = w.ir_f / c.area / c.eps_f + kSigma * (tk2 * tk2);
                } else {
                    w.ir_f = ir_f_state;
                }
            }
            begin_main_step(c, w);
            if constexpr (DIST) {
                s_sol[0][stid] = w.solar_front;
                s_sol[1][stid] = w.solar_back;
            }
        }
        __syncthreads();
        if constexpr (DIST) {
            // per-node absorbed solar of transparent constructions, scaled by dt/C (surface.rs:930-931, 766-769)
#pragma unroll
            for (int i = 0; i < NL; ++i) {
                Q[i] = 0.;
                const int node = part * NL + i - 1;
                if (wall_dist && node >= 0 && node < n) {
                    const int64_t e = tile.ext_offset + (int64_t)node * W + wl;
                    const double m = V.ext_mass[e];
                    const double vi = m >= kMassiveThreshold ? dt / m : 0.;
                    Q[i] = vi * (V.ext_af[e] * s_sol[0][wl] + V.ext_ab[e] * s_sol[1][wl]);
                }
            }
        }

        for (int s = 0; s < S; ++s) {
            // ================= scalar role, before the march =================
            if (swall) {
                const WallC& c = s_wc[stid];
                SubWeather sw;
                if (P.single_site) {
                    sw = s_w[s];
                } else {
                    const double* wtab = P.weather + (int64_t)c.site * P.weather_rows * 3;
                    const double* w0 = wtab + (int64_t)step * 3;
                    double t_out, wdir, wspeed;
                    if (P.weather_mode == HB_WEATHER_LINEAR) {
                        const double f = (double)(s + 1) / (double)S;
                        t_out = lerp_rn(w0[0], w0[3], f);
                        wdir = lerp_rn(w0[1], w0[4], f);
                        wspeed = lerp_rn(w0[2], w0[5], f);
                    } else if (P.weather_mode == HB_WEATHER_PER_SUBSTEP) {
                        w0 = wtab + ((int64_t)step * S + s) * 3;
                        t_out = w0[0]; wdir = w0[1]; wspeed = w0[2];
                    } else {
                        t_out = w0[0]; wdir = w0[1]; wspeed = w0[2];
                    }
                    sub_weather(t_out, wdir, wspeed, sw);
                }
                const double ztf = c.zf >= 0 ? s_zT[c.zf] : 0.;
                const double ztb = c.zb >= 0 ? s_zT[c.zb] : 0.;
                BoundaryLinks bl;
                scalar_pre<EXT>(c, w, sw, ztf, ztb, dt, X, V.cav, V.nm_pairs, bl);
                s_bc[stid] = bl;
            }
            __syncthreads();
            // ================= node role: RK4 of the massive nodes =================
            {
                if constexpr (EXT) {
#pragma unroll
                    for (int i = 0; i < NL; ++i) T[i] = sT[row + i * G];
                    if (has_cav) {
#pragma unroll
                        for (int i = 0; i < NL; ++i) {
                            A[i] = sA[row + i * G];
                            B[i] = sB[row + i * G];
                        }
                    }
                }
                const BoundaryLinks bl = s_bc[wl];
                if (part == 0) {
                    T[0] = bl.ghost_f;
                    B[1] = bl.bf;
                }
#pragma unroll
                for (int i = 0; i < NL; ++i) {
                    if (i == iN) A[i] = bl.ab;
                    if (i == iG) T[i] = bl.ghost_b;
                }
                double left = 0., right = 0.;
                if constexpr (G > 1) {
                    left = __shfl_up_sync(0xffffffffu, T[NL - 1], 1, G);
                    right = __shfl_down_sync(0xffffffffu, T[0], 1, G);
                }
                sweep_first<NL>(T, Vv, A, B, left, right);
#pragma unroll
                for (int i = 0; i < NL; ++i) {
                    if constexpr (DIST) Vv[i] += Q[i];
                    T[i] += Vv[i];
                }
                if constexpr (G > 1) {
                    left = __shfl_up_sync(0xffffffffu, Vv[NL - 1], 1, G);
                    right = __shfl_down_sync(0xffffffffu, Vv[0], 1, G);
                }
                sweep_next<NL>(T, Vv, A, B, left, right, 1. / 2.);
                if constexpr (G > 1) {
                    left = __shfl_up_sync(0xffffffffu, Vv[NL - 1], 1, G);
                    right = __shfl_down_sync(0xffffffffu, Vv[0], 1, G);
                }
                sweep_next<NL>(T, Vv, A, B, left, right, 1. / 6.);
                if constexpr (G > 1) {
                    left = __shfl_up_sync(0xffffffffu, Vv[NL - 1], 1, G);
                    right = __shfl_down_sync(0xffffffffu, Vv[0], 1, G);
                }
                sweep_next<NL>(T, Vv, A, B, left, right, 1. / 24.);
                if constexpr (EXT) {
#pragma unroll
                    for (int i = 0; i < NL; ++i) sT[row + i * G] = T[i];
                } else {
                    publish_ends();
                }
            }
            __syncthreads();
            // ================= scalar role, after the march =================
            if (tid >> 5 == swarp) {   // warp-uniform: the whole scalar warp takes the __syncwarp()s
                if (swall) {
                    const WallC& c = s_wc[stid];
                    double ts_f, ts_b;
                    if constexpr (EXT) {
                        ts_f = X.T(0);
                        ts_b = X.T(c.n - 1);
                    } else {
                        ts_f = s_ts[0][stid];
                        ts_b = s_ts[1][stid];
                    }
                    double pub[4];
                    scalar_post(c, w, ts_f, ts_b, dt, pub);
                    s_pub[0][stid] = pub[0]; s_pub[1][stid] = pub[1]; s_pub[2][stid] = pub[2]; s_pub[3][stid] = pub[3];
                }
                __syncwarp();
                if (szone) {
                    const double tn = zone_update(s_zc[stid], stid, s_zT[stid], dt, s_pub[0], s_pub[1], s_pub[2], s_pub[3],
                                                  s_zf, s_zb);
                    if (tn != tn) atomicOr(&V.status[(int64_t)tile.surface_begin + s_zc[stid].sb], ST_NAN_ZONE);
                    s_zT[stid] = tn;
                }
                __syncwarp();
            }
        }

        if (P.record_every > 0 && (step + 1) % P.record_every == 0) {
            if (P.zone_rec && szone) P.zone_rec[(int64_t)rec * V.n_dev_zones + sgz] = s_zT[stid];
            if (P.node_rec) {
#pragma unroll
                for (int i = 0; i < NL; ++i)
                    P.node_rec[(int64_t)rec * P.slab_len + tile.node_offset + (int64_t)i * TPB + tid] = T[i];
                if (swall) {
                    double* r = P.sc_rec + (int64_t)rec * 8 * NS;
                    r[0 * NS + sg] = w.h_f; r[1 * NS + sg] = w.h_b; r[2 * NS + sg] = w.flow_f; r[3 * NS + sg] = w.flow_b;
                    r[4 * NS + sg] = w.sol_f; r[5 * NS + sg] = w.sol_b; r[6 * NS + sg] = w.ir_f; r[7 * NS + sg] = w.ir_b;
                }
            }
            ++rec;
        }
    }

    // ---- write the state back ----
#pragma unroll
    for (int i = 0; i < NL; ++i) V.nodeT[tile.node_offset + (int64_t)i * TPB + tid] = T[i];
    if (swall) {
        V.sc[0 * NS + sg] = w.h_f; V.sc[1 * NS + sg] = w.h_b; V.sc[2 * NS + sg] = w.flow_f; V.sc[3 * NS + sg] = w.flow_b;
        V.sc[4 * NS + sg] = w.sol_f; V.sc[5 * NS + sg] = w.sol_b; V.sc[6 * NS + sg] = w.ir_f; V.sc[7 * NS + sg] = w.ir_b;
        V.energy_f[sg] += w.e_f;
        V.energy_b[sg] += w.e_b;
        V.nm_iters[sg] += w.nm_iters;
        if (w.status) atomicOr(&V.status[sg], w.status);
    }
    if (szone) V.zoneT[sgz] = s_zT[stid];
}

// ----------------------------------------------------------------------------------------------------------------
// K1/K2 (fast path)  hb_face_kernel<G, W, END>: same tile / lane-group layout and the same node march as
// hb_march_kernel, for tiles whose walls are all-massive (END = false) or whose only no-mass chunks are single nodes
// at the ends of walls (END = true; e.g. insulation | concrete | insulation).
//
// The time of a series is set by the dependent chain  scalar physics → march → scalar physics → zone  of the tiles
// with the most sub-steps, so this kernel shortens that chain:
//   * tiles are as small as the models allow (64 threads: more independent chains per SM, shorter node phase);
//   * the FRONT and BACK faces of wall j are evaluated by lane j of two different warps (two SM sub-partitions);
//   * between two node marches a face lane runs  post(s) → zone → pre(s+1)  as one straight-line region; the
//     sub-step weather table of the NEXT main step is prepared one step ahead (double buffer);
//   * "simple" tiles (every zone has exactly one wall, whose back face looks at it — a wall with its own zone):
//     the back-face lane advances its zone in registers, no shared-memory exchange and no pair barrier;
//     otherwise the zone sums go through shared memory and the front warp only signals (bar.arrive) unless one of
//     its faces looks at a zone;
//   * single-node no-mass ends are solved in the face lane's registers (solve_end_node) from the two temperatures
//     the node lanes publish per end — no shared-memory mirror of the wall;
//   * in the node phase the two edge slots of a lane are evaluated first and their values shuffled to the
//     neighbouring lanes while the six inner slots are computed.
// ----------------------------------------------------------------------------------------------------------------
#ifdef HB_PHASE_CLOCKS
#ifndef HB_CLK_MIN_S
#define HB_CLK_MIN_S 100   // tiles with at least this many sub-steps report (they run alone on their SM at the end)
#endif
__device__ unsigned long long hb_phase_clk[8];   // wait B1, node, wait B2, post, zone, pre, -, sub-steps
#define HB_CLK(k) { const long long _t = clock64(); clk[k] += _t - tlast; tlast = _t; }
#else
#define HB_CLK(k)
#endif
__device__ __forceinline__ void pair_sync() { asm volatile("bar.sync 1, 64;" ::: "memory"); }
__device__ __forceinline__ void pair_arrive() { asm volatile("bar.arrive 1, 64;" ::: "memory"); }

// Threads of a face-kernel block: the node lanes, but at least the warps the face role needs (one warp when both
// faces of every wall fit in it, 2·W <= 32; else one warp per side).
__host__ __device__ constexpr int face_threads(int G, int W) {
    return G * W >= (2 * W <= 32 ? 32 : 64) ? G * W : (2 * W <= 32 ? 32 : 64);
}

// One sub-step of RK4 for a lane's kNL slots (see sweep_first/sweep_next in hb_math.cuh for the algebra), with the
// edge slots first: their results are what the neighbouring lanes need for the next sweep.
template <int G, int NL>
__device__ __forceinline__ void node_rk4(double (&T)[NL], const double (&A)[NL], const double (&B)[NL]) {
    constexpr unsigned FULL = 0xffffffffu;
    double X[NL];
    double left = 0., right = 0.;
    if constexpr (G > 1) {
        left = __shfl_up_sync(FULL, T[NL - 1], 1, G);
        right = __shfl_down_sync(FULL, T[0], 1, G);
    }
    // sweep 1: X = M·T
    {
        const double d0 = T[1] - T[0], dl = T[0] - left;
        const double dr = right - T[NL - 1], dm = T[NL - 1] - T[NL - 2];
        X[0] = A[0] * d0 - B[0] * dl;
        X[NL - 1] = A[NL - 1] * dr - B[NL - 1] * dm;
        if constexpr (G > 1) {
            left = __shfl_up_sync(FULL, X[NL - 1], 1, G);
            right = __shfl_down_sync(FULL, X[0], 1, G);
        }
        double dprev = d0;
#pragma unroll
        for (int i = 1; i < NL - 1; ++i) {
            const double dnext = T[i + 1] - T[i];
            X[i] = A[i] * dnext - B[i] * dprev;
            dprev = dnext;
        }
#pragma unroll
        for (int i = 0; i < NL; ++i) T[i] += X[i];
    }
    // sweeps 2-4: X ← M·X, T += c·X
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const double c = k == 0 ? 1. / 2. : k == 1 ? 1. / 6. : 1. / 24.;
        const double d0 = X[1] - X[0], dl = X[0] - left;
        const double dr = right - X[NL - 1], dm = X[NL - 1] - X[NL - 2];
        const double x0 = A[0] * d0 - B[0] * dl;
        const double xl = A[NL - 1] * dr - B[NL - 1] * dm;
        if constexpr (G > 1) {
            if (k < 2) {
                left = __shfl_up_sync(FULL, xl, 1, G);
                right = __shfl_down_sync(FULL, x0, 1, G);
            }
        }
        double dprev = d0;
#pragma unroll
        for (int i = 1; i < NL - 1; ++i) {
            const double dnext = (i + 1 == NL - 1 ? X[NL - 1] : X[i + 1]) - X[i];
            X[i] = A[i] * dnext - B[i] * dprev;
            dprev = dnext;
        }
        X[0] = x0;
        X[NL - 1] = xl;
#pragma unroll
        for (int i = 0; i < NL; ++i) T[i] = fma(c, X[i], T[i]);
    }
}

#ifndef HB_FACE_OCC
#define HB_FACE_OCC 512   // resident threads per SM the face kernel is compiled for (512: 128 registers per thread)
#endif
// NL: slots per lane — kNL for the regular tiers; kNLBig / kNLHuge (G = 32 only, one block per SM) for walls beyond 254 nodes
template <int G, int W, int MODE, bool SIMPLE, int NL = kNL>
__global__ void __launch_bounds__(face_threads(G, W), NL == kNL ? HB_FACE_OCC / face_threads(G, W) : 1)
    hb_face_kernel(DevView V, SeriesDev P, const int32_t* tile_ids) {
    constexpr bool END = MODE != kFaceLean;    // single-node no-mass chunks at the ends of walls are possible
    constexpr bool LEAN = MODE != kFaceGen;    // one weather table per tile, no film override, no AmbientTemperature back
    extern __shared__ double s_dyn[];
    constexpr int NT = G * W;                   // node lanes (the tile's slab stride)
    constexpr int TPB = face_threads(G, W);     // a tile of 32 node lanes gets a second warp for the back faces
    static_assert(NL == kNL || (G == kMaxG && (NL == kNLBig || NL == kNLHuge)), "slots per lane");
    constexpr int NWARP = TPB / 32;
    constexpr bool ONE = 2 * W <= 32;           // front faces on lanes [0, W), back faces on lanes [W, 2W) of ONE warp
    constexpr bool PANE = ONE && (MODE == kFaceGen || MODE == kFacePane);   // panes: both faces in one warp
    static_assert(NT >= 32 && NT % 32 == 0 && W <= 32 && (ONE || NWARP >= 2), "tile shape");
    const Tile tile = V.tiles[tile_ids[blockIdx.x]];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t NS = V.n_dev_surfaces;

    __shared__ ZoneC s_zc[W];
    __shared__ double s_zT[W];
    __shared__ double s_pub[4][W];      // h·A and h·A·T_s of the front face, then of the back face (general mode)
    __shared__ double2 s_pub2[2][W];    // (h·A, h·A·T_s) of the front / back face (lean modes)
    __shared__ int32_t s_zf[W], s_zb[W];
    __shared__ double s_end[6][W];      // node 0, node 1, node n-2, node n-1, node 2, node n-3 (the last two: END)
    __shared__ double s_link[8][W];     // ghost_f, link_f, ghost_b, link_b, T(node 0), T(node n-1) (END)
    // sub-step weather of two consecutive main steps, [2][S] (dynamic: 2·S_max(group) rows, launch_face)
    SubWeather* const s_w = reinterpret_cast<SubWeather*>(s_dyn);
    __shared__ FaceC s_fc[2][W];        // per-face constants: read from shared memory where they are used

    const int S = tile.S;
    const double dt = tile.dt;

    // ---- node role ----
    const bool node_lane = NT == TPB || tid < NT;   // warp-uniform
    const int wl = node_lane ? tid / G : 0, part = tid % G;
    const int64_t g = (int64_t)tile.surface_begin + wl;
    const int n = V.n_nodes[g];
    const int iN = n - part * NL;        // lane-local index of the last node's slot (slot n)
    const int iG = n + 1 - part * NL;    // … of the back ghost (slot n+1)
    const int nmf = END ? (V.flags[g] >> 10) & 1 : 0, nmb = END ? (V.flags[g] >> 11) & 1 : 0;
    const int nmf2 = END ? (V.flags[g] >> 13) & 1 : 0, nmb2 = END ? (V.flags[g] >> 14) & 1 : 0;   // two-node end chunks
    double T[NL], A[NL], B[NL];
    if (node_lane) {
        const int64_t o = tile.node_offset + tid;
#pragma unroll
        for (int i = 0; i < NL; ++i) {
            T[i] = V.nodeT[o + (int64_t)i * NT];
            A[i] = V.coefA[o + (int64_t)i * NT];
            B[i] = V.coefB[o + (int64_t)i * NT];
        }
    } else {
#pragma unroll
        for (int i = 0; i < NL; ++i) T[i] = A[i] = B[i] = 0.;
    }
    // value of the lane-local slot k (0 <= k < NL) as a log2(NL)-level mux: no divergent branches
    auto slot = [&](int k) {
        double m[NL / 2];
#pragma unroll
        for (int i = 0; i < NL / 2; ++i) m[i] = (k & 1) ? T[2 * i + 1] : T[2 * i];
#pragma unroll
        for (int len = NL / 4, b = 2; len >= 1; len >>= 1, b <<= 1) {
#pragma unroll
            for (int i = 0; i < len; ++i) m[i] = (k & b) ? m[2 * i + 1] : m[2 * i];
        }
        return m[0];
    };
    static_assert((NL & (NL - 1)) == 0 && NL >= 8, "slot() is a power-of-two mux");
    const bool has_last = iN >= 0 && iN < NL, has_prev = END && iN - 1 >= 0 && iN - 1 < NL;
    const bool has_prev2 = END && iN - 2 >= 0 && iN - 2 < NL;
    double* const p_prev2 = &s_end[5][wl];
    double* const p_last = &s_end[3][wl];
    double* const p_prev = &s_end[2][wl];
    auto publish_ends = [&]() {
        if (part == 0) {
            s_end[0][wl] = T[1];
            if constexpr (END) {
                s_end[1][wl] = T[2];
                s_end[4][wl] = T[3];
            }
        }
        if constexpr (END) {
            // three muxes evaluated by every lane (one region instead of three divergent ones), predicated stores
            const double v_last = slot(iN & (NL - 1));
            const double v_prev = slot((iN - 1) & (NL - 1)), v_prev2 = slot((iN - 2) & (NL - 1));
            if (has_last) *p_last = v_last;
            if (has_prev) *p_prev = v_prev;
            if (has_prev2) *p_prev2 = v_prev2;
        } else {
            if (has_last) *p_last = slot(iN);
        }
    };
    if (node_lane) publish_ends();

    // ---- face role ----
    const int wf = (int)(blockIdx.x % NWARP);     // warp of the front faces
    const int wb = ONE ? wf : (wf + 1) % NWARP;   // warp of the back faces and of the zones
    const bool face_warp = warp == wf || warp == wb;
    const bool back = ONE ? lane >= W : warp == wb;
    const int kf = back ? 1 : 0;
    const int j = ONE ? (lane & (W - 1)) : lane;  // the face lane's wall (and, for back lanes, zone) in the tile
    const bool face_lane = face_warp && (ONE ? lane < 2 * W : lane < W);
    const int64_t sg = (int64_t)tile.surface_begin + j;
    const bool fwall = face_lane && V.n_nodes[sg] > 0;
    const bool szone = face_lane && back && j < tile.n_zones;
    const int64_t sgz = (int64_t)tile.zone_begin + j;
    const FaceC& c = s_fc[kf][j];
    FaceS w{};
    double zT = 0.;        // this lane's zone temperature
    double za0 = 0., zb0 = 0.;   // a, b of the zone's non-surface sources (constant over a series)
    double zvol = 1.;            // the zone's volume (simple tiles: register copy)
    bool any_front_zone = false, simple_ok = true;
    {
        ZoneC zc{};
        if (szone) {
            zc.vol = V.zvol[sgz]; zc.q = V.zq[sgz]; zc.inf_v = V.zinf_v[sgz]; zc.inf_t = V.zinf_t[sgz];
            zc.vent_v = V.zvent_v[sgz]; zc.vent_t = V.zvent_t[sgz]; zc.sb = V.z_surf_begin[sgz];
            zc.se = V.z_surf_end[sgz];
            zT = V.zoneT[sgz];
            s_zT[j] = zT;
            simple_ok = zc.sb == j && zc.se == j + 1;
            zone_sources(zc, za0, zb0);
            zvol = zc.vol;
        }
        if (face_lane && back) s_zc[j] = zc;
    }
    if (fwall) {
        WallC wc;
        load_wallc(V, sg, P.single_site != 0, wc);
        const int flg = V.flags[sg];
        const bool pane = PANE && ((flg >> 12) & 1);
        const bool nm = END && !pane && ((flg >> (back ? 11 : 10)) & 1);
        const bool nm2 = nm && ((flg >> (back ? 14 : 13)) & 1);
        load_facec(wc, back, back ? V.u_b[sg] : V.u_f[sg], nm, s_fc[kf][j], pane, back ? V.abL[sg] : V.af0[sg],
                   back ? V.ax_b[sg] : V.ax_f[sg], nm2, back ? V.u_b2[sg] : V.u_f2[sg]);
        w.h = V.sc[(0 + kf) * NS + sg]; w.flow = V.sc[(2 + kf) * NS + sg];
        w.sol = V.sc[(4 + kf) * NS + sg]; w.ir = V.sc[(6 + kf) * NS + sg];
        if (back) {
            s_zf[j] = wc.zf;
            s_zb[j] = wc.zb;
            simple_ok = simple_ok && wc.zb == j && szone;
        }
        any_front_zone = wc.zf >= 0;
    } else if (face_lane) {
        s_fc[kf][j] = FaceC{};   // the lean face path runs unpredicated: lanes without a wall work on zeros
        if (back) {
            s_zf[j] = -1;
            s_zb[j] = -1;
            simple_ok = simple_ok && !szone;
        }
    }
    // the front warp has to wait for the zone update only if one of its faces looks at a zone
    const bool front_waits = __syncthreads_or(any_front_zone ? 1 : 0) != 0;
    // SIMPLE tiles (TILE_SIMPLE, decided on the host): every zone has exactly one wall, wall j <-> zone j, whose
    // back face looks at it, and no front face looks at a zone.  The kernel re-checks and reports a mismatch.
    constexpr bool simple = SIMPLE;
    if (SIMPLE && !(__syncthreads_and(simple_ok ? 1 : 0) != 0 && !front_waits) && tid == 0)
        atomicOr(&V.status[tile.surface_begin], ST_NAN_ZONE);
    if (fwall) w.ts = s_end[back ? 3 : 0][j];
    // zone lanes: which faces of the tile look at my zone (bit j of zmf / zmb: front / back face of wall j)
    unsigned zmf = 0, zmb = 0;
    int zsb = 0, zcnt = 0;
    if (!SIMPLE && szone) {
        zsb = s_zc[j].sb;
        zcnt = s_zc[j].se - zsb;
        for (int q = zsb; q < zsb + zcnt; ++q) {
            if (s_zf[q] == j) zmf |= 1u << q;
            if (s_zb[q] == j) zmb |= 1u << q;
        }
    }
    // Segmented form (buildings): every zone is looked at by exactly the back faces of its own surfaces [sb, se) —
    // contiguous lanes of this warp — so all zones are summed at once by a segmented shuffle scan.
    int seg_last = -1;   // last lane of the segment this (back, zone-facing) face belongs to
    bool seg_ok_lane = LEAN && ONE && !SIMPLE;
    if (!SIMPLE && szone)
        seg_ok_lane = seg_ok_lane && zmf == 0 && zcnt <= 16 && zmb == (((1u << zcnt) - 1u) << zsb);
    if (!SIMPLE && fwall && back && c.zone >= 0) seg_last = W + s_zc[c.zone].se - 1;
    const bool seg_ok = !SIMPLE && __syncthreads_and(seg_ok_lane ? 1 : 0) != 0;
    const double ir_state = w.ir;   // front face: IR slot as found in the state (HB_IR_NET_GAIN with ε ≤ 1e-3)

    // sub-step weather of main step `step` into buffer step & 1 (all threads of the block)
    auto fill_weather = [&](int step) {
        if (!P.single_site) return;
        const double* w0 = P.weather + (int64_t)step * 3;
        if (P.weather_mode == HB_WEATHER_PER_SUBSTEP) w0 = P.weather + (int64_t)step * S * 3;
        for (int s = tid; s < S; s += TPB) {
            double t_out, wdir, wspeed;
            if (P.weather_mode == HB_WEATHER_LINEAR) {
                const double f = (double)(s + 1) / (double)S;
                t_out = lerp_rn(w0[0], w0[3], f);
                wdir = lerp_rn(w0[1], w0[4], f);
                wspeed = lerp_rn(w0[2], w0[5], f);
            } else if (P.weather_mode == HB_WEATHER_PER_SUBSTEP) {
                t_out = w0[3 * s + 0]; wdir = w0[3 * s + 1]; wspeed = w0[3 * s + 2];
            } else {
                t_out = w0[0]; wdir = w0[1]; wspeed = w0[2];
            }
            sub_weather(t_out, wdir, wspeed, s_w[(step & 1) * S + s]);
        }
    };
    // what the caller writes into SimulationState before a main step + the per-step derived quantities
    auto main_step_inputs = [&](int step) {
        if (P.input_mode != HB_INPUT_STATE) {
            const double* in = P.input_mode == HB_INPUT_SITE
                                   ? P.inputs + ((int64_t)c.site * P.n_steps + step) * 4
                                   : P.inputs + ((int64_t)step * P.n_desc_surfaces + c.desc_index) * 4;
            w.sol = in[kf];
            w.ir = in[2 + kf];
            if (!back && P.ir_mode == HB_IR_NET_GAIN) {
                // validate_wall_heat_transfer.rs:689-699 (front face only, as in the reference driver)
                if (c.eps > 1e-3) {
                    const double tk = w.ts + 273.15;
                    const double tk2 = tk * tk;
                    w.ir = w.ir / c.area / c.eps + kSigma * (tk2 * tk2);
                } else {
                    w.ir = ir_state;
                }
            }
        }
        face_begin_main_step(c, w);
    };
    // A pane's two faces (lanes j and W + j of this warp) exchange their border conditions and both solve the wall's
    // 2x2 no-mass system (solve_pane) with identical inputs; each keeps its own node.
    const unsigned pre_mask = __ballot_sync(0xffffffffu, LEAN ? face_lane : fwall);   // lanes that run pre()
    auto pane_solve = [&](double h, double rad, double trad) {
        const double sol_own = c.alpha_end * w.solar + c.alpha_x * __shfl_xor_sync(pre_mask, w.solar, W);
        const double p_air = __shfl_xor_sync(pre_mask, w.air, W), p_h = __shfl_xor_sync(pre_mask, h, W);
        const double p_rad = __shfl_xor_sync(pre_mask, rad, W), p_trad = __shfl_xor_sync(pre_mask, trad, W);
        const double p_ts = __shfl_xor_sync(pre_mask, w.ts, W), p_sol = __shfl_xor_sync(pre_mask, sol_own, W);
        if (c.fl & FC_PANE) {
            double t0 = back ? p_ts : w.ts, t1 = back ? w.ts : p_ts;
            unsigned long long it = 0;
            solve_pane<LEAN>(back ? p_air : w.air, back ? p_h : h, back ? p_rad : rad, back ? p_trad : trad,
                             back ? p_sol : sol_own, back ? w.air : p_air, back ? h : p_h, back ? rad : p_rad,
                             back ? trad : p_trad, back ? sol_own : p_sol, c.u_nb, t0, t1, it, w.status);
            if (!back) w.nm_iters += it;   // the chunk's iterations are counted once
            w.ts = back ? t1 : t0;
        }
    };
    // border conditions of sub-step s of main step `step` → links for the node lanes
    auto pre = [&](int step, int s) {
        if constexpr (LEAN) {
            // lean: single-site weather table, no film override, no AmbientTemperature back face, massive ends —
            // straight-line and branch-free (face_pre's arithmetic in face_pre's order)
            const SubWeather sw = s_w[(step & 1) * S + s];
            const int fl = c.fl;
            const bool is_out = fl & FC_OUT;
            const bool leeward =
                (fl & FC_TILTED) && !(add_rn(mul_rn(c.nx, sw.sin_wd), mul_rn(c.ny, sw.cos_wd)) > 0.0);
            w.forced = c.forced_c * (leeward ? 0.5 : 1.0) * sw.sqrt_ws;
            const double zone_t = simple ? zT : (c.zone >= 0 ? s_zT[c.zone] : 0.);
            w.air = is_out ? sw.t_out : (fl & FC_SPACE) ? zone_t : c.amb;
            const double trad = is_out ? w.trad_out : w.air;
            const double hn = tarp_lean(w.air - w.ts, c.kpos, c.kneg);
            const double h = is_out ? hn + w.forced : hn;
            if (h != h) w.status |= ST_NAN_H;
            const double xr = 273.15 + (trad + w.ts) / 2.;
            const double rad = c.eps4s * ((xr * xr) * xr);
            if constexpr (END) {
                if constexpr (PANE) pane_solve(h, rad, trad);
                if (fl & FC_NM2) {
                    double t_in = s_end[back ? 2 : 1][j];
                    solve_end_pair<true>(back, w.air, h, rad, trad, c.alpha_end * w.solar, c.u_nb, c.u_nb2,
                                         s_end[back ? 5 : 4][j], w.ts, t_in, w.nm_iters, w.status);
                    s_link[back ? 7 : 6][j] = t_in;
                } else if (fl & FC_NM)
                    w.ts = solve_end_node<true>(w.air, h, rad, trad, c.u_nb, s_end[back ? 2 : 1][j],
                                                c.alpha_end * w.solar, w.ts, w.nm_iters, w.status);
                s_link[back ? 5 : 4][j] = w.ts;
            }
            const double hs = h > 1e-12 ? h : 1e-12;
            s_link[back ? 2 : 0][j] = w.air + (rad * (trad - w.ts) + c.alpha_end * w.solar) * rcp_lean(hs);
            s_link[back ? 3 : 1][j] = hs * c.v_end;
            return;
        }
        SubWeather sw;
        if (P.single_site) {
            sw = s_w[(step & 1) * S + s];
        } else {
            const double* wtab = P.weather + (int64_t)c.site * P.weather_rows * 3;
            const double* w0 = wtab + (int64_t)step * 3;
            double t_out, wdir, wspeed;
            if (P.weather_mode == HB_WEATHER_LINEAR) {
                const double f = (double)(s + 1) / (double)S;
                t_out = lerp_rn(w0[0], w0[3], f);
                wdir = lerp_rn(w0[1], w0[4], f);
                wspeed = lerp_rn(w0[2], w0[5], f);
            } else if (P.weather_mode == HB_WEATHER_PER_SUBSTEP) {
                w0 = wtab + ((int64_t)step * S + s) * 3;
                t_out = w0[0]; wdir = w0[1]; wspeed = w0[2];
            } else {
                t_out = w0[0]; wdir = w0[1]; wspeed = w0[2];
            }
            sub_weather(t_out, wdir, wspeed, sw);
        }
        const double zone_t = simple ? zT : (c.zone >= 0 ? s_zT[c.zone] : 0.);
        double t_front = 0., ts_front = 0., t_nb = 0.;
        if (c.fl & FC_BACK_AMB) {
            t_front = (c.fl & FC_FRONT_OUT) ? sw.t_out : (c.fl & FC_FRONT_SPACE) ? s_zT[c.zone_front] : c.amb_front;
            ts_front = s_end[0][j];
        }
        if constexpr (END)
            if (c.fl & FC_NM) t_nb = s_end[back ? 2 : 1][j];
        double ghost, link;
        FaceEnv e;
        face_env(c, w, sw, zone_t, t_front, ts_front, e);
        if constexpr (PANE) pane_solve(e.h, e.rad, e.trad);
        const double t_nb2 = (END && (c.fl & FC_NM2)) ? s_end[back ? 5 : 4][j] : 0.;
        face_links(c, w, e, t_nb, t_nb2, ghost, link);
        if constexpr (END)
            if (c.fl & FC_NM2) s_link[back ? 7 : 6][j] = t_nb;
        s_link[back ? 2 : 0][j] = ghost;
        s_link[back ? 3 : 1][j] = link;
        if constexpr (END) s_link[back ? 5 : 4][j] = w.ts;
    };

    fill_weather(0);
    if (fwall) main_step_inputs(0);
    __syncthreads();
    if (LEAN ? face_lane : fwall) pre(0, 0);

#ifdef HB_PHASE_CLOCKS
    long long clk[8] = {0, 0, 0, 0, 0, 0, 0, 0}, tlast = clock64();
#endif
    int rec = 0;
    for (int step = 0; step < P.n_steps; ++step) {
        if (step + 1 < P.n_steps) fill_weather(step + 1);   // read from the next main step's first pre() on
        if (szone) {   // scheduled zone inputs: this main step's values, sources re-summed in the reference's order
            ZoneC zc = s_zc[j];
            if (load_zone_schedule(P, step, sgz, zc)) {
                s_zc[j] = zc;
                zone_sources(zc, za0, zb0);
            }
        }
        for (int s = 0; s < S; ++s) {
#ifdef HB_PHASE_CLOCKS
            tlast = clock64();
            clk[7] += 1;
#endif
            __syncthreads();
            HB_CLK(0)
            // ================= node role: RK4 of the massive nodes =================
            if (node_lane) {
                const double ghost_f = s_link[0][wl], bf = s_link[1][wl], ghost_b = s_link[2][wl], ab = s_link[3][wl];
                // no-mass end nodes as solved by the face lanes (END): loaded once, placed by selects
                double tf0 = 0., tb0 = 0., tf1 = 0., tb1 = 0.;
                if constexpr (END) {
                    tf0 = s_link[4][wl]; tb0 = s_link[5][wl]; tf1 = s_link[6][wl]; tb1 = s_link[7][wl];
                }
                if constexpr (END) {
                    const bool first = part == 0;
                    T[0] = first ? ghost_f : T[0];
                    B[1] = first ? bf : B[1];
                    T[1] = (first && nmf) ? tf0 : T[1];
                    T[2] = (first && nmf2) ? tf1 : T[2];
                } else if (part == 0) {
                    T[0] = ghost_f;
                    B[1] = bf;
                }
#pragma unroll
                for (int i = 0; i < NL; ++i) {
                    A[i] = i == iN ? ab : A[i];
                    if constexpr (END) {
                        T[i] = (i == iN && nmb) ? tb0 : T[i];
                        T[i] = (i == iN - 1 && nmb2) ? tb1 : T[i];
                    }
                    T[i] = i == iG ? ghost_b : T[i];
                }
                node_rk4<G, NL>(T, A, B);
                publish_ends();
            }
            HB_CLK(1)
            __syncthreads();
            HB_CLK(2)
            // ================= face role: post(s) → zones → pre(s+1) =================
            if (face_warp) {   // warp-uniform
                double hA = 0., hAT = 0.;
                if constexpr (LEAN) {
                    if (face_lane) {   // face_post's arithmetic, unpredicated
                        const double ts = s_end[back ? 3 : 0][j];
                        const double hn = tarp_lean(w.air - ts, c.kpos, c.kneg);
                        const double h = (c.fl & FC_OUT) ? hn + w.forced : hn;
                        if (h != h) w.status |= ST_NAN_H;
                        w.ts = ts;
                        w.h = h;
                        w.flow = (ts - w.air) * h;
                        w.e += w.flow * c.area * dt;
                        hA = h * c.area;
                        hAT = hA * ts;
                    }
                } else if (fwall) {
                    const double ts = s_end[back ? 3 : 0][j];
                    const double ts_front = (c.fl & FC_BACK_AMB) ? s_end[0][j] : 0.;
                    face_post(c, w, ts, ts_front, dt, hA, hAT);
                }
                HB_CLK(3)
                if constexpr (simple) {
                    if constexpr (LEAN) {   // the lane's own wall is the zone's only surface (back face)
                        const double inv_cc = zone_inv_c_t<true>(zvol, zT);
                        const double tn = zone_advance<true>(zT, dt, inv_cc, za0 + hAT, zb0 + hA);
                        if (szone) {
                            if (tn != tn) atomicOr(&V.status[sg], ST_NAN_ZONE);
                            zT = tn;
                        }
                    } else if (szone) {
                        const double inv_cc = zone_inv_c(s_zc[j], zT);
                        const double a = za0 + hAT, b = zb0 + hA;
                        const double tn = zone_advance(zT, dt, inv_cc, a, b);
                        if (tn != tn) atomicOr(&V.status[sg], ST_NAN_ZONE);
                        zT = tn;
                    }
                } else if (LEAN && ONE && seg_ok) {
                    double va = hAT, vb = hA;
#pragma unroll
                    for (int o = 1; o <= 8; o <<= 1) {
                        const double ta = __shfl_down_sync(0xffffffffu, va, o), tb = __shfl_down_sync(0xffffffffu, vb, o);
                        if (lane + o <= seg_last) { va += ta; vb += tb; }
                    }
                    const int src = szone && zcnt > 0 ? W + zsb : lane;
                    const double asum = __shfl_sync(0xffffffffu, va, src), bsum = __shfl_sync(0xffffffffu, vb, src);
                    if (szone) {
                        const double inv_cc = zone_inv_c_t<true>(zvol, zT);
                        const double tn = zone_advance<true>(zT, dt, inv_cc, zcnt > 0 ? za0 + asum : za0,
                                                             zcnt > 0 ? zb0 + bsum : zb0);
                        if (tn != tn) atomicOr(&V.status[(int64_t)tile.surface_begin + zsb], ST_NAN_ZONE);
                        s_zT[j] = tn;
                        zT = tn;
                    }
                    __syncwarp();
                } else if (LEAN) {
                    // the faces publish (h·A, h·A·T_s) as one 16-byte store; a zone lane sums its surfaces with a
                    // fixed, fully unrolled window of 8 (all loads in flight at once, four independent partial
                    // sums) — deterministic; differs from the reference's sequential order by rounding only
                    if (fwall) s_pub2[kf][j] = make_double2(hA, hAT);
                    if constexpr (ONE) {
                        __syncwarp();
                    } else {
                        if (back || front_waits) pair_sync(); else pair_arrive();
                    }
                    if (szone) {
                        double af = 0., bf = 0., ab = 0., bb = 0.;
#pragma unroll
                        for (int q = 0; q < 8; ++q) {
                            const int jj = zsb + q < W ? zsb + q : W - 1;
                            const double2 vf = s_pub2[0][jj], vb = s_pub2[1][jj];
                            const bool uf = q < zcnt && ((zmf >> jj) & 1u), ub = q < zcnt && ((zmb >> jj) & 1u);
                            bf += uf ? vf.x : 0.;
                            af += uf ? vf.y : 0.;
                            bb += ub ? vb.x : 0.;
                            ab += ub ? vb.y : 0.;
                        }
                        for (int q = 8; q < zcnt; ++q) {
                            const int jj = zsb + q;
                            const double2 vf = s_pub2[0][jj], vb = s_pub2[1][jj];
                            if ((zmf >> jj) & 1u) { bf += vf.x; af += vf.y; }
                            if ((zmb >> jj) & 1u) { bb += vb.x; ab += vb.y; }
                        }
                        const double inv_cc = zone_inv_c_t<true>(zvol, zT);
                        const double tn = zone_advance<true>(zT, dt, inv_cc, za0 + (af + ab), zb0 + (bf + bb));
                        if (tn != tn) atomicOr(&V.status[(int64_t)tile.surface_begin + zsb], ST_NAN_ZONE);
                        s_zT[j] = tn;
                        zT = tn;
                    }
                    if constexpr (ONE) {
                        __syncwarp();
                    } else {
                        if (front_waits) pair_sync(); else if (back) __syncwarp();
                    }
                } else {
                    if (fwall) {
                        s_pub[back ? 2 : 0][j] = hA;
                        s_pub[back ? 3 : 1][j] = hAT;
                    }
                    if constexpr (ONE) {
                        __syncwarp();
                    } else {
                        if (back || front_waits) pair_sync(); else pair_arrive();
                    }
                    if (szone) {
                        const double tn = zone_update_ab<LEAN>(s_zc[j], j, s_zT[j], dt, za0, zb0, s_pub[0], s_pub[1],
                                                               s_pub[2], s_pub[3], s_zf, s_zb);
                        if (tn != tn) atomicOr(&V.status[(int64_t)tile.surface_begin + s_zc[j].sb], ST_NAN_ZONE);
                        s_zT[j] = tn;
                        zT = tn;
                    }
                    if constexpr (ONE) {
                        __syncwarp();
                    } else {
                        if (front_waits) pair_sync(); else if (back) __syncwarp();
                    }
                }
                HB_CLK(4)
                if ((LEAN ? face_lane : fwall) && s + 1 < S) pre(step, s + 1);
                HB_CLK(5)
            }
        }

        if (P.record_every > 0 && (step + 1) % P.record_every == 0) {
            if (P.zone_rec && szone) P.zone_rec[(int64_t)rec * V.n_dev_zones + sgz] = zT;
            if (P.node_rec) {
#pragma unroll
                for (int i = 0; i < NL; ++i)
                    if (node_lane) P.node_rec[(int64_t)rec * P.slab_len + tile.node_offset + (int64_t)i * NT + tid] = T[i];
                if (fwall) {
                    double* rr = P.sc_rec + (int64_t)rec * 8 * NS;
                    rr[(0 + kf) * NS + sg] = w.h; rr[(2 + kf) * NS + sg] = w.flow;
                    rr[(4 + kf) * NS + sg] = w.sol; rr[(6 + kf) * NS + sg] = w.ir;
                }
            }
            ++rec;
        }
        if (step + 1 < P.n_steps) {
            if (fwall) main_step_inputs(step + 1);
            if (LEAN ? face_lane : fwall) pre(step + 1, 0);
        }
    }

#ifdef HB_PHASE_CLOCKS
    if (szone && j == 0 && S >= HB_CLK_MIN_S)
        for (int q = 0; q < 8; ++q) atomicAdd(&hb_phase_clk[q], (unsigned long long)clk[q]);
#endif
    // ---- write the state back ----
#pragma unroll
    for (int i = 0; i < NL; ++i)
        if (node_lane) V.nodeT[tile.node_offset + (int64_t)i * NT + tid] = T[i];
    if (fwall) {
        V.sc[(0 + kf) * NS + sg] = w.h; V.sc[(2 + kf) * NS + sg] = w.flow;
        V.sc[(4 + kf) * NS + sg] = w.sol; V.sc[(6 + kf) * NS + sg] = w.ir;
        if (back) V.energy_b[sg] += w.e; else V.energy_f[sg] += w.e;
        if (w.nm_iters) atomicAdd(&V.nm_iters[sg], w.nm_iters);
        if (w.status) atomicOr(&V.status[sg], w.status);
    }
    if (szone) V.zoneT[sgz] = zT;
}

#ifdef HB_AUX_KERNELS
// ----------------------------------------------------------------------------------------------------
// K3: ensemble reduction. out[0..8): Σzone, min zone, max zone, Σnode, min node, max node, E_f, E_b;
// iout[0..4): n_zones, n_nodes, no-mass iterations, status OR.   One block per tile (grid-stride), 256 threads.
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

constexpr int kReduceBlock = 256;

__global__ void __launch_bounds__(kReduceBlock) hb_reduce_kernel(DevView V, int n_tiles, double* out,
                                                                  unsigned long long* iout) {
    double sz = 0., mnz = CUDART_INF, mxz = -CUDART_INF, sn = 0., mnn = CUDART_INF, mxn = -CUDART_INF, ef = 0., eb = 0.;
    unsigned long long nz = 0, nn = 0, it = 0, st = 0;
    const int t = threadIdx.x;
    for (int ti = blockIdx.x; ti < n_tiles; ti += gridDim.x) {
        const Tile tile = V.tiles[ti];
        const int TPB = tile.G * tile.W;
        for (int th = t; th < TPB; th += kReduceBlock) {
            const int wl = th / tile.G, part = th % tile.G;
            const int n = V.n_nodes[(int64_t)tile.surface_begin + wl];
            for (int i = 0; i < tile.NL; ++i) {
                const int slot = part * tile.NL + i;
                // block tiles: slots 1..n hold the nodes (0 and n+1 are ghosts); warp tiles: slots 0..n-1
                if ((tile.flags & TILE_WARP) ? slot < n : (slot >= 1 && slot <= n)) {
                    const double v = V.nodeT[tile.node_offset + (int64_t)i * TPB + th];
                    sn += v;
                    mnn = fmin(mnn, v);
                    mxn = fmax(mxn, v);
                    nn += 1;
                }
            }
        }
        if (t < tile.W) {
            const int64_t g = (int64_t)tile.surface_begin + t;
            if (V.n_nodes[g] > 0) {
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

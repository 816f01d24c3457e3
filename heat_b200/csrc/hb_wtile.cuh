// hb_wtile.cuh — K1/K2 "warp tile" kernel (sm_100a): one WARP marches one tile, with no block barrier and no
// shared-memory hand-off anywhere in the sub-step loop.
//
// Why: the block-tile kernel (hb_face_kernel) is bound by the dependent chain  node march → face physics → node march
// of a tile: three of a tile's four warps wait at the block barrier while one warp runs post → zone → pre
// (profiles/r1j_summary.md: 2.9 warps stalled on the barrier per issued instruction, FP64 pipe 34 % busy), and the
// register file admits only four such tiles per SM.  Here every warp is its own tile:
//   * a wall is owned by G consecutive lanes (G = 1 … 32), each lane keeping the temperatures of NL (8 … 32) NODES in
//     registers — no ghost slots: the film links enter through the halo values of the first / last node (capacity
//     n <= G·NL); the tridiagonal coefficients sit beside them in registers (NL = 8; all-massive walls: NL <= 12) or
//     in shared memory (wt_coef_in_smem: half the registers per node);
//   * the FRONT face of a wall is evaluated by the first lane of its group, the BACK face (and, for a wall with its own
//     zone, the zone) by the last lane (G = 1: both by the one lane) — the face results reach the node lanes through
//     two warp shuffles, the end-node temperatures come back through a register mux + one shuffle;
//   * warps never wait for each other, so the FP64 pipe of a sub-partition is filled by whichever of its 2–4 resident
//     warps has an instruction ready: the kernel is bound by FP64 issue, not by a barrier chain.
// Arithmetic and operation order of the face physics are those of hb_face_kernel's lean modes (hb_math.cuh):
// model.rs:79-96,145-169,489-597,650-674, surface.rs:37-46,596-717,790-898,916-948, convection.rs:87-168.
#pragma once
#include <cuda_runtime.h>

#include "hb_math.cuh"

namespace hb {

// Registers per thread each geometry is compiled for.  Every warp of a block sits on one of the SM's four
// sub-partitions, whose register file (16 384 registers) holds 3 warps of <= 168 registers or 2 of <= 255: there is
// nothing in between.  (__maxnreg__, not the min-blocks argument of __launch_bounds__: for 32-thread blocks ptxas caps
// that form at 168 registers whatever the block count asked for.)
// Tiles with no-mass ends (MODE != lean) carry two fixed-point solvers and are, in the ensembles of BASELINE.json, the
// walls with the most sub-steps — the longest dependent chains of a series: they get all the registers there are.
#ifndef HB_WT_CSM16
#define HB_WT_CSM16 1   // all-massive tiles of 16 nodes per lane: coefficients in shared memory, 168 registers
#endif
#ifndef HB_WT_CSM_ALL
#define HB_WT_CSM_ALL 0   // experiment: every all-massive geometry with shared-memory coefficients, NL <= 12 at 128 registers
#endif
#ifndef HB_WT_END16_168
#define HB_WT_END16_168 1   // no-mass-end tiles of 16 nodes per lane at 168 registers too (590 bytes of spills, 3 warps: +3 % on configs[4])
#endif
#ifndef HB_WT_PANE_168
#define HB_WT_PANE_168 1    // building tiles of 12 nodes per lane (G >= 2) at 168 registers (310-370 bytes of spills, 3 warps: +4 % on configs[3])
#endif
#ifndef HB_WT_CSM_PANE
#define HB_WT_CSM_PANE 1  // buildings, 12 / 16 nodes per lane: shared-memory coefficients (255 registers, no spills: +10 % on configs[3])
#endif
#ifndef HB_WT_CSM_END
#define HB_WT_CSM_END 1   // walls with no-mass ends, 12 / 16 nodes per lane: shared-memory coefficients, 12 at 168 registers (+20 % on configs[4])
#endif
__host__ __device__ constexpr int wt_max_regs(int G, int NL, int MODE) {
    if (HB_WT_CSM_ALL && MODE == kFaceLean && NL <= 12) return 128;
    if (HB_WT_CSM_END && MODE == kFaceLeanEnd && NL > 8 && NL <= (HB_WT_END16_168 ? 16 : 12)) return 168;
    if (HB_WT_CSM_PANE && HB_WT_PANE_168 && MODE == kFacePane && NL > 8 && NL <= 12 && G >= 2) return 168;
    return (NL <= 12 || (HB_WT_CSM16 && NL == 16)) && MODE == kFaceLean ? 168 : 255;
}
// Geometries of more than 16 nodes per lane keep only T and the RK4 increment in registers: the coefficients A, B live
// in shared memory ([slot][lane] pairs, one 16-byte load per node and sweep, conflict-free) — half the registers per
// node, so a face evaluation is shared by twice the nodes.
__host__ __device__ constexpr bool wt_coef_in_smem(int NL, int MODE) {
    return NL > 16 || (HB_WT_CSM16 && NL == 16 && MODE == kFaceLean) || (HB_WT_CSM_ALL && MODE == kFaceLean) ||
           (HB_WT_CSM_END && MODE == kFaceLeanEnd && NL > 8) || (HB_WT_CSM_PANE && MODE == kFacePane && NL > 8);
}
// bytes of shared memory per warp: the sub-step weather table + (NL > 16) the coefficient pairs
__host__ __device__ constexpr size_t wt_smem_per_warp(int NL, int MODE, int s_rows) {
    return (size_t)s_rows * sizeof(SubWeather) + (wt_coef_in_smem(NL, MODE) ? (size_t)NL * 32 * sizeof(double2) : 0);
}

// T[k], k lane-dependent, as a select tree (registers cannot be indexed)
template <int NL>
__device__ __forceinline__ double wt_mux(const double (&T)[NL], int k) {
    constexpr int P = NL <= 8 ? 8 : NL <= 16 ? 16 : 32;
    static_assert(NL <= 32, "wt_mux covers 32 slots");
    double m[P / 2];
#pragma unroll
    for (int i = 0; i < P / 2; ++i) {
        const double a = T[2 * i < NL ? 2 * i : NL - 1], b = T[2 * i + 1 < NL ? 2 * i + 1 : NL - 1];
        m[i] = (k & 1) ? b : a;
    }
#pragma unroll
    for (int len = P / 4, bit = 2; len >= 1; len >>= 1, bit <<= 1) {
#pragma unroll
        for (int i = 0; i < len; ++i) m[i] = (k & bit) ? m[2 * i + 1] : m[2 * i];
    }
    return m[0];
}

// One sub-step of RK4 (surface.rs:228-308 in Horner / flux form, see node_rk4 in hb_kernels.cuh) for a lane's NL nodes.
// `first`: the lane holds node 0 — its left neighbour is the front ghost (temperature gf in the first sweep, 0 later:
// the ghost does not move); `edge`: the lane's last slot is the wall's last node — its right neighbour is the back ghost.
template <int G, int NL>
__device__ __forceinline__ void wt_rk4(double (&T)[NL], const double (&A)[NL], const double (&B)[NL], bool first,
                                       double gf, bool edge, double gb) {
    constexpr unsigned FULL = 0xffffffffu;
    double X[NL];
    double left = gf, right = gb;
    if constexpr (G > 1) {
        const double l = __shfl_up_sync(FULL, T[NL - 1], 1, G), r = __shfl_down_sync(FULL, T[0], 1, G);
        left = first ? gf : l;
        right = edge ? gb : r;
    }
    // sweep 1: X = M·T
    {
        const double d0 = T[1] - T[0], dl = T[0] - left;
        const double dr = right - T[NL - 1], dm = T[NL - 1] - T[NL - 2];
        X[0] = A[0] * d0 - B[0] * dl;
        X[NL - 1] = A[NL - 1] * dr - B[NL - 1] * dm;
        if constexpr (G > 1) {
            const double l = __shfl_up_sync(FULL, X[NL - 1], 1, G), r = __shfl_down_sync(FULL, X[0], 1, G);
            left = first ? 0. : l;
            right = edge ? 0. : r;
        } else {
            left = 0.;
            right = 0.;
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
                const double l = __shfl_up_sync(FULL, xl, 1, G), r = __shfl_down_sync(FULL, x0, 1, G);
                left = first ? 0. : l;
                right = edge ? 0. : r;
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

// The same with the coefficients in shared memory: ab[i * 32] = (A[i], B[i]) of this lane.
template <int G, int NL>
__device__ __forceinline__ void wt_rk4_smem(double (&T)[NL], const double2* ab, bool first, double gf, bool edge, double gb) {
    constexpr unsigned FULL = 0xffffffffu;
    double X[NL];
    double left = gf, right = gb;
    if constexpr (G > 1) {
        const double l = __shfl_up_sync(FULL, T[NL - 1], 1, G), r = __shfl_down_sync(FULL, T[0], 1, G);
        left = first ? gf : l;
        right = edge ? gb : r;
    }
    const double2 c0 = ab[0], cl = ab[(NL - 1) * 32];
    {
        const double d0 = T[1] - T[0], dl = T[0] - left;
        const double dr = right - T[NL - 1], dm = T[NL - 1] - T[NL - 2];
        X[0] = c0.x * d0 - c0.y * dl;
        X[NL - 1] = cl.x * dr - cl.y * dm;
        if constexpr (G > 1) {
            const double l = __shfl_up_sync(FULL, X[NL - 1], 1, G), r = __shfl_down_sync(FULL, X[0], 1, G);
            left = first ? 0. : l;
            right = edge ? 0. : r;
        } else {
            left = 0.;
            right = 0.;
        }
        double dprev = d0;
#pragma unroll
        for (int i = 1; i < NL - 1; ++i) {
            const double2 c = ab[i * 32];
            const double dnext = T[i + 1] - T[i];
            X[i] = c.x * dnext - c.y * dprev;
            dprev = dnext;
        }
#pragma unroll
        for (int i = 0; i < NL; ++i) T[i] += X[i];
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const double c = k == 0 ? 1. / 2. : k == 1 ? 1. / 6. : 1. / 24.;
        const double d0 = X[1] - X[0], dl = X[0] - left;
        const double dr = right - X[NL - 1], dm = X[NL - 1] - X[NL - 2];
        const double x0 = c0.x * d0 - c0.y * dl;
        const double xl = cl.x * dr - cl.y * dm;
        if constexpr (G > 1) {
            if (k < 2) {
                const double l = __shfl_up_sync(FULL, xl, 1, G), r = __shfl_down_sync(FULL, x0, 1, G);
                left = first ? 0. : l;
                right = edge ? 0. : r;
            }
        }
        double dprev = d0;
#pragma unroll
        for (int i = 1; i < NL - 1; ++i) {
            const double2 cc = ab[i * 32];
            const double dnext = (i + 1 == NL - 1 ? X[NL - 1] : X[i + 1]) - X[i];
            X[i] = cc.x * dnext - cc.y * dprev;
            dprev = dnext;
        }
        X[0] = x0;
        X[NL - 1] = xl;
#pragma unroll
        for (int i = 0; i < NL; ++i) T[i] = fma(c, X[i], T[i]);
    }
}

#ifdef HB_PHASE_CLOCKS
__device__ unsigned long long hb_wt_clk[8];   // pre, place, rk4, gather, post, zone, -, sub-steps
#define HB_WCLK(k) { const long long _t = clock64(); wclk[k] += _t - wlast; wlast = _t; }
#else
#define HB_WCLK(k)
#endif

// Results of one face's pre-march evaluation, handed to the node lanes.
struct WtLinks {
    double ghost, link;   // equivalent air temperature and h·dt/C of the end node
    double t_in;          // two-node no-mass end chunk: the chunk's inner node
};


// One warp marches tile `tile_slot` of its launch group over the whole series.
template <int G, int NL, int MODE, bool SIMPLE>
__device__ __forceinline__ void wt_march_tile(const DevView& V, const SeriesDev& P, const int32_t* tile_ids, int tile_slot,
                                              int s_rows, double* s_dyn) {
    constexpr bool END = MODE != kFaceLean;   // no-mass chunks of one or two nodes at the ends of walls are possible
    constexpr bool PANE = MODE == kFacePane;  // … and single panes (two no-mass nodes, both faces solve one 2x2 system)
    constexpr int NF = G == 1 ? 2 : 1;        // faces evaluated by a face lane
    constexpr unsigned FULL = 0xffffffffu;
    constexpr int W = 32 / G;
    static_assert(G >= 1 && G <= 32 && (G & (G - 1)) == 0 && NL >= 4, "geometry");
    // [S] sub-step weather of the current main step, one table per warp
    SubWeather* const s_w = reinterpret_cast<SubWeather*>(s_dyn) + (size_t)(threadIdx.x >> 5) * s_rows;

    const Tile tile = V.tiles[tile_ids[tile_slot]];
    const int lane = threadIdx.x & 31;
    const int64_t NS = V.n_dev_surfaces;
    const int S = tile.S;
    const double dt = tile.dt;

    // ---- node role ----
    const int wl = lane / G, part = lane % G, base = lane - part;
    const int64_t sg = (int64_t)tile.surface_begin + wl;
    const int n = V.n_nodes[sg];
    const int last = n - 1 - part * NL;   // lane-local index of the wall's last node (this lane holds it iff 0 <= last < NL)
    const bool first = part == 0, edge = last == NL - 1;
    const int nm1 = n > 0 ? n - 1 : 0, nm2 = n > 1 ? n - 2 : 0, nm3 = n > 2 ? n - 3 : 0;
    const int wflags = V.flags[sg];
    const bool nmf = END && ((wflags >> 10) & 1), nmb = END && ((wflags >> 11) & 1);
    const bool nmf2 = END && ((wflags >> 13) & 1), nmb2 = END && ((wflags >> 14) & 1);
    constexpr bool CSM = wt_coef_in_smem(NL, MODE);
    constexpr int NC = CSM ? 1 : NL;   // coefficient registers (none when they live in shared memory)
    double T[NL], A[NC], B[NC];
    // this lane's column of the warp's coefficient pairs: ab[i * 32] = (A[i], B[i])
    double2* const ab = reinterpret_cast<double2*>(reinterpret_cast<char*>(s_dyn) + (size_t)blockDim.x / 32 * s_rows * sizeof(SubWeather)) +
                        (size_t)(threadIdx.x >> 5) * NL * 32 + lane;
    {
        const int64_t o = tile.node_offset + lane;
#pragma unroll
        for (int i = 0; i < NL; ++i) {
            T[i] = V.nodeT[o + (int64_t)i * 32];
            if constexpr (CSM) {
                ab[i * 32] = make_double2(V.coefA[o + (int64_t)i * 32], V.coefB[o + (int64_t)i * 32]);
            } else {
                A[i] = V.coefA[o + (int64_t)i * 32];
                B[i] = V.coefB[o + (int64_t)i * 32];
            }
        }
    }

    // ---- face role: front face on the first lane of the group, back face (and the zone) on the last ----
    const bool bk = part == G - 1;
    const bool face_lane = first || bk;
    const bool fwall = face_lane && n > 0;
    // SIMPLE tiles: wall j has its own zone j.  Otherwise (buildings): every model of the tile has ONE zone, looked at by
    // the back faces of all its walls [sb, se) and by no front face; the zone lives on the back lane of wall sb.
    int zown = SIMPLE ? wl : -1, seg_end = wl + 1, zsrc = lane;
    if constexpr (!SIMPLE) {
        for (int j = 0; j < tile.n_zones; ++j) {
            const int sb = V.z_surf_begin[tile.zone_begin + j], se = V.z_surf_end[tile.zone_begin + j];
            if (sb == wl) zown = j;
            if (wl >= sb && wl < se) {
                seg_end = se;
                zsrc = sb * G + G - 1;
            }
        }
    }
    const bool szone = bk && zown >= 0 && zown < tile.n_zones;
    const int64_t sgz = (int64_t)tile.zone_begin + (zown >= 0 ? zown : 0);
    FaceC c[NF];
    FaceS w[NF];
    double ir_state[NF];
    double zT = 0., za0 = 0., zb0 = 0., zvol = 1.;
    bool simple_ok = true;
    if (szone) {
        ZoneC zc{};
        zc.vol = V.zvol[sgz]; zc.q = V.zq[sgz]; zc.inf_v = V.zinf_v[sgz]; zc.inf_t = V.zinf_t[sgz];
        zc.vent_v = V.zvent_v[sgz]; zc.vent_t = V.zvent_t[sgz]; zc.sb = V.z_surf_begin[sgz]; zc.se = V.z_surf_end[sgz];
        zT = V.zoneT[sgz];
        zone_sources(zc, za0, zb0);
        zvol = zc.vol;
        simple_ok = SIMPLE ? zc.sb == wl && zc.se == wl + 1 : zc.sb == wl && zc.se <= W;
    }
#pragma unroll
    for (int f = 0; f < NF; ++f) {
        const bool back = G == 1 ? f == 1 : bk;
        const int kf = back ? 1 : 0;
        w[f] = FaceS{};
        if (fwall) {
            WallC wc;
            load_wallc(V, sg, P.single_site != 0, wc);
            const bool pane = PANE && ((wflags >> 12) & 1);
            const bool win4 = PANE && ((wflags >> 15) & 1);
            const bool nm = END && !pane && ((wflags >> (back ? 11 : 10)) & 1);
            const bool nm2f = nm && ((wflags >> (back ? 14 : 13)) & 1);
            load_facec(wc, back, back ? V.u_b[sg] : V.u_f[sg], nm, c[f], pane, back ? V.abL[sg] : V.af0[sg],
                       back ? V.ax_b[sg] : V.ax_f[sg], nm2f, back ? V.u_b2[sg] : V.u_f2[sg], win4);
            w[f].h = V.sc[(0 + kf) * NS + sg]; w[f].flow = V.sc[(2 + kf) * NS + sg];
            w[f].sol = V.sc[(4 + kf) * NS + sg]; w[f].ir = V.sc[(6 + kf) * NS + sg];
            // the zone topology decided on the host (TILE_SIMPLE / one zone per model behind its walls): re-checked here
            if (back) simple_ok = simple_ok && wc.zf < 0 && (SIMPLE ? wc.zb == wl && szone : wc.zb >= 0 && wc.zb < tile.n_zones);
        } else {
            c[f] = FaceC{};   // the lean face path runs unpredicated: lanes without a face work on zeros
            if (SIMPLE && back && face_lane) simple_ok = simple_ok && !szone;
        }
        ir_state[f] = w[f].ir;
    }
    // double-glazed windows (flags bit 15): the front face lane solves the wall's four no-mass nodes (all in its slots)
    const bool win4_lane = PANE && first && n == 4 && ((wflags >> 15) & 1);
    double w_a1f = 0., w_a1b = 0., w_a2f = 0., w_a2b = 0., w_u23 = 0.;
    CavityDev w_cav{0, 0, 1., 1., 0., 1., 1.};
    if (win4_lane) {
        w_a1f = V.wa1f[sg]; w_a1b = V.wa1b[sg]; w_a2f = V.wa2f[sg]; w_a2b = V.wa2b[sg];
        w_u23 = V.u_b[sg];
        w_cav = V.cav[V.cav_off[sg]];
    }
    if (!__all_sync(FULL, simple_ok) && lane == 0) atomicOr(&V.status[tile.surface_begin], ST_NAN_ZONE);

    // end-node temperatures: node 0 (and 1, 2) on the first lane, node n-1 (n-2, n-3) wherever they live
    double ts_f = T[0], t1_f = T[1], t2_f = T[2];
    double ts_b, t1_b = 0., t2_b = 0.;
    auto gather_back = [&]() {
        const double m0 = wt_mux<NL>(T, nm1 % NL);
        ts_b = G == 1 ? m0 : __shfl_sync(FULL, m0, base + nm1 / NL);
        if constexpr (END) {
            const double m1 = wt_mux<NL>(T, nm2 % NL), m2 = wt_mux<NL>(T, nm3 % NL);
            t1_b = G == 1 ? m1 : __shfl_sync(FULL, m1, base + nm2 / NL);
            t2_b = G == 1 ? m2 : __shfl_sync(FULL, m2, base + nm3 / NL);
        }
    };
    gather_back();
#pragma unroll
    for (int f = 0; f < NF; ++f) {
        const bool back = G == 1 ? f == 1 : bk;
        w[f].ts = back ? ts_b : ts_f;
    }

    // sub-step weather of main step `step` (the whole warp), single-site tables
    auto fill_weather = [&](int step) {
        if (!P.single_site) return;
        const double* w0 = P.weather + (int64_t)step * 3;
        if (P.weather_mode == HB_WEATHER_PER_SUBSTEP) w0 = P.weather + (int64_t)step * S * 3;
        for (int s = lane; s < S; s += 32) {
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
        __syncwarp();
    };
    // per-site weather (several weather tables in one batch): evaluated by the face lane itself
    auto site_weather = [&](const FaceC& cf, int step, int s, SubWeather& sw) {
        const double* wtab = P.weather + (int64_t)cf.site * P.weather_rows * 3;
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
    };
    // what the caller writes into SimulationState before a main step + the per-step derived quantities
    auto main_step_inputs = [&](int step) {
#pragma unroll
        for (int f = 0; f < NF; ++f) {
            const bool back = G == 1 ? f == 1 : bk;
            const int kf = back ? 1 : 0;
            if (P.input_mode != HB_INPUT_STATE) {
                const double* in = P.input_mode == HB_INPUT_SITE
                                       ? P.inputs + ((int64_t)c[f].site * P.n_steps + step) * 4
                                       : P.inputs + ((int64_t)step * P.n_desc_surfaces + c[f].desc_index) * 4;
                w[f].sol = in[kf];
                w[f].ir = in[2 + kf];
                if (!back && P.ir_mode == HB_IR_NET_GAIN) {
                    // validate_wall_heat_transfer.rs:689-699 (front face only, as in the reference driver)
                    if (c[f].eps > 1e-3) {
                        const double tk = w[f].ts + 273.15;
                        const double tk2 = tk * tk;
                        w[f].ir = w[f].ir / c[f].area / c[f].eps + kSigma * (tk2 * tk2);
                    } else {
                        w[f].ir = ir_state[f];
                    }
                }
            }
            face_begin_main_step(c[f], w[f]);
        }
    };
    // border conditions of sub-step s of main step `step` (old surface temperatures) → links for the node lanes
    WtLinks lk[NF];
    double zface = zT;   // temperature of the zone this lane's (back) face looks at
    if constexpr (!SIMPLE) zface = __shfl_sync(FULL, zT, zsrc);
    auto pre = [&](int step, int s) {
        double h_[NF], rad_[NF], trad_[NF], asol_[NF];
#pragma unroll
        for (int f = 0; f < NF; ++f) {
            const FaceC& cf = c[f];
            FaceS& wf = w[f];
            SubWeather sw;
            if (P.single_site) sw = s_w[s]; else site_weather(cf, step, s, sw);
            const int fl = cf.fl;
            const bool is_out = fl & FC_OUT;
            const bool leeward = (fl & FC_TILTED) && !(add_rn(mul_rn(cf.nx, sw.sin_wd), mul_rn(cf.ny, sw.cos_wd)) > 0.0);
            wf.forced = cf.forced_c * (leeward ? 0.5 : 1.0) * sw.sqrt_ws;
            wf.air = is_out ? sw.t_out : (fl & FC_SPACE) ? zface : cf.amb;
            trad_[f] = is_out ? wf.trad_out : wf.air;
            const double hn = tarp_lean(wf.air - wf.ts, cf.kpos, cf.kneg);
            h_[f] = is_out ? hn + wf.forced : hn;
            if (h_[f] != h_[f]) wf.status |= ST_NAN_H;
            const double xr = 273.15 + (trad_[f] + wf.ts) / 2.;
            rad_[f] = cf.eps4s * ((xr * xr) * xr);
            asol_[f] = cf.alpha_end * wf.solar;
        }
        if constexpr (PANE) {
            // A pane's two faces (first and last lane of the group; G = 1: the lane's two faces) exchange their border
            // conditions and both solve the wall's 2x2 no-mass system (solve_pane) with identical inputs; each keeps its node.
            if constexpr (G == 1) {
                const double sol0 = c[0].alpha_end * w[0].solar + c[0].alpha_x * w[1].solar;
                const double sol1 = c[1].alpha_end * w[1].solar + c[1].alpha_x * w[0].solar;
                if (c[0].fl & FC_PANE) {
                    double t0 = w[0].ts, t1 = w[1].ts;
                    unsigned long long it = 0;
                    solve_pane<true>(w[0].air, h_[0], rad_[0], trad_[0], sol0, w[1].air, h_[1], rad_[1], trad_[1], sol1,
                                     c[0].u_nb, t0, t1, it, w[0].status);
                    w[0].nm_iters += it;
                    w[0].ts = t0;
                    w[1].ts = t1;
                }
                if (win4_lane) {
                    const double sol[4] = {sol0, w_a1f * w[0].solar + w_a1b * w[1].solar,
                                           w_a2f * w[0].solar + w_a2b * w[1].solar, sol1};
                    double t[4] = {T[0], T[1], T[2], T[3]};
                    unsigned long long it = 0;
                    solve_window4(w[0].air, h_[0], rad_[0], trad_[0], w[1].air, h_[1], rad_[1], trad_[1], sol, c[0].u_nb,
                                  w_u23, w_cav, t, it, w[0].status);
                    w[0].nm_iters += it;
                    T[0] = t[0]; T[1] = t[1]; T[2] = t[2]; T[3] = t[3];
                    w[0].ts = t[0];
                    w[1].ts = t[3];
                }
            } else {
                const FaceC& cf = c[0];
                FaceS& wf = w[0];
                const int px = G - 1;   // partner lane = lane ^ (G - 1) for the first and the last lane of a group
                const double p_solar = __shfl_xor_sync(FULL, wf.solar, px);
                const double sol_own = cf.alpha_end * wf.solar + cf.alpha_x * p_solar;
                const double p_air = __shfl_xor_sync(FULL, wf.air, px), p_h = __shfl_xor_sync(FULL, h_[0], px);
                const double p_rad = __shfl_xor_sync(FULL, rad_[0], px), p_trad = __shfl_xor_sync(FULL, trad_[0], px);
                const double p_ts = __shfl_xor_sync(FULL, wf.ts, px), p_sol = __shfl_xor_sync(FULL, sol_own, px);
                if (cf.fl & FC_PANE) {
                    double t0 = bk ? p_ts : wf.ts, t1 = bk ? wf.ts : p_ts;
                    unsigned long long it = 0;
                    solve_pane<true>(bk ? p_air : wf.air, bk ? p_h : h_[0], bk ? p_rad : rad_[0], bk ? p_trad : trad_[0],
                                     bk ? p_sol : sol_own, bk ? wf.air : p_air, bk ? h_[0] : p_h, bk ? rad_[0] : p_rad,
                                     bk ? trad_[0] : p_trad, bk ? sol_own : p_sol, cf.u_nb, t0, t1, it, wf.status);
                    if (!bk) wf.nm_iters += it;   // the chunk's iterations are counted once
                    wf.ts = bk ? t1 : t0;
                }
                if (win4_lane) {   // front lane: own values are the front face's, the partner's the back face's
                    const double sol[4] = {sol_own, w_a1f * wf.solar + w_a1b * p_solar, w_a2f * wf.solar + w_a2b * p_solar,
                                           p_sol};
                    double t[4] = {T[0], T[1], T[2], T[3]};
                    unsigned long long it = 0;
                    solve_window4(wf.air, h_[0], rad_[0], trad_[0], p_air, p_h, p_rad, p_trad, sol, cf.u_nb, w_u23, w_cav, t,
                                  it, wf.status);
                    wf.nm_iters += it;
                    T[0] = t[0]; T[1] = t[1]; T[2] = t[2]; T[3] = t[3];
                    wf.ts = t[0];
                }
            }
        }
#pragma unroll
        for (int f = 0; f < NF; ++f) {
            const bool back = G == 1 ? f == 1 : bk;
            const FaceC& cf = c[f];
            FaceS& wf = w[f];
            const int fl = cf.fl;
            const double h = h_[f], rad = rad_[f], trad = trad_[f], asol = asol_[f];
            lk[f].t_in = 0.;
            // (1/h does not wait for the no-mass solve below: started here, it overlaps with it)
            const double hs = h > 1e-12 ? h : 1e-12;
            const double inv_hs = rcp_lean(hs);
            if constexpr (END) {
                if (fl & FC_NM2) {
                    double t_in = back ? t1_b : t1_f;
                    solve_end_pair<true>(back, wf.air, h, rad, trad, asol, cf.u_nb, cf.u_nb2, back ? t2_b : t2_f, wf.ts,
                                         t_in, wf.nm_iters, wf.status);
                    lk[f].t_in = t_in;
                } else if (fl & FC_NM) {
                    wf.ts = solve_end_node<true>(wf.air, h, rad, trad, cf.u_nb, back ? t1_b : t1_f, asol, wf.ts,
                                                 wf.nm_iters, wf.status);
                }
            }
            // a no-mass end node takes no film link (v_end = 0): its temperature is placed, the ghost is never read
            lk[f].ghost = wf.air + (rad * (trad - wf.ts) + asol) * inv_hs;
            lk[f].link = hs * cf.v_end;
        }
    };

#ifdef HB_PHASE_CLOCKS
    long long wclk[8] = {0, 0, 0, 0, 0, 0, 0, 0}, wlast = clock64();
#endif
    int rec = P.record_every > 0 ? P.step_begin / P.record_every : 0;
    for (int step = P.step_begin; step < P.step_end; ++step) {
        fill_weather(step);
        main_step_inputs(step);
        // scheduled zone inputs: this main step's values, sources re-summed in the reference's order
        // (the zone's constants are re-read here rather than kept in registers over the series: 14 registers)
        if (szone && (P.zone_gain || P.zone_inf_v || P.zone_inf_t || P.zone_vent_v || P.zone_vent_t)) {
            ZoneC zs;
            zs.vol = V.zvol[sgz]; zs.q = V.zq[sgz]; zs.inf_v = V.zinf_v[sgz]; zs.inf_t = V.zinf_t[sgz];
            zs.vent_v = V.zvent_v[sgz]; zs.vent_t = V.zvent_t[sgz]; zs.sb = 0; zs.se = 0;
            load_zone_schedule(P, step, sgz, zs);
            zone_sources(zs, za0, zb0);
        }
        for (int s = 0; s < S; ++s) {
#ifdef HB_PHASE_CLOCKS
            wlast = clock64();
            wclk[7] += 1;
#endif
            // ================= face role: border conditions with the old surface temperatures =================
            pre(step, s);
            HB_WCLK(0)
            // ================= face lanes → node lanes =================
            const double gf = lk[0].ghost, lf = lk[0].link;
            double gb = lk[NF - 1].ghost, lb = lk[NF - 1].link;
            double tf0 = w[0].ts, tf1 = lk[0].t_in, tb0 = w[NF - 1].ts, tb1 = lk[NF - 1].t_in;
            if constexpr (G > 1) {
                gb = __shfl_sync(FULL, gb, base + G - 1);
                lb = __shfl_sync(FULL, lb, base + G - 1);
                if constexpr (END) {
                    tb0 = __shfl_sync(FULL, tb0, base + G - 1);
                    tb1 = __shfl_sync(FULL, tb1, base + G - 1);
                }
            }
            if constexpr (CSM) {
                if (first) ab[0].y = lf;
                if (last >= 0 && last < NL) ab[last * 32].x = lb;
            } else {
                B[0] = first ? lf : B[0];
            }
            if constexpr (END) {
                T[0] = (first && nmf) ? tf0 : T[0];
                T[1] = (first && nmf2) ? tf1 : T[1];
            }
#pragma unroll
            for (int i = 0; i < NL; ++i) {
                if constexpr (!CSM) A[i] = i == last ? lb : A[i];
                if constexpr (END) {
                    T[i] = (i == last && nmb) ? tb0 : T[i];
                    T[i] = (i == last - 1 && nmb2) ? tb1 : T[i];
                }
                T[i] = i == last + 1 ? gb : T[i];   // the slot behind the last node stands for the back ghost
            }
            HB_WCLK(1)
            // ================= node role: RK4 of the massive nodes =================
            if constexpr (CSM) wt_rk4_smem<G, NL>(T, ab, first, gf, edge, gb);
            else wt_rk4<G, NL>(T, A, B, first, gf, edge, gb);
            HB_WCLK(2)
            // ================= node lanes → face lanes =================
            ts_f = T[0];
            if constexpr (END) {
                t1_f = T[1];
                t2_f = T[2];
            }
            gather_back();
            HB_WCLK(3)
            // ================= face role: new film coefficients, flows, zone =================
            double hA_b = 0., hAT_b = 0.;
#pragma unroll
            for (int f = 0; f < NF; ++f) {
                const bool back = G == 1 ? f == 1 : bk;
                const FaceC& cf = c[f];
                FaceS& wf = w[f];
                const double ts = back ? ts_b : ts_f;
                const double hn = tarp_lean(wf.air - ts, cf.kpos, cf.kneg);
                const double h = (cf.fl & FC_OUT) ? hn + wf.forced : hn;
                if (h != h) wf.status |= ST_NAN_H;
                wf.ts = ts;
                wf.h = h;
                wf.flow = (ts - wf.air) * h;
                wf.e += wf.flow * cf.area * dt;
                if (back && (SIMPLE || (cf.fl & FC_SPACE))) {
                    hA_b = h * cf.area;
                    hAT_b = hA_b * ts;
                }
            }
            HB_WCLK(4)
            if constexpr (SIMPLE) {   // the lane's own wall is the zone's only surface (its back face): model.rs:489-597, 650-674
                const double inv_cc = zone_inv_c_t<true>(zvol, zT);
                const double tn = zone_advance<true>(zT, dt, inv_cc, za0 + hAT_b, zb0 + hA_b);
                if (szone) {
                    if (tn != tn) atomicOr(&V.status[sg], ST_NAN_ZONE);
                    zT = tn;
                }
                zface = zT;
            } else {
                // the back faces of a model's walls [sb, se) sit on every G-th lane: all zones of the tile are summed at once
                // by a segmented shuffle reduction (deterministic; differs from the reference's sequential order,
                // model.rs:556-590, by rounding only), advanced by the back lane of wall sb and broadcast to the model's faces
                double va = hAT_b, vb = hA_b;
#pragma unroll
                for (int o = 1; o < W; o <<= 1) {
                    const double ta = __shfl_down_sync(FULL, va, o * G), tb = __shfl_down_sync(FULL, vb, o * G);
                    if (wl + o < seg_end) {
                        va += ta;
                        vb += tb;
                    }
                }
                const double inv_cc = zone_inv_c_t<true>(zvol, zT);
                const double tn = zone_advance<true>(zT, dt, inv_cc, za0 + va, zb0 + vb);
                if (szone) {
                    if (tn != tn) atomicOr(&V.status[sg], ST_NAN_ZONE);
                    zT = tn;
                }
                zface = __shfl_sync(FULL, zT, zsrc);
            }
            HB_WCLK(5)
        }

        if (P.record_every > 0 && (step + 1) % P.record_every == 0) {
            if (P.zone_rec && szone) P.zone_rec[(int64_t)rec * V.n_dev_zones + sgz] = zT;
            if (P.node_rec) {
#pragma unroll
                for (int i = 0; i < NL; ++i)
                    P.node_rec[(int64_t)rec * P.slab_len + tile.node_offset + (int64_t)i * 32 + lane] = T[i];
                if (fwall) {
                    double* rr = P.sc_rec + (int64_t)rec * 8 * NS;
#pragma unroll
                    for (int f = 0; f < NF; ++f) {
                        const int kf = (G == 1 ? f == 1 : bk) ? 1 : 0;
                        rr[(0 + kf) * NS + sg] = w[f].h; rr[(2 + kf) * NS + sg] = w[f].flow;
                        rr[(4 + kf) * NS + sg] = w[f].sol; rr[(6 + kf) * NS + sg] = w[f].ir;
                    }
                }
            }
            ++rec;
        }
    }

#ifdef HB_PHASE_CLOCKS
    if (lane == 0 && tile_slot == 0)
        for (int q = 0; q < 8; ++q) atomicAdd(&hb_wt_clk[q], (unsigned long long)wclk[q]);
#endif
    // ---- write the state back ----
#pragma unroll
    for (int i = 0; i < NL; ++i) V.nodeT[tile.node_offset + (int64_t)i * 32 + lane] = T[i];
    if (fwall) {
#pragma unroll
        for (int f = 0; f < NF; ++f) {
            const bool back = G == 1 ? f == 1 : bk;
            const int kf = back ? 1 : 0;
            V.sc[(0 + kf) * NS + sg] = w[f].h; V.sc[(2 + kf) * NS + sg] = w[f].flow;
            V.sc[(4 + kf) * NS + sg] = w[f].sol; V.sc[(6 + kf) * NS + sg] = w[f].ir;
            if (back) V.energy_b[sg] += w[f].e; else V.energy_f[sg] += w[f].e;
            if (w[f].nm_iters) atomicAdd(&V.nm_iters[sg], w[f].nm_iters);
            if (w[f].status) atomicOr(&V.status[sg], w[f].status);
        }
    }
    if (szone) V.zoneT[sgz] = zT;
}

template <int G, int NL, int MODE, bool SIMPLE>
__global__ void __launch_bounds__(32 * kWtWarpsMax) __maxnreg__(wt_max_regs(G, NL, MODE))
    hb_wtile_kernel(DevView V, SeriesDev P, const int32_t* tile_ids, int n_tiles, int s_rows) {
    extern __shared__ double s_dyn[];
    if (P.started != nullptr && threadIdx.x == 0) {   // tell the host this block has its SM (launch_series waits for it)
        unsigned long long ns;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(ns));
        unsigned smid;
        asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
        // non-zero; SM id and a ~µs stamp for HB_GROUP_TIMELINE
        P.started[blockIdx.x] = (int32_t)(((ns >> 10) & 0xfffffull) | ((unsigned long long)(smid & 0x3ff) << 20) | (1ull << 30));
        __threadfence_system();
    }
    const int tile_slot = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (tile_slot < n_tiles) wt_march_tile<G, NL, MODE, SIMPLE>(V, P, tile_ids, tile_slot, s_rows, s_dyn);   // whole warp
    // The launches that take an SM per block run as clusters of two blocks — the two SMs of one TPC (hb_wtile_tier.cu):
    // a warp of such a block that shares its TPC with a bulk-loaded SM runs a third slower (measured: 3170 against 2370
    // cycles per sub-step).  A block that is done (or the one that pads an odd grid) keeps its SM until its partner is
    // done too.  Every other launch is a cluster of one: the barrier falls through.
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

}  // namespace hb

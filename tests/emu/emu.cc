// emu.cc — TEST INFRASTRUCTURE: a sequential host re-enactment of hb_march_kernel's data flow.
//
// It calls the very same HB_HD functions the CUDA kernel calls (heat_b200/csrc/hb_math.cuh) on the very same packed
// buffers (hb::pack_arrays), with the block barriers replaced by phase loops and the warp shuffles by array reads.
// Purpose: debug the device arithmetic and the packing against the oracle on a machine without a GPU.  It is built
// and used by tests/ only; the product (heat_b200/) never links or calls it and has no CPU path.
#include <cmath>
#include <cstring>
#include <stdexcept>
#include <string>
#include <type_traits>
#include <vector>

#include "hb_math.cuh"

using namespace hb;

namespace {
std::string g_err;
struct WallLinksEmu { double ghost, link, t_in; };   // hb_wtile.cuh: WtLinks

void march_tile(const DevView& V, const SeriesDev& P, const Tile& tile, double* zone_rec_base) {
    const int G = tile.G, W = tile.W, TPB = G * W, NL = kNL, ROW = G * NL + 1;
    const bool EXT = tile.flags & TILE_EXT, DIST = tile.flags & TILE_DIST, has_cav = tile.flags & TILE_CAV;
    const int S = tile.S;
    const double dt = tile.dt;
    const int64_t NS = V.n_dev_surfaces;
    std::vector<double> T(TPB * NL), A(TPB * NL), B(TPB * NL), Vv(TPB * NL), Q(TPB * NL, 0.);
    std::vector<double> sT(W * ROW, 0.), sA(W * ROW, 0.), sB(W * ROW, 0.);
    std::vector<WallC> wc(W);
    std::vector<ZoneC> zc(W);
    std::vector<WallS> ws(W);
    std::vector<ExtWall> xs(W);
    std::vector<double> zT(W, 0.), pub0(W, 0.), pub1(W, 0.), pub2(W, 0.), pub3(W, 0.), ts0(W, 0.), ts1(W, 0.);
    std::vector<int32_t> zf(W), zb(W);
    std::vector<BoundaryLinks> bc(W, BoundaryLinks{0., 0., 0., 0.});
    std::vector<SubWeather> swt(kMaxSubsteps);
    std::vector<double> ir_state(W, 0.);

    auto rowof = [&](int tid) { return (tid / G) * ROW + tid % G; };
    for (int tid = 0; tid < TPB; ++tid)
        for (int i = 0; i < NL; ++i) {
            const int64_t o = tile.node_offset + (int64_t)i * TPB + tid;
            T[tid * NL + i] = V.nodeT[o];
            A[tid * NL + i] = V.coefA[o];
            B[tid * NL + i] = V.coefB[o];
            sT[rowof(tid) + i * G] = V.nodeT[o];
            sA[rowof(tid) + i * G] = V.coefA[o];
            sB[rowof(tid) + i * G] = V.coefB[o];
        }
    for (int t = 0; t < W; ++t) {
        const int64_t sg = (int64_t)tile.surface_begin + t;
        WallC c;
        load_wallc(V, sg, P.single_site != 0, c);
        wc[t] = c;
        zf[t] = c.zf;
        zb[t] = c.zb;
        const int64_t gz = (int64_t)tile.zone_begin + t;
        ZoneC z{};
        if (t < tile.n_zones) {
            z.vol = V.zvol[gz]; z.q = V.zq[gz]; z.inf_v = V.zinf_v[gz]; z.inf_t = V.zinf_t[gz];
            z.vent_v = V.zvent_v[gz]; z.vent_t = V.zvent_t[gz]; z.sb = V.z_surf_begin[gz]; z.se = V.z_surf_end[gz];
            zT[t] = V.zoneT[gz];
        }
        zc[t] = z;
    }
    auto publish_ends = [&]() {
        for (int tid = 0; tid < TPB; ++tid) {
            const int wl = tid / G, part = tid % G;
            const int n = V.n_nodes[(int64_t)tile.surface_begin + wl];
            if (part == 0) ts0[wl] = T[tid * NL + 1];
            for (int i = 0; i < NL; ++i)
                if (i == n - part * NL) ts1[wl] = T[tid * NL + i];
        }
    };
    publish_ends();
    for (int t = 0; t < W; ++t) {
        const int64_t sg = (int64_t)tile.surface_begin + t;
        WallS w{};
        if (wc[t].n > 0) {
            w.h_f = V.sc[0 * NS + sg]; w.h_b = V.sc[1 * NS + sg]; w.flow_f = V.sc[2 * NS + sg]; w.flow_b = V.sc[3 * NS + sg];
            w.sol_f = V.sc[4 * NS + sg]; w.sol_b = V.sc[5 * NS + sg]; w.ir_f = V.sc[6 * NS + sg]; w.ir_b = V.sc[7 * NS + sg];
            w.ts_f = ts0[t];
            w.ts_b = ts1[t];
        }
        ir_state[t] = w.ir_f;
        ws[t] = w;
        ExtWall X{};
        if (EXT) {
            X.t = sT.data() + t * ROW; X.a = sA.data() + t * ROW; X.b = sB.data() + t * ROW;
            X.u = V.ext_u + tile.ext_offset + t; X.mass = V.ext_mass + tile.ext_offset + t;
            X.af = V.ext_af + tile.ext_offset + t; X.ab = V.ext_ab + tile.ext_offset + t;
            X.G = G; X.W = W;
        }
        xs[t] = X;
    }

    auto weather_at = [&](int site, int step, int s, SubWeather& out) {
        const double* wtab = P.weather + (int64_t)site * P.weather_rows * 3;
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
        sub_weather(t_out, wdir, wspeed, out);
    };

    int rec = 0;
    std::vector<double> left(TPB), right(TPB);
    for (int step = 0; step < P.n_steps; ++step) {
        for (int t = 0; t < tile.n_zones; ++t)   // scheduled zone inputs of this main step
            load_zone_schedule(P, step, (int64_t)tile.zone_begin + t, zc[t]);
        if (P.single_site)
            for (int s = 0; s < S; ++s) weather_at(0, step, s, swt[s]);
        for (int t = 0; t < W; ++t) {
            const WallC& c = wc[t];
            if (c.n == 0) continue;
            WallS& w = ws[t];
            if (P.input_mode == HB_INPUT_SITE) {
                const double* in = P.inputs + ((int64_t)c.site * P.n_steps + step) * 4;
                w.sol_f = in[0]; w.sol_b = in[1]; w.ir_f = in[2]; w.ir_b = in[3];
            } else if (P.input_mode == HB_INPUT_PER_SURFACE) {
                const double* in = P.inputs + ((int64_t)step * P.n_desc_surfaces + c.desc_index) * 4;
                w.sol_f = in[0]; w.sol_b = in[1]; w.ir_f = in[2]; w.ir_b = in[3];
            }
            if (P.input_mode != HB_INPUT_STATE && P.ir_mode == HB_IR_NET_GAIN) {
                if (c.eps_f > 1e-3) {
                    const double tk = w.ts_f + 273.15;
                    const double tk2 = tk * tk;
                    w.ir_f = w.ir_f / c.area / c.eps_f + kSigma * (tk2 * tk2);
                } else {
                    w.ir_f = ir_state[t];
                }
            }
            begin_main_step(c, w);
        }
        if (DIST)
            for (int tid = 0; tid < TPB; ++tid) {
                const int wl = tid / G, part = tid % G;
                const int64_t g = (int64_t)tile.surface_begin + wl;
                const int n = V.n_nodes[g];
                const bool wall_dist = (V.flags[g] >> 8) & 1;
                for (int i = 0; i < NL; ++i) {
                    Q[tid * NL + i] = 0.;
                    const int node = part * NL + i - 1;
                    if (wall_dist && node >= 0 && node < n) {
                        const int64_t e = tile.ext_offset + (int64_t)node * W + wl;
                        const double m = V.ext_mass[e];
                        const double vi = m >= kMassiveThreshold ? dt / m : 0.;
                        Q[tid * NL + i] = vi * (V.ext_af[e] * ws[wl].solar_front + V.ext_ab[e] * ws[wl].solar_back);
                    }
                }
            }
        for (int s = 0; s < S; ++s) {
            for (int t = 0; t < W; ++t) {
                const WallC& c = wc[t];
                if (c.n == 0) continue;
                SubWeather sw;
                if (P.single_site) sw = swt[s];
                else weather_at(c.site, step, s, sw);
                const double ztf = c.zf >= 0 ? zT[c.zf] : 0.;
                const double ztb = c.zb >= 0 ? zT[c.zb] : 0.;
                if (EXT) scalar_pre<true>(c, ws[t], sw, ztf, ztb, dt, xs[t], V.cav, V.nm_pairs, bc[t]);
                else scalar_pre<false>(c, ws[t], sw, ztf, ztb, dt, xs[t], V.cav, V.nm_pairs, bc[t]);
            }
            // ---- node role ----
            for (int tid = 0; tid < TPB; ++tid) {
                const int wl = tid / G, part = tid % G;
                const int n = V.n_nodes[(int64_t)tile.surface_begin + wl];
                double* t = &T[tid * NL];
                if (EXT) {
                    for (int i = 0; i < NL; ++i) t[i] = sT[rowof(tid) + i * G];
                    if (has_cav)
                        for (int i = 0; i < NL; ++i) {
                            A[tid * NL + i] = sA[rowof(tid) + i * G];
                            B[tid * NL + i] = sB[rowof(tid) + i * G];
                        }
                }
                const BoundaryLinks bl = bc[wl];
                if (part == 0) {
                    t[0] = bl.ghost_f;
                    B[tid * NL + 1] = bl.bf;
                }
                for (int i = 0; i < NL; ++i) {
                    if (i == n - part * NL) A[tid * NL + i] = bl.ab;
                    if (i == n + 1 - part * NL) t[i] = bl.ghost_b;
                }
            }
            auto halos = [&](const std::vector<double>& x) {
                for (int tid = 0; tid < TPB; ++tid) {
                    const int part = tid % G;
                    left[tid] = part > 0 ? x[(tid - 1) * NL + NL - 1] : x[tid * NL + NL - 1];
                    right[tid] = part < G - 1 ? x[(tid + 1) * NL] : x[tid * NL];
                }
            };
            using Arr = double(&)[kNL];
            halos(T);
            for (int tid = 0; tid < TPB; ++tid) {
                sweep_first<kNL>((Arr)T[tid * NL], (Arr)Vv[tid * NL], (Arr)A[tid * NL], (Arr)B[tid * NL], left[tid], right[tid]);
                for (int i = 0; i < NL; ++i) {
                    if (DIST) Vv[tid * NL + i] += Q[tid * NL + i];
                    T[tid * NL + i] += Vv[tid * NL + i];
                }
            }
            const double cs[3] = {1. / 2., 1. / 6., 1. / 24.};
            for (int k = 0; k < 3; ++k) {
                halos(Vv);
                for (int tid = 0; tid < TPB; ++tid)
                    sweep_next<kNL>((Arr)T[tid * NL], (Arr)Vv[tid * NL], (Arr)A[tid * NL], (Arr)B[tid * NL], left[tid],
                                    right[tid], cs[k]);
            }
            if (EXT) {
                for (int tid = 0; tid < TPB; ++tid)
                    for (int i = 0; i < NL; ++i) sT[rowof(tid) + i * G] = T[tid * NL + i];
            } else {
                publish_ends();
            }
            // ---- scalar role, after ----
            for (int t = 0; t < W; ++t) {
                const WallC& c = wc[t];
                if (c.n == 0) continue;
                double ts_f, ts_b;
                if (EXT) {
                    ts_f = xs[t].T(0);
                    ts_b = xs[t].T(c.n - 1);
                } else {
                    ts_f = ts0[t];
                    ts_b = ts1[t];
                }
                double pub[4];
                scalar_post(c, ws[t], ts_f, ts_b, dt, pub);
                pub0[t] = pub[0]; pub1[t] = pub[1]; pub2[t] = pub[2]; pub3[t] = pub[3];
            }
            std::vector<double> znew(zT);
            for (int t = 0; t < tile.n_zones; ++t) {
                const double tn = zone_update(zc[t], t, zT[t], dt, pub0.data(), pub1.data(), pub2.data(), pub3.data(),
                                              zf.data(), zb.data());
                if (tn != tn) V.status[(int64_t)tile.surface_begin + zc[t].sb] |= ST_NAN_ZONE;
                znew[t] = tn;
            }
            zT = znew;
        }
        if (P.record_every > 0 && (step + 1) % P.record_every == 0) {
            if (zone_rec_base)
                for (int t = 0; t < tile.n_zones; ++t)
                    zone_rec_base[(int64_t)rec * V.n_dev_zones + tile.zone_begin + t] = zT[t];
            if (P.node_rec) {
                for (int tid = 0; tid < TPB; ++tid)
                    for (int i = 0; i < NL; ++i)
                        P.node_rec[(int64_t)rec * P.slab_len + tile.node_offset + (int64_t)i * TPB + tid] = T[tid * NL + i];
                for (int t = 0; t < W; ++t) {
                    if (wc[t].n == 0) continue;
                    const int64_t sg = (int64_t)tile.surface_begin + t;
                    const WallS& w = ws[t];
                    double* r = P.sc_rec + (int64_t)rec * 8 * NS;
                    r[0 * NS + sg] = w.h_f; r[1 * NS + sg] = w.h_b; r[2 * NS + sg] = w.flow_f; r[3 * NS + sg] = w.flow_b;
                    r[4 * NS + sg] = w.sol_f; r[5 * NS + sg] = w.sol_b; r[6 * NS + sg] = w.ir_f; r[7 * NS + sg] = w.ir_b;
                }
            }
            ++rec;
        }
    }
    for (int tid = 0; tid < TPB; ++tid)
        for (int i = 0; i < NL; ++i) V.nodeT[tile.node_offset + (int64_t)i * TPB + tid] = T[tid * NL + i];
    for (int t = 0; t < W; ++t) {
        if (wc[t].n == 0) continue;
        const int64_t sg = (int64_t)tile.surface_begin + t;
        const WallS& w = ws[t];
        V.sc[0 * NS + sg] = w.h_f; V.sc[1 * NS + sg] = w.h_b; V.sc[2 * NS + sg] = w.flow_f; V.sc[3 * NS + sg] = w.flow_b;
        V.sc[4 * NS + sg] = w.sol_f; V.sc[5 * NS + sg] = w.sol_b; V.sc[6 * NS + sg] = w.ir_f; V.sc[7 * NS + sg] = w.ir_b;
        V.energy_f[sg] += w.e_f;
        V.energy_b[sg] += w.e_b;
        V.nm_iters[sg] += w.nm_iters;
        V.status[sg] |= w.status;
    }
    for (int t = 0; t < tile.n_zones; ++t) V.zoneT[(int64_t)tile.zone_begin + t] = zT[t];
}

// hb_wtile_kernel's data flow (hb_wtile.cuh): one warp = one tile, G lanes per wall, NL NODES per lane (no ghost slots),
// front face on the first lane of a group, back face + zone on the last; shuffles become array reads.
void march_tile_warp(const DevView& V, const SeriesDev& P, const Tile& tile, double* zone_rec_base) {
    const int G = tile.G, NL = tile.NL, NF = G == 1 ? 2 : 1;
    const bool END = tile.flags & TILE_END;
    const int S = tile.S;
    const double dt = tile.dt;
    const int64_t NS = V.n_dev_surfaces;
    const bool SIMPLE = tile.flags & TILE_SIMPLE, PANE = !SIMPLE;   // buildings run the pane-capable kernel
    const int W = 32 / G;
    if (G * tile.W != 32) throw std::runtime_error("emu: warp tile shape");
    std::vector<double> T(32 * NL), A(32 * NL), B(32 * NL), X(32 * NL);
    struct Lane {
        int wl, part, base, n, last, nm1, nm2, nm3;
        bool first, edge, bk, face_lane, fwall, szone, nmf, nmb, nmf2, nmb2;
        int64_t sg, sgz;
        FaceC c[2];
        FaceS w[2];
        double ir_state[2];
        WallLinksEmu lk[2];
        double zT, za0, zb0, zvol, zface;
        ZoneC zc;
        double ts_f, t1_f, t2_f, ts_b, t1_b, t2_b;
        int zown, seg_end, zsrc;
        double h_[2], rad_[2], trad_[2], asol_[2];
        bool win4_lane;
        double w_a1f, w_a1b, w_a2f, w_a2b, w_u23;
        CavityDev w_cav;
    };
    std::vector<Lane> L(32);
    std::vector<SubWeather> swt(kMaxSubsteps);
    for (int lane = 0; lane < 32; ++lane)
        for (int i = 0; i < NL; ++i) {
            const int64_t o = tile.node_offset + (int64_t)i * 32 + lane;
            T[lane * NL + i] = V.nodeT[o];
            A[lane * NL + i] = V.coefA[o];
            B[lane * NL + i] = V.coefB[o];
        }
    auto gather_back = [&]() {
        for (int lane = 0; lane < 32; ++lane) {
            Lane& l = L[lane];
            l.ts_f = T[lane * NL + 0]; l.t1_f = T[lane * NL + 1]; l.t2_f = T[lane * NL + 2];
            l.ts_b = T[(l.base + l.nm1 / NL) * NL + l.nm1 % NL];
            l.t1_b = T[(l.base + l.nm2 / NL) * NL + l.nm2 % NL];
            l.t2_b = T[(l.base + l.nm3 / NL) * NL + l.nm3 % NL];
        }
    };
    for (int lane = 0; lane < 32; ++lane) {
        Lane& l = L[lane];
        l.wl = lane / G; l.part = lane % G; l.base = lane - l.part;
        l.sg = (int64_t)tile.surface_begin + l.wl;
        l.n = V.n_nodes[l.sg];
        l.last = l.n - 1 - l.part * NL;
        l.first = l.part == 0; l.edge = l.last == NL - 1;
        l.nm1 = l.n > 0 ? l.n - 1 : 0; l.nm2 = l.n > 1 ? l.n - 2 : 0; l.nm3 = l.n > 2 ? l.n - 3 : 0;
        const int wflags = V.flags[l.sg];
        l.nmf = END && ((wflags >> 10) & 1); l.nmb = END && ((wflags >> 11) & 1);
        l.nmf2 = END && ((wflags >> 13) & 1); l.nmb2 = END && ((wflags >> 14) & 1);
        l.bk = l.part == G - 1;
        l.face_lane = l.first || l.bk;
        l.fwall = l.face_lane && l.n > 0;
        l.win4_lane = PANE && l.first && l.n == 4 && ((wflags >> 15) & 1);
        l.w_a1f = l.w_a1b = l.w_a2f = l.w_a2b = l.w_u23 = 0.;
        l.w_cav = CavityDev{0, 0, 1., 1., 0., 1., 1.};
        if (l.win4_lane) {
            l.w_a1f = V.wa1f[l.sg]; l.w_a1b = V.wa1b[l.sg]; l.w_a2f = V.wa2f[l.sg]; l.w_a2b = V.wa2b[l.sg];
            l.w_u23 = V.u_b[l.sg];
            l.w_cav = V.cav[V.cav_off[l.sg]];
        }
        l.zown = SIMPLE ? l.wl : -1; l.seg_end = l.wl + 1; l.zsrc = lane;
        if (!SIMPLE)
            for (int j = 0; j < tile.n_zones; ++j) {
                const int sb = V.z_surf_begin[tile.zone_begin + j], se = V.z_surf_end[tile.zone_begin + j];
                if (sb == l.wl) l.zown = j;
                if (l.wl >= sb && l.wl < se) { l.seg_end = se; l.zsrc = sb * G + G - 1; }
            }
        l.szone = l.bk && l.zown >= 0 && l.zown < tile.n_zones;
        l.sgz = (int64_t)tile.zone_begin + (l.zown >= 0 ? l.zown : 0);
        l.zT = 0.; l.za0 = 0.; l.zb0 = 0.; l.zvol = 1.;
        l.zc = ZoneC{};
        if (l.szone) {
            ZoneC& z = l.zc;
            const int64_t gz = l.sgz;
            z.vol = V.zvol[gz]; z.q = V.zq[gz]; z.inf_v = V.zinf_v[gz]; z.inf_t = V.zinf_t[gz];
            z.vent_v = V.zvent_v[gz]; z.vent_t = V.zvent_t[gz]; z.sb = V.z_surf_begin[gz]; z.se = V.z_surf_end[gz];
            l.zT = V.zoneT[gz];
            zone_sources(z, l.za0, l.zb0);
            l.zvol = z.vol;
            if (SIMPLE ? !(z.sb == l.wl && z.se == l.wl + 1) : !(z.sb == l.wl && z.se <= W))
                throw std::runtime_error("emu: warp tile zone topology");
        }
        for (int f = 0; f < NF; ++f) {
            const bool back = G == 1 ? f == 1 : l.bk;
            const int kf = back ? 1 : 0;
            l.w[f] = FaceS{};
            l.c[f] = FaceC{};
            if (l.fwall) {
                WallC wc;
                load_wallc(V, l.sg, P.single_site != 0, wc);
                const bool pane = PANE && ((wflags >> 12) & 1);
                const bool win4 = PANE && ((wflags >> 15) & 1);
                const bool nm = END && !pane && ((wflags >> (back ? 11 : 10)) & 1);
                const bool nm2f = nm && ((wflags >> (back ? 14 : 13)) & 1);
                load_facec(wc, back, back ? V.u_b[l.sg] : V.u_f[l.sg], nm, l.c[f], pane, back ? V.abL[l.sg] : V.af0[l.sg],
                           back ? V.ax_b[l.sg] : V.ax_f[l.sg], nm2f, back ? V.u_b2[l.sg] : V.u_f2[l.sg], win4);
                l.w[f].h = V.sc[(0 + kf) * NS + l.sg]; l.w[f].flow = V.sc[(2 + kf) * NS + l.sg];
                l.w[f].sol = V.sc[(4 + kf) * NS + l.sg]; l.w[f].ir = V.sc[(6 + kf) * NS + l.sg];
            }
            l.ir_state[f] = l.w[f].ir;
        }
    }
    gather_back();
    for (Lane& l : L)
        for (int f = 0; f < NF; ++f) l.w[f].ts = (G == 1 ? f == 1 : l.bk) ? l.ts_b : l.ts_f;
    for (Lane& l : L) l.zface = SIMPLE ? l.zT : L[l.zsrc].zT;

    auto weather_at = [&](int site, int step, int s, SubWeather& out) {
        const double* wtab = P.weather + (int64_t)site * P.weather_rows * 3;
        const double* w0 = wtab + (int64_t)step * 3;
        double t_out, wdir, wspeed;
        if (P.weather_mode == HB_WEATHER_LINEAR) {
            const double f = (double)(s + 1) / (double)S;
            t_out = lerp_rn(w0[0], w0[3], f); wdir = lerp_rn(w0[1], w0[4], f); wspeed = lerp_rn(w0[2], w0[5], f);
        } else if (P.weather_mode == HB_WEATHER_PER_SUBSTEP) {
            w0 = wtab + ((int64_t)step * S + s) * 3;
            t_out = w0[0]; wdir = w0[1]; wspeed = w0[2];
        } else {
            t_out = w0[0]; wdir = w0[1]; wspeed = w0[2];
        }
        sub_weather(t_out, wdir, wspeed, out);
    };
    std::vector<double> left(32), right(32);
    int rec = 0;
    for (int step = 0; step < P.n_steps; ++step) {
        if (P.single_site)
            for (int s = 0; s < S; ++s) weather_at(0, step, s, swt[s]);
        for (Lane& l : L) {
            for (int f = 0; f < NF; ++f) {
                const bool back = G == 1 ? f == 1 : l.bk;
                const int kf = back ? 1 : 0;
                const FaceC& c = l.c[f];
                FaceS& w = l.w[f];
                if (P.input_mode != HB_INPUT_STATE) {
                    const double* in = P.input_mode == HB_INPUT_SITE
                                           ? P.inputs + ((int64_t)c.site * P.n_steps + step) * 4
                                           : P.inputs + ((int64_t)step * P.n_desc_surfaces + c.desc_index) * 4;
                    w.sol = in[kf];
                    w.ir = in[2 + kf];
                    if (!back && P.ir_mode == HB_IR_NET_GAIN) {
                        if (c.eps > 1e-3) {
                            const double tk = w.ts + 273.15;
                            const double tk2 = tk * tk;
                            w.ir = w.ir / c.area / c.eps + kSigma * (tk2 * tk2);
                        } else {
                            w.ir = l.ir_state[f];
                        }
                    }
                }
                face_begin_main_step(c, w);
            }
            if (l.szone && load_zone_schedule(P, step, l.sgz, l.zc)) zone_sources(l.zc, l.za0, l.zb0);
        }
        for (int s = 0; s < S; ++s) {
            // face role: border conditions with the old surface temperatures
            for (Lane& l : L)
                for (int f = 0; f < NF; ++f) {
                    const FaceC& cf = l.c[f];
                    FaceS& wf = l.w[f];
                    SubWeather sw;
                    if (P.single_site) sw = swt[s]; else weather_at(cf.site, step, s, sw);
                    const int fl = cf.fl;
                    const bool is_out = fl & FC_OUT;
                    const bool leeward = (fl & FC_TILTED) && !(add_rn(mul_rn(cf.nx, sw.sin_wd), mul_rn(cf.ny, sw.cos_wd)) > 0.0);
                    wf.forced = cf.forced_c * (leeward ? 0.5 : 1.0) * sw.sqrt_ws;
                    wf.air = is_out ? sw.t_out : (fl & FC_SPACE) ? l.zface : cf.amb;
                    l.trad_[f] = is_out ? wf.trad_out : wf.air;
                    const double hn = tarp_lean(wf.air - wf.ts, cf.kpos, cf.kneg);
                    l.h_[f] = is_out ? hn + wf.forced : hn;
                    if (l.h_[f] != l.h_[f]) wf.status |= ST_NAN_H;
                    const double xr = 273.15 + (l.trad_[f] + wf.ts) / 2.;
                    l.rad_[f] = cf.eps4s * ((xr * xr) * xr);
                    l.asol_[f] = cf.alpha_end * wf.solar;
                }
            if (PANE) {   // both faces of a pane solve the wall's 2x2 system with identical inputs
                std::vector<Lane> Q = L;   // values before the exchange (shuffles read the partners' registers)
                for (int lane = 0; lane < 32; ++lane) {
                    Lane& l = L[lane];
                    if (l.win4_lane) {   // the front lane solves the window's four no-mass nodes
                        const Lane& me = Q[lane];
                        const int fb = G == 1 ? 1 : 0;                     // index of the back face's values …
                        const Lane& p = G == 1 ? me : Q[lane ^ (G - 1)];   // … on this lane (G = 1) or on the partner
                        const double sf = me.w[0].solar, sb = p.w[fb].solar;
                        const double sol[4] = {me.c[0].alpha_end * sf + me.c[0].alpha_x * sb, l.w_a1f * sf + l.w_a1b * sb,
                                               l.w_a2f * sf + l.w_a2b * sb, p.c[fb].alpha_end * sb + p.c[fb].alpha_x * sf};
                        double t[4] = {T[lane * NL + 0], T[lane * NL + 1], T[lane * NL + 2], T[lane * NL + 3]};
                        unsigned long long it = 0;
                        solve_window4(me.w[0].air, me.h_[0], me.rad_[0], me.trad_[0], p.w[fb].air, p.h_[fb], p.rad_[fb],
                                      p.trad_[fb], sol, me.c[0].u_nb, l.w_u23, l.w_cav, t, it, l.w[0].status);
                        l.w[0].nm_iters += it;
                        for (int q = 0; q < 4; ++q) T[lane * NL + q] = t[q];
                        l.w[0].ts = t[0];
                        if (G == 1) l.w[1].ts = t[3];
                        continue;
                    }
                    if (G == 1) {
                        if (!(l.c[0].fl & FC_PANE)) continue;
                        const double sol0 = l.c[0].alpha_end * l.w[0].solar + l.c[0].alpha_x * l.w[1].solar;
                        const double sol1 = l.c[1].alpha_end * l.w[1].solar + l.c[1].alpha_x * l.w[0].solar;
                        double t0 = l.w[0].ts, t1 = l.w[1].ts;
                        unsigned long long it = 0;
                        solve_pane<true>(l.w[0].air, l.h_[0], l.rad_[0], l.trad_[0], sol0, l.w[1].air, l.h_[1], l.rad_[1],
                                         l.trad_[1], sol1, l.c[0].u_nb, t0, t1, it, l.w[0].status);
                        l.w[0].nm_iters += it;
                        l.w[0].ts = t0;
                        l.w[1].ts = t1;
                    } else {
                        if (!(l.c[0].fl & FC_PANE)) continue;
                        const Lane& p = Q[lane ^ (G - 1)];
                        const Lane& me = Q[lane];
                        const double sol_own = me.c[0].alpha_end * me.w[0].solar + me.c[0].alpha_x * p.w[0].solar;
                        const double p_sol = p.c[0].alpha_end * p.w[0].solar + p.c[0].alpha_x * me.w[0].solar;
                        const bool bk = l.bk;
                        double t0 = bk ? p.w[0].ts : me.w[0].ts, t1 = bk ? me.w[0].ts : p.w[0].ts;
                        unsigned long long it = 0;
                        solve_pane<true>(bk ? p.w[0].air : me.w[0].air, bk ? p.h_[0] : me.h_[0], bk ? p.rad_[0] : me.rad_[0],
                                         bk ? p.trad_[0] : me.trad_[0], bk ? p_sol : sol_own, bk ? me.w[0].air : p.w[0].air,
                                         bk ? me.h_[0] : p.h_[0], bk ? me.rad_[0] : p.rad_[0], bk ? me.trad_[0] : p.trad_[0],
                                         bk ? sol_own : p_sol, me.c[0].u_nb, t0, t1, it, l.w[0].status);
                        if (!bk) l.w[0].nm_iters += it;
                        l.w[0].ts = bk ? t1 : t0;
                    }
                }
            }
            for (Lane& l : L)
                for (int f = 0; f < NF; ++f) {
                    const bool back = G == 1 ? f == 1 : l.bk;
                    const FaceC& cf = l.c[f];
                    FaceS& wf = l.w[f];
                    const int fl = cf.fl;
                    const double h = l.h_[f], rad = l.rad_[f], trad = l.trad_[f], asol = l.asol_[f];
                    l.lk[f].t_in = 0.;
                    if (END) {
                        if (fl & FC_NM2) {
                            double t_in = back ? l.t1_b : l.t1_f;
                            solve_end_pair<true>(back, wf.air, h, rad, trad, asol, cf.u_nb, cf.u_nb2, back ? l.t2_b : l.t2_f,
                                                 wf.ts, t_in, wf.nm_iters, wf.status);
                            l.lk[f].t_in = t_in;
                        } else if (fl & FC_NM) {
                            wf.ts = solve_end_node<true>(wf.air, h, rad, trad, cf.u_nb, back ? l.t1_b : l.t1_f, asol, wf.ts,
                                                         wf.nm_iters, wf.status);
                        }
                    }
                    const double hs = h > 1e-12 ? h : 1e-12;
                    l.lk[f].ghost = wf.air + (rad * (trad - wf.ts) + asol) * rcp_lean(hs);
                    l.lk[f].link = hs * cf.v_end;
                }
            // face lanes -> node lanes, placement
            std::vector<double> gfv(32), gbv(32);
            for (int lane = 0; lane < 32; ++lane) {
                Lane& l = L[lane];
                const Lane& lb = L[l.base + G - 1];
                const double gf = l.lk[0].ghost, lf = l.lk[0].link;
                const double gb = lb.lk[NF - 1].ghost, lbk = lb.lk[NF - 1].link;
                const double tf0 = l.w[0].ts, tf1 = l.lk[0].t_in, tb0 = lb.w[NF - 1].ts, tb1 = lb.lk[NF - 1].t_in;
                double* t = &T[lane * NL];
                double* a = &A[lane * NL];
                double* b = &B[lane * NL];
                b[0] = l.first ? lf : b[0];
                if (END) {
                    t[0] = (l.first && l.nmf) ? tf0 : t[0];
                    t[1] = (l.first && l.nmf2) ? tf1 : t[1];
                }
                for (int i = 0; i < NL; ++i) {
                    a[i] = i == l.last ? lbk : a[i];
                    if (END) {
                        t[i] = (i == l.last && l.nmb) ? tb0 : t[i];
                        t[i] = (i == l.last - 1 && l.nmb2) ? tb1 : t[i];
                    }
                    t[i] = i == l.last + 1 ? gb : t[i];
                }
                gfv[lane] = gf;
                gbv[lane] = gb;
            }
            // node role: wt_rk4
            auto halos = [&](const std::vector<double>& x, bool first_sweep) {
                for (int lane = 0; lane < 32; ++lane) {
                    const Lane& l = L[lane];
                    double lv = first_sweep ? gfv[lane] : 0., rv = first_sweep ? gbv[lane] : 0.;
                    if (G > 1) {
                        const double sl = l.part > 0 ? x[(lane - 1) * NL + NL - 1] : x[lane * NL + NL - 1];
                        const double sr = l.part < G - 1 ? x[(lane + 1) * NL] : x[lane * NL];
                        lv = l.first ? lv : sl;
                        rv = l.edge ? rv : sr;
                    }
                    left[lane] = lv;
                    right[lane] = rv;
                }
            };
            halos(T, true);
            for (int lane = 0; lane < 32; ++lane) {
                const double *t = &T[lane * NL], *a = &A[lane * NL], *b = &B[lane * NL];
                double* x = &X[lane * NL];
                for (int i = 0; i < NL; ++i) {
                    const double dn = (i + 1 < NL ? t[i + 1] : right[lane]) - t[i];
                    const double dp = t[i] - (i > 0 ? t[i - 1] : left[lane]);
                    x[i] = a[i] * dn - b[i] * dp;
                }
            }
            for (int q = 0; q < 32 * NL; ++q) T[q] += X[q];
            for (int k = 0; k < 3; ++k) {
                const double cc = k == 0 ? 1. / 2. : k == 1 ? 1. / 6. : 1. / 24.;
                halos(X, false);
                std::vector<double> Y(32 * NL);
                for (int lane = 0; lane < 32; ++lane) {
                    const double *x = &X[lane * NL], *a = &A[lane * NL], *b = &B[lane * NL];
                    for (int i = 0; i < NL; ++i) {
                        const double dn = (i + 1 < NL ? x[i + 1] : right[lane]) - x[i];
                        const double dp = x[i] - (i > 0 ? x[i - 1] : left[lane]);
                        Y[lane * NL + i] = a[i] * dn - b[i] * dp;
                    }
                }
                X = Y;
                for (int q = 0; q < 32 * NL; ++q) T[q] = fma(cc, X[q], T[q]);
            }
            gather_back();
            // face role: new film coefficients, flows, zone
            std::vector<double> va(32, 0.), vb(32, 0.);
            for (int lane = 0; lane < 32; ++lane) {
                Lane& l = L[lane];
                for (int f = 0; f < NF; ++f) {
                    const bool back = G == 1 ? f == 1 : l.bk;
                    const FaceC& cf = l.c[f];
                    FaceS& wf = l.w[f];
                    const double ts = back ? l.ts_b : l.ts_f;
                    const double hn = tarp_lean(wf.air - ts, cf.kpos, cf.kneg);
                    const double h = (cf.fl & FC_OUT) ? hn + wf.forced : hn;
                    if (h != h) wf.status |= ST_NAN_H;
                    wf.ts = ts;
                    wf.h = h;
                    wf.flow = (ts - wf.air) * h;
                    wf.e += wf.flow * cf.area * dt;
                    if (back && (SIMPLE || (cf.fl & FC_SPACE))) {
                        vb[lane] = h * cf.area;
                        va[lane] = vb[lane] * ts;
                    }
                }
            }
            if (!SIMPLE)   // segmented shuffle reduction over the back lanes of each model (stride G)
                for (int o = 1; o < W; o <<= 1) {
                    const std::vector<double> pa = va, pb = vb;
                    for (int lane = 0; lane < 32; ++lane) {
                        const Lane& l = L[lane];
                        if (l.wl + o < l.seg_end) {
                            const int src = lane + o * G < 32 ? lane + o * G : lane;
                            va[lane] = pa[lane] + pa[src];
                            vb[lane] = pb[lane] + pb[src];
                        }
                    }
                }
            for (int lane = 0; lane < 32; ++lane) {
                Lane& l = L[lane];
                if (l.szone) {
                    const double inv_cc = zone_inv_c_t<true>(l.zvol, l.zT);
                    const double tn = zone_advance<true>(l.zT, dt, inv_cc, l.za0 + va[lane], l.zb0 + vb[lane]);
                    if (tn != tn) V.status[l.sg] |= ST_NAN_ZONE;
                    l.zT = tn;
                }
            }
            for (Lane& l : L) l.zface = SIMPLE ? l.zT : L[l.zsrc].zT;
        }
        if (P.record_every > 0 && (step + 1) % P.record_every == 0) {
            for (const Lane& l : L)
                if (zone_rec_base && l.szone) zone_rec_base[(int64_t)rec * V.n_dev_zones + l.sgz] = l.zT;
            if (P.node_rec) {
                for (int lane = 0; lane < 32; ++lane)
                    for (int i = 0; i < NL; ++i)
                        P.node_rec[(int64_t)rec * P.slab_len + tile.node_offset + (int64_t)i * 32 + lane] = T[lane * NL + i];
                for (const Lane& l : L) {
                    if (!l.fwall) continue;
                    double* rr = P.sc_rec + (int64_t)rec * 8 * NS;
                    for (int f = 0; f < NF; ++f) {
                        const int kf = (G == 1 ? f == 1 : l.bk) ? 1 : 0;
                        rr[(0 + kf) * NS + l.sg] = l.w[f].h; rr[(2 + kf) * NS + l.sg] = l.w[f].flow;
                        rr[(4 + kf) * NS + l.sg] = l.w[f].sol; rr[(6 + kf) * NS + l.sg] = l.w[f].ir;
                    }
                }
            }
            ++rec;
        }
    }
    for (int lane = 0; lane < 32; ++lane)
        for (int i = 0; i < NL; ++i) V.nodeT[tile.node_offset + (int64_t)i * 32 + lane] = T[lane * NL + i];
    for (const Lane& l : L) {
        if (l.fwall)
            for (int f = 0; f < NF; ++f) {
                const bool back = G == 1 ? f == 1 : l.bk;
                const int kf = back ? 1 : 0;
                V.sc[(0 + kf) * NS + l.sg] = l.w[f].h; V.sc[(2 + kf) * NS + l.sg] = l.w[f].flow;
                V.sc[(4 + kf) * NS + l.sg] = l.w[f].sol; V.sc[(6 + kf) * NS + l.sg] = l.w[f].ir;
                if (back) V.energy_b[l.sg] += l.w[f].e; else V.energy_f[l.sg] += l.w[f].e;
                V.nm_iters[l.sg] += l.w[f].nm_iters;
                V.status[l.sg] |= l.w[f].status;
            }
        if (l.szone) V.zoneT[l.sgz] = l.zT;
    }
}

// hb_face_kernel's data flow (all-massive tiles and tiles whose no-mass chunks are single end nodes): per-face lanes.
void march_tile_face(const DevView& V, const SeriesDev& P, const Tile& tile, double* zone_rec_base) {
    const int G = tile.G, W = tile.W, NT = G * W, NL = tile.NL;
    const bool END = tile.flags & TILE_END;
    const int S = tile.S;
    const double dt = tile.dt;
    const int64_t NS = V.n_dev_surfaces;
    std::vector<double> T(NT * NL), A(NT * NL), B(NT * NL), Vv(NT * NL);
    std::vector<FaceC> fc(2 * W);
    std::vector<FaceS> fs(2 * W);
    std::vector<char> have(W, 0);
    std::vector<ZoneC> zc(W);
    std::vector<double> zT(W, 0.), pub0(W, 0.), pub1(W, 0.), pub2(W, 0.), pub3(W, 0.);
    std::vector<double> e0(W, 0.), e1(W, 0.), e2(W, 0.), e3(W, 0.), e4(W, 0.), e5(W, 0.), link(8 * W, 0.), ir_state(W, 0.);
    std::vector<int32_t> zf(W, -1), zb(W, -1);
    std::vector<SubWeather> swt(kMaxSubsteps);
    for (int tid = 0; tid < NT; ++tid)
        for (int i = 0; i < NL; ++i) {
            const int64_t o = tile.node_offset + (int64_t)i * NT + tid;
            T[tid * NL + i] = V.nodeT[o];
            A[tid * NL + i] = V.coefA[o];
            B[tid * NL + i] = V.coefB[o];
        }
    auto publish_ends = [&]() {
        for (int tid = 0; tid < NT; ++tid) {
            const int wl = tid / G, part = tid % G;
            const int n = V.n_nodes[(int64_t)tile.surface_begin + wl];
            const int iN = n - part * NL;
            if (part == 0) { e0[wl] = T[tid * NL + 1]; e1[wl] = T[tid * NL + 2]; e4[wl] = T[tid * NL + 3]; }
            for (int i = 0; i < NL; ++i) {
                if (i == iN) e3[wl] = T[tid * NL + i];
                if (i == iN - 1) e2[wl] = T[tid * NL + i];
                if (i == iN - 2) e5[wl] = T[tid * NL + i];
            }
        }
    };
    publish_ends();
    for (int t = 0; t < W; ++t) {
        const int64_t sg = (int64_t)tile.surface_begin + t;
        if (V.n_nodes[sg] > 0) {
            have[t] = 1;
            WallC c;
            load_wallc(V, sg, P.single_site != 0, c);
            const int flg = V.flags[sg];
            for (int k = 0; k < 2; ++k) {
                const bool pane = (tile.flags & TILE_PANE) && ((flg >> 12) & 1);
                const bool nm = END && !pane && ((flg >> (k ? 11 : 10)) & 1);
                const bool nm2 = nm && ((flg >> (k ? 14 : 13)) & 1);
                load_facec(c, k == 1, k ? V.u_b[sg] : V.u_f[sg], nm, fc[2 * t + k], pane, k ? V.abL[sg] : V.af0[sg],
                           k ? V.ax_b[sg] : V.ax_f[sg], nm2, k ? V.u_b2[sg] : V.u_f2[sg]);
                FaceS w{};
                w.h = V.sc[(0 + k) * NS + sg]; w.flow = V.sc[(2 + k) * NS + sg];
                w.sol = V.sc[(4 + k) * NS + sg]; w.ir = V.sc[(6 + k) * NS + sg];
                w.ts = k ? e3[t] : e0[t];
                fs[2 * t + k] = w;
            }
            zf[t] = c.zf; zb[t] = c.zb;
            ir_state[t] = fs[2 * t].ir;
        }
        const int64_t gz = (int64_t)tile.zone_begin + t;
        ZoneC z{};
        if (t < tile.n_zones) {
            z.vol = V.zvol[gz]; z.q = V.zq[gz]; z.inf_v = V.zinf_v[gz]; z.inf_t = V.zinf_t[gz];
            z.vent_v = V.zvent_v[gz]; z.vent_t = V.zvent_t[gz]; z.sb = V.z_surf_begin[gz]; z.se = V.z_surf_end[gz];
            zT[t] = V.zoneT[gz];
        }
        zc[t] = z;
    }
    auto weather_at = [&](int site, int step, int s, SubWeather& out) {
        const double* wtab = P.weather + (int64_t)site * P.weather_rows * 3;
        const double* w0 = wtab + (int64_t)step * 3;
        double t_out, wdir, wspeed;
        if (P.weather_mode == HB_WEATHER_LINEAR) {
            const double f = (double)(s + 1) / (double)S;
            t_out = lerp_rn(w0[0], w0[3], f); wdir = lerp_rn(w0[1], w0[4], f); wspeed = lerp_rn(w0[2], w0[5], f);
        } else if (P.weather_mode == HB_WEATHER_PER_SUBSTEP) {
            w0 = wtab + ((int64_t)step * S + s) * 3;
            t_out = w0[0]; wdir = w0[1]; wspeed = w0[2];
        } else {
            t_out = w0[0]; wdir = w0[1]; wspeed = w0[2];
        }
        sub_weather(t_out, wdir, wspeed, out);
    };
    int rec = 0;
    std::vector<double> left(NT), right(NT);
    for (int step = 0; step < P.n_steps; ++step) {
        for (int t = 0; t < tile.n_zones; ++t)   // scheduled zone inputs of this main step
            load_zone_schedule(P, step, (int64_t)tile.zone_begin + t, zc[t]);
        if (P.single_site)
            for (int s = 0; s < S; ++s) weather_at(0, step, s, swt[s]);
        for (int t = 0; t < W; ++t) {
            if (!have[t]) continue;
            for (int k = 0; k < 2; ++k) {
                const FaceC& c = fc[2 * t + k];
                FaceS& w = fs[2 * t + k];
                if (P.input_mode != HB_INPUT_STATE) {
                    const double* in = P.input_mode == HB_INPUT_SITE
                                           ? P.inputs + ((int64_t)c.site * P.n_steps + step) * 4
                                           : P.inputs + ((int64_t)step * P.n_desc_surfaces + c.desc_index) * 4;
                    w.sol = in[k]; w.ir = in[2 + k];
                    if (k == 0 && P.ir_mode == HB_IR_NET_GAIN) {
                        if (c.eps > 1e-3) {
                            const double tk = w.ts + 273.15;
                            const double tk2 = tk * tk;
                            w.ir = w.ir / c.area / c.eps + kSigma * (tk2 * tk2);
                        } else {
                            w.ir = ir_state[t];
                        }
                    }
                }
                face_begin_main_step(c, w);
            }
        }
        for (int s = 0; s < S; ++s) {
            for (int t = 0; t < W; ++t) {
                if (!have[t]) continue;
                FaceEnv env[2];
                double t_nb[2] = {0., 0.};
                for (int k = 0; k < 2; ++k) {
                    const FaceC& c = fc[2 * t + k];
                    FaceS& w = fs[2 * t + k];
                    SubWeather sw;
                    if (P.single_site) sw = swt[s];
                    else weather_at(c.site, step, s, sw);
                    const double zone_t = c.zone >= 0 ? zT[c.zone] : 0.;
                    double t_front = 0., ts_front = 0.;
                    if (c.fl & FC_BACK_AMB) {
                        t_front = (c.fl & FC_FRONT_OUT) ? sw.t_out : (c.fl & FC_FRONT_SPACE) ? zT[c.zone_front] : c.amb_front;
                        ts_front = e0[t];
                    }
                    if (c.fl & FC_NM) t_nb[k] = k ? e2[t] : e1[t];
                    face_env(c, w, sw, zone_t, t_front, ts_front, env[k]);
                }
                if (fc[2 * t].fl & FC_PANE) {   // the two faces of a pane solve one 2x2 system (hb_face_kernel: pane_solve)
                    FaceS &wf = fs[2 * t], &wb = fs[2 * t + 1];
                    const FaceC &cf = fc[2 * t], &cb = fc[2 * t + 1];
                    const double sol0 = cf.alpha_end * wf.solar + cf.alpha_x * wb.solar;
                    const double sol1 = cb.alpha_end * wb.solar + cb.alpha_x * wf.solar;
                    double t0 = wf.ts, t1 = wb.ts;
                    solve_pane(wf.air, env[0].h, env[0].rad, env[0].trad, sol0, wb.air, env[1].h, env[1].rad, env[1].trad,
                               sol1, cf.u_nb, t0, t1, wf.nm_iters, wf.status);
                    wf.ts = t0;
                    wb.ts = t1;
                }
                for (int k = 0; k < 2; ++k) {
                    const FaceC& c = fc[2 * t + k];
                    FaceS& w = fs[2 * t + k];
                    double ghost, lk;
                    face_links(c, w, env[k], t_nb[k], k ? e5[t] : e4[t], ghost, lk);
                    if (c.fl & FC_NM2) link[(k ? 7 : 6) * W + t] = t_nb[k];
                    link[(k ? 2 : 0) * W + t] = ghost;
                    link[(k ? 3 : 1) * W + t] = lk;
                    link[(k ? 5 : 4) * W + t] = w.ts;
                }
            }
            for (int tid = 0; tid < NT; ++tid) {
                const int wl = tid / G, part = tid % G;
                const int64_t g = (int64_t)tile.surface_begin + wl;
                const int n = V.n_nodes[g];
                const int nmf = END ? (V.flags[g] >> 10) & 1 : 0, nmb = END ? (V.flags[g] >> 11) & 1 : 0;
                const int nmf2 = END ? (V.flags[g] >> 13) & 1 : 0, nmb2 = END ? (V.flags[g] >> 14) & 1 : 0;
                double* t = &T[tid * NL];
                if (part == 0) {
                    t[0] = link[0 * W + wl];
                    B[tid * NL + 1] = link[1 * W + wl];
                    if (nmf) t[1] = link[4 * W + wl];
                    if (nmf2) t[2] = link[6 * W + wl];
                }
                for (int i = 0; i < NL; ++i) {
                    if (i == n - part * NL) {
                        A[tid * NL + i] = link[3 * W + wl];
                        if (nmb) t[i] = link[5 * W + wl];
                    }
                    if (i == n + 1 - part * NL) t[i] = link[2 * W + wl];
                    if (i == n - 1 - part * NL && nmb2) t[i] = link[7 * W + wl];
                }
            }
            auto halos = [&](const std::vector<double>& x) {
                for (int tid = 0; tid < NT; ++tid) {
                    const int part = tid % G;
                    left[tid] = part > 0 ? x[(tid - 1) * NL + NL - 1] : x[tid * NL + NL - 1];
                    right[tid] = part < G - 1 ? x[(tid + 1) * NL] : x[tid * NL];
                }
            };
            auto sweeps = [&](auto nl) {   // the four RK4 sweeps for lanes of nl slots
                constexpr int N = decltype(nl)::value;
                using Arr = double(&)[N];
                halos(T);
                for (int tid = 0; tid < NT; ++tid) {
                    sweep_first<N>((Arr)T[tid * N], (Arr)Vv[tid * N], (Arr)A[tid * N], (Arr)B[tid * N], left[tid], right[tid]);
                    for (int i = 0; i < N; ++i) T[tid * N + i] += Vv[tid * N + i];
                }
                const double cs[3] = {1. / 2., 1. / 6., 1. / 24.};
                for (int k = 0; k < 3; ++k) {
                    halos(Vv);
                    for (int tid = 0; tid < NT; ++tid)
                        sweep_next<N>((Arr)T[tid * N], (Arr)Vv[tid * N], (Arr)A[tid * N], (Arr)B[tid * N], left[tid],
                                      right[tid], cs[k]);
                }
            };
            if (NL == kNL) sweeps(std::integral_constant<int, kNL>{});
            else if (NL == kNLBig) sweeps(std::integral_constant<int, kNLBig>{});
            else sweeps(std::integral_constant<int, kNLHuge>{});
            publish_ends();
            for (int t = 0; t < W; ++t) {
                if (!have[t]) continue;
                for (int k = 0; k < 2; ++k) {
                    double hA, hAT;
                    face_post(fc[2 * t + k], fs[2 * t + k], k ? e3[t] : e0[t], e0[t], dt, hA, hAT);
                    (k ? pub2 : pub0)[t] = hA;
                    (k ? pub3 : pub1)[t] = hAT;
                }
            }
            std::vector<double> znew(zT);
            for (int t = 0; t < tile.n_zones; ++t) {
                const double tn = zone_update(zc[t], t, zT[t], dt, pub0.data(), pub1.data(), pub2.data(), pub3.data(),
                                              zf.data(), zb.data());
                if (tn != tn) V.status[(int64_t)tile.surface_begin + zc[t].sb] |= ST_NAN_ZONE;
                znew[t] = tn;
            }
            zT = znew;
        }
        if (P.record_every > 0 && (step + 1) % P.record_every == 0) {
            if (zone_rec_base)
                for (int t = 0; t < tile.n_zones; ++t)
                    zone_rec_base[(int64_t)rec * V.n_dev_zones + tile.zone_begin + t] = zT[t];
            if (P.node_rec) {
                for (int tid = 0; tid < NT; ++tid)
                    for (int i = 0; i < NL; ++i)
                        P.node_rec[(int64_t)rec * P.slab_len + tile.node_offset + (int64_t)i * NT + tid] = T[tid * NL + i];
                for (int t = 0; t < W; ++t) {
                    if (!have[t]) continue;
                    const int64_t sg = (int64_t)tile.surface_begin + t;
                    double* r = P.sc_rec + (int64_t)rec * 8 * NS;
                    for (int k = 0; k < 2; ++k) {
                        const FaceS& w = fs[2 * t + k];
                        r[(0 + k) * NS + sg] = w.h; r[(2 + k) * NS + sg] = w.flow;
                        r[(4 + k) * NS + sg] = w.sol; r[(6 + k) * NS + sg] = w.ir;
                    }
                }
            }
            ++rec;
        }
    }
    for (int tid = 0; tid < NT; ++tid)
        for (int i = 0; i < NL; ++i) V.nodeT[tile.node_offset + (int64_t)i * NT + tid] = T[tid * NL + i];
    for (int t = 0; t < W; ++t) {
        if (!have[t]) continue;
        const int64_t sg = (int64_t)tile.surface_begin + t;
        for (int k = 0; k < 2; ++k) {
            const FaceS& w = fs[2 * t + k];
            V.sc[(0 + k) * NS + sg] = w.h; V.sc[(2 + k) * NS + sg] = w.flow;
            V.sc[(4 + k) * NS + sg] = w.sol; V.sc[(6 + k) * NS + sg] = w.ir;
            (k ? V.energy_b : V.energy_f)[sg] += w.e;
            V.nm_iters[sg] += w.nm_iters;
            V.status[sg] |= w.status;
        }
    }
    for (int t = 0; t < tile.n_zones; ++t) V.zoneT[(int64_t)tile.zone_begin + t] = zT[t];
}
}  // namespace

extern "C" {

const char* emu_last_error() { return g_err.c_str(); }

// state: in/out, reference layout (NULL-initialised state = the initial values).  zone inputs may be NULL.
// zone_record [n_rec][zones], state_record [n_rec][state_len] may be NULL.  Returns 0 on success.
int emu_run(const hb_batch_desc* desc, int use_initial_state, double* state, const double* q_internal,
            const double* inf_v, const double* inf_t, const double* vent_v, const double* vent_t,
            const hb_series* s, long long* nomass_iterations, long long* status_or) {
    try {
        BatchPlan B;
        plan_batch(desc, B);
        HostArrays H;
        pack_arrays(B, H);
        if (!use_initial_state) scatter_state(B, state, H.nodeT.data(), H.sc.data(), H.zoneT.data());
        auto scatter = [&](const double* src, std::vector<double>& dst) {
            if (src)
                for (size_t z = 0; z < B.dev_zone_of.size(); ++z) dst[B.dev_zone_of[z]] = src[z];
        };
        scatter(q_internal, H.zq); scatter(inf_v, H.zinf_v); scatter(inf_t, H.zinf_t);
        scatter(vent_v, H.zvent_v); scatter(vent_t, H.zvent_t);
        bool single = true;
        if (s->model_site)
            for (size_t gd = 0; gd < B.dev_surface_of.size(); ++gd) {
                H.site[B.dev_surface_of[gd]] = s->model_site[B.surface_model[gd]];
                if (s->model_site[B.surface_model[gd]] != 0) single = false;
            }
        DevView V = H.view(B);
        SeriesDev P{};
        P.n_steps = s->n_steps; P.weather_mode = s->weather_mode; P.input_mode = s->input_mode; P.ir_mode = s->ir_mode;
        P.weather = s->weather; P.weather_rows = s->weather_rows; P.inputs = s->inputs;
        P.n_desc_surfaces = (int64_t)B.surfaces.size();
        P.single_site = single ? 1 : 0;
        P.slab_len = B.node_slab_len;
        P.n_dev_zones = B.n_dev_zones;
        std::vector<double> sched_dev[5];
        {
            const double* src[5] = {s->zone_gain, s->zone_infiltration_volume, s->zone_infiltration_temperature,
                                    s->zone_ventilation_volume, s->zone_ventilation_temperature};
            const double** dst[5] = {&P.zone_gain, &P.zone_inf_v, &P.zone_inf_t, &P.zone_vent_v, &P.zone_vent_t};
            const int64_t nz = (int64_t)B.dev_zone_of.size();
            for (int k = 0; k < 5; ++k) {
                if (!src[k] || s->n_steps <= 0) continue;
                sched_dev[k].assign((size_t)s->n_steps * B.n_dev_zones, k == 2 || k == 4 ? NAN : 0.);
                for (int64_t t = 0; t < s->n_steps; ++t)
                    for (int64_t z = 0; z < nz; ++z) sched_dev[k][(size_t)t * B.n_dev_zones + B.dev_zone_of[z]] = src[k][t * nz + z];
                *dst[k] = sched_dev[k].data();
            }
        }
        const int n_rec = s->record_every > 0 ? s->n_steps / s->record_every : 0;
        std::vector<double> zr, nr, sr;
        P.record_every = n_rec > 0 && (s->zone_t_record || s->state_record) ? s->record_every : 0;
        if (P.record_every) {
            zr.resize((size_t)n_rec * B.n_dev_zones);
            if (s->state_record) {
                nr.resize((size_t)n_rec * B.node_slab_len);
                sr.resize((size_t)n_rec * 8 * B.n_dev_surfaces);
                P.node_rec = nr.data();
                P.sc_rec = sr.data();
            }
        }
        for (const Tile& t : B.tiles) {
            if (t.flags & TILE_WARP) march_tile_warp(V, P, t, zr.empty() ? nullptr : zr.data());
            else if (t.flags & (TILE_EXT | TILE_DIST)) march_tile(V, P, t, zr.empty() ? nullptr : zr.data());
            else march_tile_face(V, P, t, zr.empty() ? nullptr : zr.data());
        }
        if (P.record_every) {
            const int64_t nz_total = (int64_t)B.zone_volume.size();
            for (int r = 0; r < n_rec; ++r) {
                const double* z = zr.data() + (size_t)r * B.n_dev_zones;
                if (s->zone_t_record)
                    for (int64_t k = 0; k < nz_total; ++k) s->zone_t_record[(int64_t)r * nz_total + k] = z[B.dev_zone_of[k]];
                if (s->state_record)
                    gather_state(B, nr.data() + (size_t)r * B.node_slab_len, sr.data() + (size_t)r * 8 * B.n_dev_surfaces, z,
                                 s->state_record + (int64_t)r * B.state_len_total);
            }
        }
        gather_state(B, H.nodeT.data(), H.sc.data(), H.zoneT.data(), state);
        long long it = 0, st = 0;
        for (auto v : H.nm_iters) it += (long long)v;
        for (auto v : H.status) st |= v;
        if (nomass_iterations) *nomass_iterations = it;
        if (status_or) *status_or = st;
        return 0;
    } catch (const std::exception& e) {
        g_err = e.what();
        return 1;
    }
}

// tiles of the plan the library builds for `desc` (what hb_reduce_noalloc's HB_REDUCE_NOALLOC_MAX_TILES counts)
long long emu_n_tiles(const hb_batch_desc* desc) {
    try {
        BatchPlan B;
        plan_batch(desc, B);
        return (long long)B.tiles.size();
    } catch (const std::exception& e) {
        g_err = e.what();
        return -1;
    }
}

// (G, NL, W, flags, S) of up to `cap` tiles of that plan; returns the number of tiles
long long emu_tile_table(const hb_batch_desc* desc, int* out, long long cap) {
    try {
        BatchPlan B;
        plan_batch(desc, B);
        for (size_t i = 0; i < B.tiles.size() && (long long)i < cap; ++i) {
            const Tile& t = B.tiles[i];
            out[5 * i + 0] = t.G; out[5 * i + 1] = t.NL; out[5 * i + 2] = t.W; out[5 * i + 3] = t.flags; out[5 * i + 4] = t.S;
        }
        return (long long)B.tiles.size();
    } catch (const std::exception& e) {
        g_err = e.what();
        return -1;
    }
}

long long emu_state_len(const hb_batch_desc* desc) {
    try {
        BatchPlan B;
        plan_batch(desc, B);
        return B.state_len_total;
    } catch (const std::exception& e) {
        g_err = e.what();
        return -1;
    }
}
}

// hb_dev.hpp — views of the packed SoA buffers (shared by the host packer, the launchers and the kernels).
#pragma once
#include <cstdint>
#include <cstdlib>
#include <vector>

#include "hb_internal.hpp"

namespace hb {

struct CavityDev {
    int32_t node, gas;
    double thickness, height, angle, eout, ein;
};

// Pointers into the packed buffers.  On the GPU they are device pointers (hb_capi.cu uploads HostArrays verbatim).
struct DevView {
    // ---- per device surface slot (tile.surface_begin + wall) ----
    const int32_t* n_nodes;   // 0 on padding slots
    const int32_t* flags;     // bits 0-3 front kind, 4-7 back kind, 8 distributed solar, 9 has cavity,
                              // 15: double-glazed window solved by the front face lane (warp tiles; 10 / 11 clear)
                              // 10 / 11: node 0 / node n-1 is a single-node no-mass chunk (TILE_END tiles)
                              // 13 / 14: the front / back end is a TWO-node no-mass chunk (nodes 0,1 / n-2,n-1; 10 / 11 set too)
                              // 12: the wall is a pane (two no-mass nodes, chunk (0,2); TILE_PANE tiles; 10 and 11 set too)
    const int32_t* zone_f;    // tile-local zone index or -1
    const int32_t* zone_b;
    const int32_t* site;
    const int32_t* desc_index;  // descriptor-order surface index (for per-surface inputs)
    const double* area;
    const double* nx;
    const double* ny;
    const double* cos_tilt;
    const double* forced_c;   // 2.537·1.67·sqrt(perimeter·wind_modifier/area)  (convection.rs:157-163)
    const double* eps_f;
    const double* eps_b;
    const double* amb_f;
    const double* amb_b;
    const double* hs_ov_f;
    const double* hs_ov_b;
    const double* af0;        // front solar absorptance of node 0
    const double* abL;        // back solar absorptance of node n-1
    const double* v0;         // dt/C of node 0 (0 when it has no mass)
    const double* vL;         // dt/C of node n-1
    const double* u_f;        // conductance of the link node 0 → 1 (TILE_END)
    const double* u_b;        // conductance of the link node n-2 → n-1
    const double* u_f2;       // conductance of the link node 1 → 2 (two-node no-mass chunk at the front end)
    const double* u_b2;       // conductance of the link node n-3 → n-2 (… at the back end)
    const double* ax_f;       // panes (flags bit 12): absorptance of node 0 for the BACK irradiance
    const double* ax_b;       // panes: absorptance of node n-1 for the FRONT irradiance
    // double-glazed windows on warp tiles (flags bit 15: pane | gas | pane, four no-mass nodes): absorptances of the two
    // inner nodes for the front / back irradiance
    const double* wa1f;
    const double* wa1b;
    const double* wa2f;
    const double* wa2b;
    const int32_t* nm_off;    // CSR over no-mass chunks: pairs (ini, fin)
    const int32_t* nm_pairs;
    const int32_t* cav_off;   // CSR over cavity links
    const CavityDev* cav;
    // ---- [kNL][G*W] slabs per tile (tile.node_offset) ----
    double* nodeT;            // slot temperatures (ghost/padding slots hold scratch values)
    const double* coefA;      // u_i·dt/C_i   towards slot+1 (0 on ghosts, padding, no-mass nodes and the last node)
    const double* coefB;      // u_{i-1}·dt/C_i towards slot-1 (0 on …, and the first node)
    // ---- [node][W] wall-major arrays of TILE_EXT tiles (tile.ext_offset) ----
    const double* ext_u;
    const double* ext_mass;
    const double* ext_af;
    const double* ext_ab;
    // ---- state scalars [8][n_dev_surfaces]: h_f, h_b, flow_f, flow_b, solar_f, solar_b, ir_f, ir_b ----
    double* sc;
    int64_t n_dev_surfaces;
    double* energy_f;
    double* energy_b;
    unsigned long long* nm_iters;
    int32_t* status;
    // ---- zones (tile.zone_begin + local zone) ----
    double* zoneT;
    const double* zvol;
    const double* zq;
    const double* zinf_v;
    const double* zinf_t;
    const double* zvent_v;
    const double* zvent_t;
    const int32_t* z_surf_begin;  // tile-local wall range of the zone's model
    const int32_t* z_surf_end;
    int64_t n_dev_zones;
    const Tile* tiles;
};

// Host image of every packed buffer (what hb_create uploads).
struct HostArrays {
    std::vector<int32_t> n_nodes, flags, zone_f, zone_b, site, desc_index, nm_off, nm_pairs, cav_off;
    std::vector<CavityDev> cav;
    std::vector<double> area, nx, ny, cos_tilt, forced_c, eps_f, eps_b, amb_f, amb_b, hs_ov_f, hs_ov_b, af0, abL, v0, vL,
        u_f, u_b, u_f2, u_b2, ax_f, ax_b, wa1f, wa1b, wa2f, wa2b;
    std::vector<double> nodeT, coefA, coefB, ext_u, ext_mass, ext_af, ext_ab, sc, energy_f, energy_b;
    std::vector<unsigned long long> nm_iters;
    std::vector<int32_t> status;
    std::vector<double> zoneT, zvol, zq, zinf_v, zinf_t, zvent_v, zvent_t;
    std::vector<int32_t> z_surf_begin, z_surf_end;
    DevView view(const BatchPlan& B);   // pointers into the host vectors (used by the test emulator)
};
void pack_arrays(const BatchPlan& B, HostArrays& H);
// scatter / gather between the reference-layout state vector and the packed buffers
void scatter_state(const BatchPlan& B, const double* state, double* nodeT, double* sc, double* zoneT);
void gather_state(const BatchPlan& B, const double* nodeT, const double* sc, const double* zoneT, double* state);
int64_t read_epw(const char* path, double* weather, double* inputs, int64_t capacity);

struct SeriesDev {
    int32_t n_steps;
    int32_t weather_mode;
    int32_t input_mode;
    int32_t ir_mode;
    const double* weather;   // [site][rows][3]
    int64_t weather_rows;
    const double* inputs;
    int64_t n_desc_surfaces;
    int32_t record_every;
    int32_t single_site;     // every model is on site 0: tiles share one weather table per main step
    double* zone_rec;        // [rec][n_dev_zones]
    double* node_rec;        // [rec][slab_len]
    double* sc_rec;          // [rec][8][n_dev_surfaces]
    int64_t slab_len;
    const double* zone_gain; // [n_steps][n_dev_zones] (device zone order) or nullptr: q of the zones during main step t
    int64_t n_dev_zones;
    // likewise infiltration / ventilation volume flow and temperature (model.rs:524-544), nullptr = constants stay
    const double* zone_inf_v;
    const double* zone_inf_t;
    const double* zone_vent_v;
    const double* zone_vent_t;
    // launches that take an SM per block (wt_isolated): block b sets started[b] = 1 when it begins — host-mapped memory
    // the host polls before it launches the bulk, so that such a block never waits for an SM to drain.  nullptr otherwise.
    volatile int32_t* started;
    // hb_wtile_kernel marches the main steps [step_begin, step_end) of the series (a group of short tiles is launched
    // as several such segments one after the other, launch_series); the other kernels always march [0, n_steps)
    int32_t step_begin, step_end;
};

// Warps per block of hb_wtile_kernel: the warps of a block are spread over the SM's four sub-partitions (one register
// file and one FP64 pipe each); every warp is an independent tile and never synchronises with the others.
constexpr int kWtWarps = 4;
// … and of its launches that take an SM per block (wt_isolated): two latency-bound warps per sub-partition
// (HB_WT_ISO_WARPS, diagnostic: 4 or 8)
constexpr int kWtWarpsMax = 8;
inline int wt_isolated_warps() {
    static const int w = std::getenv("HB_WT_ISO_WARPS") ? std::atoi(std::getenv("HB_WT_ISO_WARPS")) : kWtWarpsMax;
    return w == 4 ? 4 : kWtWarpsMax;
}
// blocks of such a launch: pairs (clusters of two, the two SMs of a TPC — see the end of hb_wtile_kernel)
inline int wt_isolated_blocks(int n_tiles) { return ((n_tiles + wt_isolated_warps() - 1) / wt_isolated_warps() + 1) & ~1; }

// A handful of warp tiles with the longest chains of a series END it (S x n_steps sub-steps in sequence): their launch
// asks for most of an SM's shared memory per block, so that such a block has its SM to itself instead of time-slicing the
// FP64 pipes with the throughput-bound bulk (HB_WT_ISOLATE=0 switches it off).  mode: 0 all-massive (never isolated).
inline bool wt_isolated(int mode, int s_max, int n_tiles) {
    static const bool on = !(std::getenv("HB_WT_ISOLATE") && std::getenv("HB_WT_ISOLATE")[0] == '0');
    return on && mode != 0 && s_max >= 96 && n_tiles <= 64;
}

enum { ST_NAN_H = 1, ST_NAN_ERR = 2, ST_NAN_ZONE = 4, ST_ITER_CAP = 8, ST_BAD_GAMMA = 16 };

}  // namespace hb

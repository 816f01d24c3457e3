// hb_dev.hpp — views of the packed SoA buffers (shared by the host packer, the launchers and the kernels).
#pragma once
#include <cstdint>
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
};

enum { ST_NAN_H = 1, ST_NAN_ERR = 2, ST_NAN_ZONE = 4, ST_ITER_CAP = 8, ST_BAD_GAMMA = 16 };

}  // namespace hb

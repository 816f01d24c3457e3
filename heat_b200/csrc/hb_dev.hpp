// hb_dev.hpp — device-side views of the SoA buffers (shared by the launchers and the kernels).
#pragma once
#include <cstdint>

#include "hb_internal.hpp"

namespace hb {

struct CavityDev {
    int32_t node, gas;
    double thickness, height, angle, eout, ein;
};

struct DevView {
    // per device surface slot
    const int32_t* n_nodes;
    const int32_t* flags;     // bits 0-3 front kind, 4-7 back kind, 8 distributed solar, 9 has cavity
    const int32_t* zone_f;    // tile-local zone index or -1
    const int32_t* zone_b;
    const int32_t* S;         // dt_subdivisions of the surface's model
    const int32_t* site;
    const int32_t* desc_index;  // descriptor-order surface index (for per-surface inputs)
    const double* dt;
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
    const int32_t* nm_off;    // CSR over no-mass chunks: pairs (ini, fin)
    const int32_t* nm_pairs;
    const int32_t* cav_off;   // CSR over cavity links
    const CavityDev* cav;
    // [node][kBlock] slabs per tile
    double* nodeT;
    const double* mass;
    const double* u;
    const double* alpha_f;
    const double* alpha_b;
    // state scalars [8][n_dev_surfaces]: h_f, h_b, flow_f, flow_b, solar_f, solar_b, ir_f, ir_b
    double* sc;
    int64_t n_dev_surfaces;
    double* energy_f;
    double* energy_b;
    unsigned long long* nm_iters;
    int32_t* status;
    // zones
    double* zoneT;
    const double* zvol;
    const double* zq;
    const double* zinf_v;
    const double* zinf_t;
    const double* zvent_v;
    const double* zvent_t;
    const double* zdt;
    const int32_t* zS;
    const int32_t* z_surf_begin;  // tile-local thread range of the zone's model
    const int32_t* z_surf_end;
    int64_t n_dev_zones;
    const Tile* tiles;
};

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
    double* zone_rec;        // [rec][n_dev_zones]
    double* node_rec;        // [rec][slab_len]
    double* sc_rec;          // [rec][8][n_dev_surfaces]
    int64_t slab_len;
};

enum { ST_NAN_H = 1, ST_NAN_ERR = 2, ST_NAN_ZONE = 4, ST_ITER_CAP = 8, ST_BAD_GAMMA = 16 };

}  // namespace hb

// hb_internal.hpp — shared declarations between the host packer, the CUDA kernels and the C ABI.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include "heat_b200.h"

namespace hb {

constexpr int kBlock = 128;          // threads per tile (= CUDA block); one thread per surface
constexpr int kMaxNomassChunk = 8;   // longest no-mass chunk the kernels solve
constexpr int kMaxNodes = 256;       // largest tier (tiers > 32 keep node arrays in local memory)

// Constants of the reference (kept in one place; cited where used)
constexpr double kSigma = 5.670374419e-8;  // lib.rs:49
constexpr double kMaxRs = 0.05;            // discretization.rs:21
constexpr double kMaxDx = 0.04;            // model.rs:236
constexpr double kMinDt = 60.0;            // model.rs:237
constexpr int kSafety = 2;                 // model.rs:329
constexpr double kDefaultEm = 0.84;        // surface.rs:449, discretization.rs:265
constexpr double kMassiveThreshold = 1e-5; // discretization.rs:149,155
constexpr double kInitialH = 1.739658084820765;  // surface_trait.rs:231,248
constexpr double kInitialT = 22.0;               // surface_trait.rs:365-370, zone.rs:45-50

struct CavityParams {   // cavity.rs:28-50
    int32_t node;       // link between node and node+1
    int32_t gas;
    double thickness, height, angle, eout, ein;
};

// Host-side result of Discretization::new + ThermalSurfaceData::new for one surface.
struct SurfacePlan {
    int32_t n_nodes = 0;
    int32_t tstep_subdivision = 1;
    std::vector<int32_t> n_elements;
    std::vector<double> mass;        // per node
    std::vector<double> u;           // per node: link to next node (solid k/dx; cavity -1; last 0)
    std::vector<CavityParams> cavities;
    std::vector<int32_t> massive_chunks, nomass_chunks;  // [ini,fin] pairs
    std::vector<double> front_alpha, back_alpha;
    double front_emissivity = 0, back_emissivity = 0;
    double wind_speed_modifier = 0;
    double cos_tilt = 0;
    bool distributed_solar = false;  // some alpha off the two end nodes
};

struct ModelPlan {
    int32_t zone_begin = 0, n_zones = 0;
    int32_t surface_begin = 0, n_surfaces = 0;
    int32_t dt_subdivisions = 1;
    double dt = 0;
    int64_t state_offset = 0;
    int32_t state_len = 0;
    int64_t nodes = 0;
    int32_t nmax = 0;
};

// A tile = the models one CUDA block marches. Threads ↔ surfaces, first n_zones threads ↔ zones.
struct Tile {
    int32_t surface_begin;   // index into the DEVICE ordering of surfaces
    int32_t n_surfaces;
    int32_t zone_begin;      // index into the DEVICE ordering of zones
    int32_t n_zones;
    int32_t nmax;            // max n_nodes in the tile
    int32_t smax;            // max dt_subdivisions in the tile
    int64_t node_offset;     // offset (in doubles) of the tile's [node][kBlock] slabs
};

struct BatchPlan {
    std::vector<hb_layer> layers;
    std::vector<int32_t> construction_offset;
    std::vector<hb_surface> surfaces;   // descriptor order
    std::vector<double> zone_volume;
    std::vector<int32_t> model_terrain;
    int32_t steps_per_hour = 1;
    std::vector<ModelPlan> models;      // descriptor order
    std::vector<SurfacePlan> plans;     // descriptor order (global surface index)
    std::vector<int64_t> surface_state_offset;  // descriptor order: index of front-h slot
    std::vector<int32_t> surface_model;         // descriptor order
    int64_t state_len_total = 0;
    int64_t nodes_total = 0;
    int64_t updates_per_step = 0;

    // device ordering
    std::vector<int32_t> model_order;      // device position → descriptor model index
    std::vector<int32_t> dev_surface_of;   // descriptor surface → device surface slot
    std::vector<int32_t> dev_zone_of;      // descriptor zone → device zone slot
    std::vector<Tile> tiles;
    int64_t node_slab_len = 0;             // total doubles in one [node][thread] slab array
    int64_t n_dev_surfaces = 0;            // tiles.size() * kBlock (padded)
    int64_t n_dev_zones = 0;
};

// Host discretization + planning (hb_host.cc). Throws std::runtime_error with the reference-style message.
void plan_surface(const hb_layer* layers, int n_layers, const hb_surface& s, int terrain, double main_dt,
                  SurfacePlan& out);
void plan_batch(const hb_batch_desc* desc, BatchPlan& out);

}  // namespace hb

// hb_internal.hpp — shared declarations between the host packer, the CUDA kernels and the C ABI.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include "heat_b200.h"

namespace hb {

// ---- device geometry of a tile (= one CUDA block) -----------------------------------------------------------
// A wall (surface) is marched by a LANE GROUP of G consecutive lanes, each lane holding kNL consecutive "slots" of
// the wall in registers.  Slot 0 is the front ghost node (equivalent outdoor/indoor air temperature), slots
// 1..n hold the n nodes, slot n+1 is the back ghost node, the rest is padding: n + 2 <= G * kNL.
// A tile holds W walls -> G*W threads.  One warp of the block additionally plays the SCALAR role: lane j of that
// warp evaluates the per-wall scalar physics (film coefficients, no-mass chunks, zone update) of wall j.
constexpr int kNL = 8;               // slots per lane
constexpr int kMaxG = 32;            // lanes per wall: 1, 2, 4, 8, 16, 32
constexpr int kMaxNodes = kMaxG * kNL - 2;  // 254: the largest wall of the regular tiers
// Walls beyond 254 nodes (thick, low-diffusivity layers under a short sub-step another surface of the model imposes)
// keep 32 lanes and take 16 or 32 slots per lane (hb_face_kernel only; spills to local memory, rare by construction).
constexpr int kNLBig = 16, kNLHuge = 32;
constexpr int kMaxNodesBig = kMaxG * kNLHuge - 2;  // 1022
constexpr int kMaxWalls = 32;        // walls (and zones) per tile; == warp size (scalar role)
constexpr int kMaxNomassChunk = 32;  // longest no-mass chunk the kernels solve (tridiagonal elimination in local arrays)
constexpr int kMaxSubsteps = 128;    // shared weather table rows (S <= 120 for n >= 1 step/hour)

// Tile widths.  walls_of_group: the largest tile of a tier (256 node lanes) and the width of the shared-memory-mirror
// kernel (hb_march_kernel).  pref_walls_of_group: what hb_face_kernel tiles use unless a model needs more slots —
// at most 128 node lanes (one warp per SM sub-partition: the shortest node phase) and at most 16 walls, so that the
// front and back faces of all walls fit in ONE warp (dense FP64 issue of the scalar chain).  A model needs all its
// surfaces and zones in one tile: tiles grow in powers of two up to walls_of_group(G) when a model asks for it.
inline constexpr int walls_of_group(int G) { return G <= 8 ? 32 : G == 16 ? 16 : 8; }
inline constexpr int pref_walls_of_group(int G) { return G == 1 ? 32 : G <= 8 ? 16 : 128 / G; }
inline int walls_for_model(int G, int n_surfaces, int n_zones) {
    int W = pref_walls_of_group(G);
    while (W < walls_of_group(G) && (W < n_surfaces || W < n_zones)) W *= 2;
    return W;
}
inline int group_of_nodes(int n) {
    int G = 1;
    while (G < kMaxG && G * kNL < n + 2) G *= 2;
    return G;
}
inline int slots_of_nodes(int n) {   // slots per lane of the tier that holds a wall of n nodes
    return n + 2 <= kMaxG * kNL ? kNL : n + 2 <= kMaxG * kNLBig ? kNLBig : kNLHuge;
}

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

// ---- warp tiles (hb_wtile_kernel): one warp = one tile of 32 / G walls, NL NODES per lane, no ghost slots ----------
// Geometries in order of capacity G·NL; a wall takes the first one that holds it (cost per wall and sub-step in FP64
// warp instructions ≈ (16.5·NL + 115·faces per lane)·G / 32 grows with the capacity, so the smallest fit is the cheapest).
struct WtGeom { int cap, G, NL; };
constexpr int kWtMaxNodes = 512;
// Smallest-capacity geometry (G lanes x NL nodes) that holds a wall of n nodes, NL taken from `nls` (ascending);
// ties go to the larger NL (fewer lanes per wall: the face work is shared by more walls).
// Geometries the kernels exist for: more than 16 nodes per lane (coefficients in shared memory, all-massive walls only)
// with at most 4 lanes per wall, 32 nodes per lane with one lane per wall only (registers).
inline bool wt_geom_exists(int G, int NL) { return NL <= 16 || (G <= 4 && (NL < 32 || G == 1)); }
inline bool wt_geom_of_nodes(int n, const std::vector<int>& nls, WtGeom& out, int min_G = 1) {
    bool found = false;
    for (int G = min_G; G <= 32; G *= 2)
        for (int NL : nls) {
            const int cap = G * NL;
            if (n > cap || cap > kWtMaxNodes || !wt_geom_exists(G, NL)) continue;
            if (!found || cap < out.cap || (cap == out.cap && NL > out.NL)) out = WtGeom{cap, G, NL};
            found = true;
        }
    return found;
}

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
    int32_t variant = 0;             // TILE_* flags needed by its surfaces
    int32_t walls = 0;               // wall slots of the tiles this model goes into
    int32_t NL = kNL;                // slots per lane of those tiles
    int32_t G = 1;                   // lanes per wall of those tiles
};

enum : int32_t {
    TILE_EXT = 1,    // some wall has no-mass chunks or cavities: node temperatures are mirrored in shared memory
    TILE_DIST = 2,   // some wall absorbs solar radiation on inner nodes (transparent layers)
    TILE_CAV = 4,    // some wall has a gas cavity (temperature-dependent link)
    TILE_END = 8,    // the only no-mass chunks are single nodes at the ends of walls (solved by the face lanes)
    TILE_PANE = 32,  // face-kernel tile with single-pane windows (two no-mass nodes solved jointly by the face lanes)
    TILE_SIMPLE = 64,  // face-kernel tile whose models are one wall with its own zone behind it (zone j <-> wall j)
    TILE_GEN = 16,   // face-kernel tile with a film-coefficient override or an AmbientTemperature back boundary
    TILE_WARP = 128, // warp tile (hb_wtile_kernel): G·W = 32, NL nodes per lane, node i in slot i (no ghost slots)
};

// A tile = the models one CUDA block marches.
struct Tile {
    int32_t surface_begin;   // index into the DEVICE ordering of surfaces (tile owns W consecutive slots)
    int32_t n_surfaces;
    int32_t zone_begin;      // index into the DEVICE ordering of zones (tile owns W consecutive slots)
    int32_t n_zones;
    int32_t nmax;            // max n_nodes in the tile
    int32_t S;               // dt_subdivisions (uniform over the tile)
    int32_t G;               // lanes per wall
    int32_t W;               // wall slots
    int32_t flags;           // TILE_*
    int32_t NL;              // slots per lane (kNL; kNLBig / kNLHuge for walls beyond kMaxNodes)
    int64_t node_offset;     // offset (doubles) of the tile's [NL][G*W] slabs (node T, coefficient A, coefficient B)
    int64_t ext_offset;      // offset (doubles) of the tile's [nmax][W] wall-major arrays (TILE_EXT only, else -1)
    double dt;               // sub-step length [s] (uniform over the tile)
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
    std::vector<int32_t> surface_tile;     // descriptor surface → tile index
    std::vector<int32_t> zone_tile;        // descriptor zone → tile index
    std::vector<Tile> tiles;
    int64_t node_slab_len = 0;             // total doubles in one slab array
    int64_t ext_len = 0;                   // total doubles in one wall-major ext array
    int64_t n_dev_surfaces = 0;            // Σ W over tiles (padded)
    int64_t n_dev_zones = 0;

    // position of node i of descriptor surface gd inside the slab arrays
    int64_t slab_index(int64_t gd, int node) const {
        const Tile& t = tiles[surface_tile[gd]];
        const int wl = dev_surface_of[gd] - t.surface_begin;
        const int slot = (t.flags & TILE_WARP) ? node : node + 1;   // block tiles keep slot 0 for the front ghost
        return t.node_offset + (int64_t)(slot % t.NL) * (t.G * t.W) + (int64_t)wl * t.G + slot / t.NL;
    }
};

// Host discretization + planning (hb_host.cc). Throws std::runtime_error with the reference-style message.
void plan_surface(const hb_layer* layers, int n_layers, const hb_surface& s, int terrain, double main_dt,
                  SurfacePlan& out);
void plan_batch(const hb_batch_desc* desc, BatchPlan& out, bool build_tiles = true);
// dt·λmax of the explicit march of a wall's massive chunks (see hb_surface_rk4_number in heat_b200.h)
double rk4_number(const SurfacePlan& P, double dt, double film);

}  // namespace hb

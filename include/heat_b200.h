/*
 * heat_b200.h — C ABI of the B200-native wall-conduction time-stepping path.
 *
 * Drop-in boundary for SIMPLE-BuildingSimulation/heat's `ThermalModel`
 * (reference: src/model.rs:188-428, trait communication_protocols::SimulationModel):
 *
 *   reference                                           this ABI
 *   ---------------------------------------------------------------------------------------
 *   ThermalModel::new(meta, (), model, header, n)        hb_create()            model.rs:215-354
 *   SimulationStateHeader::take_values()                 hb_get_state() after create (initial values)
 *   ThermalModel::allocate_memory()                      (inside hb_create; scratch is device memory) :193-207
 *   ThermalModel::march(date, weather, model, state, m)  hb_march_series(n_steps = 1)  model.rs:359-427
 *   caller reads/writes SimulationState slots            hb_set_state()/hb_get_state(), same slot order
 *   ThermalModel::dt_subdivisions(), .dt                 hb_model_info                  model.rs:73-76,434
 *   Discretization {segments,tstep_subdivision,n_elements} hb_surface_info/_arrays      discretization.rs:71-87
 *   Err(String) / panic!                                 hb_status != HB_OK + hb_last_error()
 *
 * plus the batched entry point the north star adds: one handle holds MANY independent models
 * (walls with their own zone, shoebox buildings, ...) and hb_march_series() advances all of them over
 * many main time steps in one persistent kernel launch.
 *
 * Host work that stays on the host (discretization.rs:95-544 — massive/no-mass classification, element
 * counts, tstep_subdivision; surface.rs:463-540 — chunks, solar alphas, wind modifier) is done inside
 * hb_create() in C++; the results are packed into SoA device buffers.  All compute after that runs in
 * hand-written CUDA for sm_100a.  There is NO CPU fallback: every entry point that needs the device
 * fails with HB_ERR_CUDA when no CUDA device is usable.
 *
 * Limits of this implementation where the reference has none (hb_create / hb_march_series return HB_ERR_INVALID with a
 * message; nothing is silently approximated):
 *   - a model's surfaces and zones must fit one tile: 32 surfaces / 32 zones when its largest wall has <= 62 nodes,
 *     16 up to 126 nodes, 8 beyond (ThermalModel::new, model.rs:245-323, loops over any number);
 *   - a no-mass chunk (consecutive no-mass nodes, anywhere in a wall) may have at most 32 nodes (surface.rs:790-898
 *     solves any length); chunks of one or two nodes at the ends of walls, single panes and pane | gas | pane windows
 *     are on the fast path, everything else on the general kernel;
 *   - at most 128 sub-steps per main step (n >= 1 step per hour gives <= 120: min_dt = 60 s, model.rs:237,326-331);
 *   - walls of more than 1022 nodes; more than 254 nodes only for all-massive walls / walls with no-mass ends.
 * Not limits: the no-mass fixed point gives up where the reference does (99 199 000 passes, surface.rs:869-873 —
 * HB_ERR_PHYSICS "Excessive number of iterations"), and a wall on which the reference's explicit RK4 is unstable at the
 * sub-step its discretization rule picks (see hb_surface_rk4_number) ends, as in the reference, in NaN → HB_ERR_PHYSICS.
 * A physics failure names the surface (batch order), its model and the device slot.
 *
 * Conventions: plain pointers and sizes only; all `double*` are HOST pointers unless the name ends in
 * `_dev`; a handle is not re-entrant; no global state except the thread-local last-error string.
 */
#ifndef HEAT_B200_H
#define HEAT_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef int32_t hb_status;
enum {
    HB_OK = 0,
    HB_ERR_INVALID = 1,     /* bad descriptor / argument (reference: Err(String)) */
    HB_ERR_CUDA = 2,        /* CUDA runtime error or no device */
    HB_ERR_PHYSICS = 3,     /* a wall reported NaN / unsupported boundary (reference: panic!) */
    HB_ERR_UNSUPPORTED = 4  /* e.g. Boundary::Ground (reference: unimplemented!, model.rs:92) */
};

/* simple_model::Boundary as matched at surface.rs:611-702 */
enum { HB_BOUNDARY_OUTDOOR = 0, HB_BOUNDARY_SPACE = 1, HB_BOUNDARY_AMBIENT = 2, HB_BOUNDARY_GROUND = 3 };
/* simple_model::Substance::{Normal, Gas} */
enum { HB_LAYER_SOLID = 0, HB_LAYER_GAS = 1 };
/* gas.rs:45-74 */
enum { HB_GAS_AIR = 0, HB_GAS_ARGON = 1, HB_GAS_KRYPTON = 2, HB_GAS_XENON = 3 };
/* simple_model::TerrainClass, surface.rs:151-163 (NONE: no SiteDetails → Urban; UNSET: details without terrain) */
enum { HB_TERRAIN_NONE = -1, HB_TERRAIN_COUNTRY = 0, HB_TERRAIN_SUBURBS = 1, HB_TERRAIN_CITY = 2,
       HB_TERRAIN_OCEAN = 3, HB_TERRAIN_URBAN = 4, HB_TERRAIN_UNSET = 5 };

/* One material layer = simple_model Material + its Substance. NaN in an optional field = "not set". */
typedef struct {
    int32_t kind;                 /* HB_LAYER_* */
    int32_t gas;                  /* HB_GAS_* when kind == HB_LAYER_GAS */
    double thickness;             /* [m] */
    double conductivity;          /* [W/m.K] */
    double density;               /* [kg/m3] */
    double specific_heat;         /* [J/kg.K] */
    double front_emissivity;      /* default 0.84 (surface.rs:449) */
    double back_emissivity;       /* default 0.84 */
    double front_solar_abs;       /* default 0.84 (glazing.rs:86) */
    double back_solar_abs;        /* default 0.84 (glazing.rs:87) */
    double solar_transmittance;   /* default 0    (glazing.rs:85) */
} hb_layer;

/* One simple_model Surface or Fenestration with what ThermalModel::new derives from its geometry
 * (model.rs:248-256).  Zone indices are LOCAL to the surface's model. */
typedef struct {
    int32_t construction;
    int32_t is_fenestration;      /* within a model: all surfaces first, then all fenestrations */
    int32_t front_kind, back_kind;
    int32_t front_zone, back_zone;
    double front_ambient_t, back_ambient_t;
    double area, perimeter;
    double centroid_z;
    double normal[3];
    double front_hs_override;     /* NaN = none; mirrors the debug-only `front_hs: Option<Float>` (surface.rs:372-380) */
    double back_hs_override;
    double front_emissivity_override; /* NaN = none; mirrors writes to the pub field (validate_wall_heat_transfer.rs:630) */
    double back_emissivity_override;
} hb_surface;

/* A batch of independent models sharing one pool of constructions. */
typedef struct {
    int32_t n_layers;
    const hb_layer* layers;
    int32_t n_constructions;
    const int32_t* construction_offset;   /* n_constructions + 1 */
    int32_t n_models;
    const int32_t* model_zone_offset;     /* n_models + 1 */
    const int32_t* model_surface_offset;  /* n_models + 1 */
    const double* zone_volume;            /* [total zones] */
    const hb_surface* surfaces;           /* [total surfaces] */
    const int32_t* model_terrain;         /* [n_models] HB_TERRAIN_* or NULL (= NONE) */
    int32_t steps_per_hour;               /* `n` of ThermalModel::new */
} hb_batch_desc;

typedef struct hb_handle hb_handle;

const char* hb_last_error(void);
const char* hb_version(void);

/* ThermalModel::new for every model of the batch, on CUDA device `device`.
 * device == HB_DEVICE_NONE builds a host-only planning handle: the discretization/layout queries below work,
 * every entry point that touches the device returns HB_ERR_CUDA (there is no CPU fallback). */
#define HB_DEVICE_NONE (-1)
hb_status hb_create(const hb_batch_desc* desc, int32_t device, hb_handle** out);
void hb_destroy(hb_handle* h);

/* ---- layout / discretization queries ---- */
typedef struct {
    int32_t n_zones, n_surfaces;          /* of this model (surfaces + fenestrations) */
    int32_t zone_begin, surface_begin;    /* global indices */
    int32_t dt_subdivisions;              /* substeps per main step, model.rs:331 */
    double dt;                            /* [s] model.rs:326-330 */
    int64_t state_offset;                 /* into the concatenated state vector */
    int32_t state_len;                    /* reference SimulationState length for this model */
    int64_t nodes;                        /* Σ n_nodes over its surfaces */
} hb_model_info;

typedef struct {
    int32_t n_nodes;
    int32_t n_layers;
    int32_t tstep_subdivision;            /* discretization.rs:83 */
    int32_t n_massive_chunks, n_nomass_chunks;
    int64_t state_offset;                 /* index of the front-h slot in the concatenated state vector */
    double front_emissivity, back_emissivity;
    double wind_speed_modifier;
    double cos_tilt;
} hb_surface_info;

int32_t hb_n_models(const hb_handle* h);
int64_t hb_n_surfaces_total(const hb_handle* h);
int64_t hb_n_zones_total(const hb_handle* h);
int64_t hb_state_len_total(const hb_handle* h);
int64_t hb_nodes_total(const hb_handle* h);
/* Σ over surfaces of n_nodes × dt_subdivisions(model): node-substep updates done by ONE main step */
int64_t hb_updates_per_step(const hb_handle* h);
hb_status hb_get_model_info(const hb_handle* h, int32_t model, hb_model_info* out);
hb_status hb_get_surface_info(const hb_handle* h, int64_t surface, hb_surface_info* out);
/* arrays of the surface: n_elements[n_layers]; mass[n_nodes]; u[n_nodes] (solid k/dx, cavity -1, last 0);
 * front_alpha/back_alpha[n_nodes]; massive/nomass chunks as [ini,fin] pairs. Any pointer may be NULL. */
hb_status hb_get_surface_arrays(const hb_handle* h, int64_t surface, int32_t* n_elements, double* mass, double* u,
                                double* front_alpha, double* back_alpha, int32_t* massive_chunks,
                                int32_t* nomass_chunks);

/* Diagnostic with no counterpart in the reference: dt·λmax(C⁻¹K) of the surface's massive chunks at the model's
 * sub-step dt, films of `film` W/m²K (convective + linearised IR) on both faces.  The reference's explicit RK4
 * (surface.rs:228-308) is bounded only below 2.7853; its discretization rule (discretization.rs:95-544) looks at whole
 * layers and lets a few stacks through whose half-cell interface nodes exceed that — such a wall marches to NaN in the
 * reference ("Future temperatures is NaN" / NaN film coefficients) and, identically, here (HB_ERR_PHYSICS).  Generators
 * of synthetic ensembles use this to re-draw them (heat_b200/cases.py). */
hb_status hb_surface_rk4_number(const hb_handle* h, int64_t surface, double film, double* out);

/* ---- state: reference slot order, models concatenated ----
 * per model: [zone T × n_zones] then per surface (surfaces, then fenestrations):
 * front h, back h, front conv. flow, back conv. flow, front solar irr., back solar irr., front IR irr.,
 * back IR irr., node T × n_nodes   (zone.rs:43-50, surface.rs:428-442) */
hb_status hb_set_state(hb_handle* h, const double* state);
hb_status hb_get_state(hb_handle* h, double* state);

/* Per-zone inputs that are SimulationState elements in the reference (model.rs:502-544):
 * Σ HVAC+luminaire power, infiltration / ventilation volume flow and temperature (NaN temperature = absent).
 * Arrays over all zones of the batch; NULL = none. */
hb_status hb_set_zone_inputs(hb_handle* h, const double* q_internal, const double* infiltration_volume,
                             const double* infiltration_temperature, const double* ventilation_volume,
                             const double* ventilation_temperature);

/* ---- marching ---- */
enum {
    HB_WEATHER_STEP_CONSTANT = 0, /* row t holds for every substep of main step t (validate driver, :676-679) */
    HB_WEATHER_LINEAR = 1,        /* substep s of step t samples row[t] + (row[t+1]-row[t])·(s+1)/S; n_steps+1 rows */
    HB_WEATHER_PER_SUBSTEP = 2    /* rows [t][s], s < S; all models of a site must share S (arbitrary `Weather` impl) */
};
enum {
    HB_INPUT_STATE = 0,           /* keep the irradiance slots as they are in the state */
    HB_INPUT_SITE = 1,            /* per site rows [t][4] = front solar, back solar, front IR, back IR, written
                                     into every surface of the site's models at the start of main step t */
    HB_INPUT_PER_SURFACE = 2      /* rows [t][surface][4] */
};
enum {
    HB_IR_IRRADIANCE = 0,         /* IR columns are irradiances [W/m2] */
    HB_IR_NET_GAIN = 1            /* IR columns are net thermal radiation gains [W]: irradiance =
                                     gain/area/ε + σ(T_face+273.15)^4 evaluated at the start of the main step,
                                     skipped when ε ≤ 1e-3 (validate_wall_heat_transfer.rs:689-699) */
};

typedef struct {
    int32_t n_steps;              /* main time steps to advance */
    int32_t weather_mode;         /* HB_WEATHER_* */
    int32_t n_sites;              /* number of weather tables (≥ 1) */
    const int32_t* model_site;    /* [n_models] or NULL (all models on site 0) */
    const double* weather;        /* [site][row][3]: dry bulb [C], wind direction [deg], wind speed [m/s] */
    int64_t weather_rows;         /* rows per site */
    int32_t input_mode;           /* HB_INPUT_* */
    int32_t ir_mode;              /* HB_IR_* */
    const double* inputs;         /* per mode above; NULL for HB_INPUT_STATE */
    int32_t record_every;         /* 0: no recording; k: record after every k-th main step */
    double* zone_t_record;        /* [n_steps/record_every][total zones] or NULL */
    double* state_record;         /* [n_steps/record_every][state_len_total] or NULL (small batches only) */
    const double* zone_gain;      /* schedule of the zones' HVAC + luminaire power (model.rs:502-516): HOST table
                                     [n_steps][total zones] [W], row t replaces q_internal of hb_set_zone_inputs during
                                     main step t (the reference reads these state elements once per sub-step and the
                                     caller rewrites them between march calls); NULL = the constants stay.  Host
                                     memory in hb_march_series AND hb_march_series_dev. */
    /* Schedules of the zones' infiltration and ventilation (model.rs:524-544: volume flow [m3/s] and temperature [C]
     * of each stream are state elements the reference reads every sub-step and the caller rewrites between march
     * calls): HOST tables [n_steps][total zones]; row t replaces the constants of hb_set_zone_inputs during main step
     * t; a NaN temperature = element absent in that step.  NULL = the constant stays.  A temperature table needs its
     * volume table (or a volume set with hb_set_zone_inputs). */
    const double* zone_infiltration_volume;
    const double* zone_infiltration_temperature;
    const double* zone_ventilation_volume;
    const double* zone_ventilation_temperature;
} hb_series;

/* Advance every model by series->n_steps main steps (each = dt_subdivisions substeps, model.rs:369-424).
 * Host buffers in, host buffers out; the state stays resident on the device between calls. */
hb_status hb_march_series(hb_handle* h, const hb_series* series);

/* Same, with inputs already resident on the device (weather_dev/inputs_dev uploaded once with hb_upload) and
 * no recording; `elapsed_ms` (may be NULL) receives the CUDA-event time of the kernel(s) on the handle's stream. */
hb_status hb_upload(hb_handle* h, const double* host, int64_t count, double** dev_out);
hb_status hb_free_dev(hb_handle* h, double* dev);
hb_status hb_march_series_dev(hb_handle* h, const hb_series* series_with_dev_pointers, float* elapsed_ms);

/* Number of kernel launches issued by this handle so far (bench.py's gpu_launches). */
int64_t hb_launch_count(const hb_handle* h);

/* ---- ensemble reductions (K3): Σ front/back convective energy is integrated in-kernel ---- */
typedef struct {
    double sum_zone_t, min_zone_t, max_zone_t;
    double sum_node_t, min_node_t, max_node_t;
    double front_conv_energy_j;   /* Σ_surfaces Σ_substeps flow_front·A·dt since create/reset (model.rs:161) */
    double back_conv_energy_j;
    int64_t n_zones, n_nodes;
    int64_t nomass_iterations;    /* total fixed-point iterations (surface.rs:808-896), branch-parity diagnostic */
    int64_t status_flags;         /* OR of per-surface status words (non-zero = HB_ERR_PHYSICS) */
} hb_ensemble_stats;
/* Local (this handle / this GPU) statistics. */
hb_status hb_reduce(hb_handle* h, hb_ensemble_stats* out);
hb_status hb_reset_energy(hb_handle* h);

/* hb_reduce for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  No counterpart in the reference.
 * Same statistics as hb_reduce, from a variant of the K3 kernel whose scratch — one 96-byte partial per tile, folded in
 * tile order by a second one-block kernel, so the sums do not depend on atomic order — is a file-scope device array of
 * (HB_REDUCE_NOALLOC_MAX_TILES + 1) x 96 bytes = 1.5 MiB that every GPU loading the library holds until exit.  The call
 * makes no cudaMalloc / cudaFree and uses no allocating library routine; the first call on a device also resolves the
 * array's device address (cudaGetSymbolAddress) — make that one outside a capture.
 *   out_host  host memory, filled after a stream synchronize; may be NULL
 *   out_dev   DEVICE memory of the caller (sizeof(hb_ensemble_stats), 8-byte aligned), written by the kernel with no
 *             synchronize — ordered after the handle's earlier work; may be NULL (not both)
 * Limit: a batch of more than HB_REDUCE_NOALLOC_MAX_TILES tiles (about 3.4·10^5 walls of configs[2]) is rejected with
 * HB_ERR_INVALID — use hb_reduce.  All calls, on all handles of a device, share the one array, which keeps the last
 * call's partials (every word a call reads it has written first: nothing needs clearing): calls that overlap on the
 * GPU — two handles, two host threads — race; serialise them. */
#define HB_REDUCE_NOALLOC_MAX_TILES 16384
hb_status hb_reduce_noalloc(hb_handle* h, hb_ensemble_stats* out_host, hb_ensemble_stats* out_dev);

/* ---- the path's only collective: the ensemble statistics over all GPUs of a job ----
 * One process per GPU, one handle per process; walls and buildings never exchange anything while they march
 * (a building's surfaces couple only through its own zone, model.rs:556-590).  NCCL (libnccl.so.2, resolved at run
 * time) lives INSIDE the library: a Rust / C host needs no NCCL binding of its own.
 *   rank 0:      hb_comm_unique_id(id)  → the host ships the 128 bytes to the other ranks (file, socket, MPI, ...)
 *   every rank:  hb_comm_init(h, id, rank, world)      (ncclCommInitRank on the handle's device)
 *   every rank:  hb_reduce_all(h, &stats)               (hb_reduce + ncclAllReduce: sums, minima, maxima; the status
 *                                                        words are OR-ed bit by bit) — every rank gets the same result
 * Without hb_comm_init (or with world == 1) hb_reduce_all equals hb_reduce. */
#define HB_COMM_ID_BYTES 128
hb_status hb_comm_unique_id(void* id_out);
hb_status hb_comm_init(hb_handle* h, const void* id, int32_t rank, int32_t world);
hb_status hb_comm_destroy(hb_handle* h);
hb_status hb_reduce_all(hb_handle* h, hb_ensemble_stats* out);

/* ---- EPW ingestion (host only; the reference reads weather through the `weather` crate, model.rs:372) ----
 * Hourly records of an EnergyPlus weather file in the layouts hb_series consumes:
 * weather [row][3] = dry bulb [C] (EPW field 7), wind direction [deg] (21), wind speed [m/s] (22);
 * inputs  [row][4] = global horizontal [W/m2] (14) as front solar, 0, horizontal infrared [W/m2] (13) as front IR, 0
 * (HB_INPUT_SITE + HB_IR_IRRADIANCE).  Either array may be NULL; at most `capacity` rows are written, `*n_rows`
 * receives the number of records in the file (call with capacity 0 to size the buffers).  For a periodic year with
 * HB_WEATHER_LINEAR append row 0 after the last one. */
hb_status hb_read_epw(const char* path, double* weather, double* inputs, int64_t capacity, int64_t* n_rows);

/* ---- device probes used by bench.py ---- */
/* Dependent-free DFMA micro-benchmark: returns measured FP64 TFLOP/s (FMA = 2 FLOP) on `device`. */
hb_status hb_measure_fp64_peak(int32_t device, double* tflops);
/* Writes `bytes` of device memory (L2 flush between timed iterations). */
hb_status hb_flush_l2(hb_handle* h, int64_t bytes);

#ifdef __cplusplus
}
#endif
#endif

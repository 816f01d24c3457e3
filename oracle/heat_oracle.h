/*
 * heat_oracle.h — C interface of the CPU ORACLE (TEST INFRASTRUCTURE ONLY).
 *
 * This library is a CPU restatement of the wall-conduction time-stepping path of
 * SIMPLE-BuildingSimulation/heat (reference: /root/reference, Rust, f64).  It is the
 * checker for the CUDA path in heat_b200/; it is never shipped, never linked by the
 * product and never measured as the product.  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may load it.
 *
 * PARITY STATUS: the reference is a Rust crate with 10 un-vendored git dependencies
 * and no Rust toolchain exists in this environment, so the reference itself cannot be
 * executed here.  The oracle is pinned against every asserted known-answer test the
 * reference holds for this path (tests/test_oracle_kat.py lists them with file:line);
 * pieces that no reference test pins are marked "parity unpinned" below and in
 * DESIGN.md (discretize_construction integers, `matrix` crate rounding order,
 * `weather` interpolation).
 *
 * The flat model description mirrors what the reference reads out of `SimpleModel`
 * (simple_model crate, un-vendored): substances+materials are merged into layers,
 * constructions are layer ranges, spaces are zones.
 */
#ifndef HEAT_ORACLE_H
#define HEAT_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Boundary kinds — simple_model::Boundary as matched at src/surface.rs:611-702 */
enum { HO_BOUNDARY_OUTDOOR = 0, HO_BOUNDARY_SPACE = 1, HO_BOUNDARY_AMBIENT = 2, HO_BOUNDARY_GROUND = 3 };
/* Layer kinds — simple_model::Substance::{Normal,Gas} */
enum { HO_LAYER_SOLID = 0, HO_LAYER_GAS = 1 };
/* Gases — src/gas.rs:45-74 */
enum { HO_GAS_AIR = 0, HO_GAS_ARGON = 1, HO_GAS_KRYPTON = 2, HO_GAS_XENON = 3 };
/* Terrain — src/surface.rs:151-157; NONE = no site details → Urban default (:159-163);
 * UNSET = site details present but no terrain → alpha=delta=0 (:142-158) */
enum { HO_TERRAIN_NONE = -1, HO_TERRAIN_COUNTRY = 0, HO_TERRAIN_SUBURBS = 1, HO_TERRAIN_CITY = 2,
       HO_TERRAIN_OCEAN = 3, HO_TERRAIN_URBAN = 4, HO_TERRAIN_UNSET = 5 };

typedef struct {
    int32_t kind;                 /* HO_LAYER_* */
    int32_t gas;                  /* HO_GAS_* when kind == GAS */
    double thickness;             /* Material::thickness [m] */
    double conductivity;          /* [W/m.K] */
    double density;               /* [kg/m3] */
    double specific_heat;         /* [J/kg.K] */
    double front_emissivity;      /* front_thermal_absorbtance; NaN → default 0.84 */
    double back_emissivity;       /* back_thermal_absorbtance;  NaN → default 0.84 */
    double front_solar_abs;       /* NaN → default 0.84 (src/glazing.rs:86) */
    double back_solar_abs;        /* NaN → default 0.84 (src/glazing.rs:87) */
    double solar_transmittance;   /* NaN → 0 (src/glazing.rs:85) */
} ho_layer;

typedef struct {
    int32_t construction;         /* index into construction_offset */
    int32_t is_fenestration;      /* 0 = Surface, 1 = Fenestration (state order: surfaces first) */
    int32_t front_kind, back_kind;/* HO_BOUNDARY_* */
    int32_t front_zone, back_zone;/* zone index when kind == SPACE */
    double front_ambient_t;       /* AmbientTemperature{temperature} */
    double back_ambient_t;
    double area, perimeter;       /* src/model.rs:254-255 */
    double centroid_z;            /* height for wind_speed_modifier, src/model.rs:272 */
    double normal[3];             /* unit normal, src/model.rs:248 */
    double front_hs_override;     /* NaN = none. debug_assertions-only field src/surface.rs:372-380 */
    double back_hs_override;
    double front_emissivity_override; /* NaN = none (tests write pub field, validate_wall_heat_transfer.rs:630) */
    double back_emissivity_override;
} ho_surface;

typedef struct {
    int32_t n_layers;
    const ho_layer* layers;
    int32_t n_constructions;
    const int32_t* construction_offset;   /* n_constructions + 1 entries into layers */
    int32_t n_zones;
    const double* zone_volume;
    int32_t n_surfaces;                   /* surfaces followed by fenestrations */
    const ho_surface* surfaces;
    int32_t terrain;                      /* HO_TERRAIN_* */
    int32_t steps_per_hour;               /* `n` of ThermalModel::new, src/model.rs:220 */
} ho_model_desc;

typedef struct ho_model ho_model;

/* ThermalModel::new (src/model.rs:215-354).  Returns NULL on error (see ho_last_error). */
ho_model* ho_model_new(const ho_model_desc* desc);
/* Same with the arguments the reference's unit tests hand to Discretization::new (src/surface.rs:1124-1127 etc.):
 * main_dt_override <= 0 keeps main_dt = 3600/steps_per_hour. */
ho_model* ho_model_new_ex(const ho_model_desc* desc, double main_dt_override, double max_dx, double min_dt);
void ho_model_free(ho_model* m);
const char* ho_last_error(void);

/* Layout queries */
int32_t ho_dt_subdivisions(const ho_model* m);        /* src/model.rs:434 */
double ho_dt(const ho_model* m);
int32_t ho_state_len(const ho_model* m);
int32_t ho_surface_n_nodes(const ho_model* m, int32_t surface);
int32_t ho_surface_state_offset(const ho_model* m, int32_t surface);  /* index of front h slot */
int32_t ho_surface_tstep_subdivision(const ho_model* m, int32_t surface);
/* n_elements per layer of the surface's construction; returns count written */
int32_t ho_surface_n_elements(const ho_model* m, int32_t surface, int32_t* out, int32_t cap);
/* node masses and link U (solid: k/dx, cavity: -1, back: 0); arrays of n_nodes */
int32_t ho_surface_segments(const ho_model* m, int32_t surface, double* mass, double* u);
/* chunk ranges: out = [ini0,fin0,ini1,fin1,...]; returns number of chunks */
int32_t ho_surface_chunks(const ho_model* m, int32_t surface, int32_t massive, int32_t* out, int32_t cap);
int32_t ho_surface_alphas(const ho_model* m, int32_t surface, double* front, double* back);
double ho_surface_wind_modifier(const ho_model* m, int32_t surface);
double ho_surface_emissivity(const ho_model* m, int32_t surface, int32_t back);

/* Initial state (SimulationStateHeader::take_values analogue): zone T = 22, h = 1.739658084820765,
 * flows = irradiances = 0, node T = 22 (src/zone.rs:45-50, src/surface_trait.rs:223-378). */
void ho_initial_state(const ho_model* m, double* state);

/* Per-zone inputs that live in SimulationState in the reference (src/model.rs:502-544).
 * Arrays of n_zones; NaN temperature = element absent. May be NULL (= none). */
typedef struct {
    const double* q_internal;     /* Σ HVAC + luminaire power [W] */
    const double* infiltration_volume;
    const double* infiltration_temperature;
    const double* ventilation_volume;
    const double* ventilation_temperature;
} ho_zone_inputs;

/* ThermalModel::march (src/model.rs:359-427): one main timestep = dt_subdivisions substeps.
 * weather: dt_subdivisions rows of {dry_bulb [C], wind_direction [deg], wind_speed [m/s]},
 * row s = Weather::get_weather_data(date after (s+1)*dt).  Returns 0 on success. */
int32_t ho_march(const ho_model* m, double* state, const double* weather, const ho_zone_inputs* zin);

/* iterate_surfaces only (src/model.rs:102-180), used by the reference unit tests in src/surface.rs. */
int32_t ho_iterate_surfaces(const ho_model* m, double* state, double wind_direction_rad, double wind_speed,
                            double t_out, double dt);
/* calculate_zones_abc (src/model.rs:489-597) */
int32_t ho_zones_abc(const ho_model* m, const double* state, const ho_zone_inputs* zin, double* a, double* b, double* c);

/* Count of no-mass fixed-point iterations executed so far (diagnostic for branch-parity reports) */
int64_t ho_nomass_iterations(const ho_model* m);

/* ---- pure-function KAT hooks ---- */
double ho_nusselt(double ra, double gamma, double a_gi);                       /* src/gas.rs:197-221 */
double ho_gas_property(int32_t gas, int32_t which, double temp_k);             /* 0 λ,1 μ,2 cp,3 M,4 ρ */
double ho_cavity_u_value(int32_t gas, double thickness, double height, double angle, double eout, double ein,
                         double t_front, double t_back);                       /* src/cavity.rs:59-69 */
double ho_tarp_natural(double t_air, double t_surf, double cos_tilt);          /* src/convection.rs:87-110 */
double ho_tarp_forced_plus_natural(double t_air, double t_surf, double cos_tilt, double air_speed,
                                   double area, double perimeter, int32_t windward); /* :151-168 */
int32_t ho_is_windward(double wind_direction_rad, double cos_tilt, const double* normal); /* src/surface.rs:37-46 */
double ho_wind_speed_modifier(double height, int32_t terrain);                 /* src/surface.rs:135-166 */
/* rearrange_k + n× rk4 on a dense system (src/surface.rs:168-308); k row-major n×n, c diagonal n */
void ho_rk4_steps(int32_t n, double* k, const double* c, double* q, double* temps, double dt, int32_t nsteps,
                  int32_t rearrange);
/* discretize_construction + build on a bare construction (src/discretization.rs:95-114) */
int32_t ho_discretize(const ho_layer* layers, int32_t n_layers, double model_dt, double max_dx, double min_dt,
                      double height, double angle, int32_t* tstep_subdivision, int32_t* n_elements,
                      double* mass, double* u, int32_t cap_nodes);
/* Discretization::build on GIVEN element counts (src/discretization.rs:163-298); returns the node count */
int32_t ho_build(const ho_layer* layers, int32_t n_layers, int32_t tstep_subdivision, const int32_t* n_elements,
                 double height, double angle, double* mass, double* u, int32_t cap_nodes);
/* Discretization::get_chunks on bare node masses (src/discretization.rs:117-160); returns the chunk count */
int32_t ho_chunks(const double* mass, int32_t n, int32_t massive, int32_t* out, int32_t cap);
/* Discretization::get_k_q on solid segments (mass[i], Solid(u[i])), last Back (src/discretization.rs:596-700);
 * k_out dense (fin-ini)^2 row-major, q_out fin-ini */
int32_t ho_get_k_q(int32_t n, const double* mass, const double* u, int32_t ini, int32_t fin, const double* temps,
                   double front_air, double front_rad_t, double front_hs, double front_rad_hs, double back_air,
                   double back_rad_t, double back_hs, double back_rad_hs, double* k_out, double* q_out);
/* ISO 9050 combination of n panes (src/glazing.rs:180-230): out = tau, rho_f, rho_b, alpha_f, alpha_b */
int32_t ho_glazing_combine(const double* tau, const double* rho_front, const double* rho_back, int32_t n, double* out);
/* glazing alphas (src/glazing.rs:259-286) for tau/rho_front/rho_back triples */
int32_t ho_glazing_alphas(const double* tau, const double* rho_front, const double* rho_back, int32_t n, double* out);

/* ---- CPU baseline: run `n_models` independent copies of the hot path on `threads` host threads ----
 * Every model marches `n_steps` main steps with the series inputs below; returns node-substep updates done.
 * series layout: weather[step][3] shared by all models; surf_inputs[step][4] = front solar, back solar,
 * front IR, back IR applied to every surface (NULL = zeros). */
int64_t ho_run_ensemble(ho_model* const* models, int32_t n_models, int32_t n_steps, const double* weather,
                        const double* surf_inputs, int32_t weather_linear, int32_t threads,
                        double* final_state_concat);
/* … with the zones' HVAC + luminaire power [W] per zone, models concatenated (NULL = none) */
int64_t ho_run_ensemble_q(ho_model* const* models, int32_t n_models, int32_t n_steps, const double* weather,
                          const double* surf_inputs, int32_t weather_linear, int32_t threads,
                          double* final_state_concat, const double* q_internal_concat);

#ifdef __cplusplus
}
#endif
#endif

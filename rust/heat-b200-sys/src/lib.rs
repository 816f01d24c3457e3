// UNCOMPILED IN THIS ENVIRONMENT (no cargo / rustc in the image).
// heat-b200-sys/src/lib.rs — 1:1 with include/heat_b200.h.  The C side's struct layouts are asserted by
// tests/test_c_abi.py; tests/test_rust_bindings.py parses the extern "C" block below and checks every name and
// argument count against the header.
#![allow(non_camel_case_types)]
use std::os::raw::{c_char, c_double, c_float, c_void};

pub type hb_status = i32;
pub const HB_OK: hb_status = 0;
pub const HB_ERR_INVALID: hb_status = 1;
pub const HB_ERR_CUDA: hb_status = 2;
pub const HB_ERR_PHYSICS: hb_status = 3;
pub const HB_ERR_UNSUPPORTED: hb_status = 4;
pub const HB_DEVICE_NONE: i32 = -1;
pub const HB_REDUCE_NOALLOC_MAX_TILES: usize = 16384;
pub const HB_COMM_ID_BYTES: usize = 128;

pub const HB_BOUNDARY_OUTDOOR: i32 = 0;
pub const HB_BOUNDARY_SPACE: i32 = 1;
pub const HB_BOUNDARY_AMBIENT: i32 = 2;
pub const HB_BOUNDARY_GROUND: i32 = 3;
pub const HB_LAYER_SOLID: i32 = 0;
pub const HB_LAYER_GAS: i32 = 1;
pub const HB_WEATHER_STEP_CONSTANT: i32 = 0;
pub const HB_WEATHER_LINEAR: i32 = 1;
pub const HB_WEATHER_PER_SUBSTEP: i32 = 2;
pub const HB_INPUT_STATE: i32 = 0;
pub const HB_INPUT_SITE: i32 = 1;
pub const HB_INPUT_PER_SURFACE: i32 = 2;
pub const HB_IR_IRRADIANCE: i32 = 0;
pub const HB_IR_NET_GAIN: i32 = 1;

#[repr(C)]
#[derive(Clone, Copy)]
pub struct hb_layer {
    pub kind: i32,
    pub gas: i32,
    pub thickness: c_double,
    pub conductivity: c_double,
    pub density: c_double,
    pub specific_heat: c_double,
    pub front_emissivity: c_double,
    pub back_emissivity: c_double,
    pub front_solar_abs: c_double,
    pub back_solar_abs: c_double,
    pub solar_transmittance: c_double,
}

#[repr(C)]
#[derive(Clone, Copy)]
pub struct hb_surface {
    pub construction: i32,
    pub is_fenestration: i32,
    pub front_kind: i32,
    pub back_kind: i32,
    pub front_zone: i32,
    pub back_zone: i32,
    pub front_ambient_t: c_double,
    pub back_ambient_t: c_double,
    pub area: c_double,
    pub perimeter: c_double,
    pub centroid_z: c_double,
    pub normal: [c_double; 3],
    pub front_hs_override: c_double,
    pub back_hs_override: c_double,
    pub front_emissivity_override: c_double,
    pub back_emissivity_override: c_double,
}

#[repr(C)]
pub struct hb_batch_desc {
    pub n_layers: i32,
    pub layers: *const hb_layer,
    pub n_constructions: i32,
    pub construction_offset: *const i32,
    pub n_models: i32,
    pub model_zone_offset: *const i32,
    pub model_surface_offset: *const i32,
    pub zone_volume: *const c_double,
    pub surfaces: *const hb_surface,
    pub model_terrain: *const i32,
    pub steps_per_hour: i32,
}

#[repr(C)]
#[derive(Default, Clone, Copy)]
pub struct hb_model_info {
    pub n_zones: i32,
    pub n_surfaces: i32,
    pub zone_begin: i32,
    pub surface_begin: i32,
    pub dt_subdivisions: i32,
    pub dt: c_double,
    pub state_offset: i64,
    pub state_len: i32,
    pub nodes: i64,
}

#[repr(C)]
#[derive(Default, Clone, Copy)]
pub struct hb_surface_info {
    pub n_nodes: i32,
    pub n_layers: i32,
    pub tstep_subdivision: i32,
    pub n_massive_chunks: i32,
    pub n_nomass_chunks: i32,
    pub state_offset: i64,
    pub front_emissivity: c_double,
    pub back_emissivity: c_double,
    pub wind_speed_modifier: c_double,
    pub cos_tilt: c_double,
}

#[repr(C)]
pub struct hb_series {
    pub n_steps: i32,
    pub weather_mode: i32,
    pub n_sites: i32,
    pub model_site: *const i32,
    pub weather: *const c_double,
    pub weather_rows: i64,
    pub input_mode: i32,
    pub ir_mode: i32,
    pub inputs: *const c_double,
    pub record_every: i32,
    pub zone_t_record: *mut c_double,
    pub state_record: *mut c_double,
    /// [n_steps][zones] HVAC + luminaire power schedule (host), or null (model.rs:502-516)
    pub zone_gain: *const c_double,
    /// [n_steps][zones] infiltration / ventilation schedules (host), or null (model.rs:524-544)
    pub zone_infiltration_volume: *const c_double,
    pub zone_infiltration_temperature: *const c_double,
    pub zone_ventilation_volume: *const c_double,
    pub zone_ventilation_temperature: *const c_double,
}

#[repr(C)]
#[derive(Default, Clone, Copy)]
pub struct hb_ensemble_stats {
    pub sum_zone_t: c_double,
    pub min_zone_t: c_double,
    pub max_zone_t: c_double,
    pub sum_node_t: c_double,
    pub min_node_t: c_double,
    pub max_node_t: c_double,
    pub front_conv_energy_j: c_double,
    pub back_conv_energy_j: c_double,
    pub n_zones: i64,
    pub n_nodes: i64,
    pub nomass_iterations: i64,
    pub status_flags: i64,
}

#[repr(C)]
pub struct hb_handle {
    _private: [u8; 0],
}

extern "C" {
    pub fn hb_last_error() -> *const c_char;
    pub fn hb_version() -> *const c_char;
    pub fn hb_create(desc: *const hb_batch_desc, device: i32, out: *mut *mut hb_handle) -> hb_status;
    pub fn hb_destroy(h: *mut hb_handle);
    pub fn hb_n_models(h: *const hb_handle) -> i32;
    pub fn hb_n_surfaces_total(h: *const hb_handle) -> i64;
    pub fn hb_n_zones_total(h: *const hb_handle) -> i64;
    pub fn hb_state_len_total(h: *const hb_handle) -> i64;
    pub fn hb_nodes_total(h: *const hb_handle) -> i64;
    pub fn hb_updates_per_step(h: *const hb_handle) -> i64;
    pub fn hb_get_model_info(h: *const hb_handle, model: i32, out: *mut hb_model_info) -> hb_status;
    pub fn hb_get_surface_info(h: *const hb_handle, surface: i64, out: *mut hb_surface_info) -> hb_status;
    pub fn hb_get_surface_arrays(h: *const hb_handle, surface: i64, n_elements: *mut i32, mass: *mut c_double,
                                 u: *mut c_double, front_alpha: *mut c_double, back_alpha: *mut c_double,
                                 massive_chunks: *mut i32, nomass_chunks: *mut i32) -> hb_status;
    pub fn hb_surface_rk4_number(h: *const hb_handle, surface: i64, film: c_double, out: *mut c_double) -> hb_status;
    pub fn hb_set_state(h: *mut hb_handle, state: *const c_double) -> hb_status;
    pub fn hb_get_state(h: *mut hb_handle, state: *mut c_double) -> hb_status;
    pub fn hb_set_zone_inputs(h: *mut hb_handle, q_internal: *const c_double, infiltration_volume: *const c_double,
                              infiltration_temperature: *const c_double, ventilation_volume: *const c_double,
                              ventilation_temperature: *const c_double) -> hb_status;
    pub fn hb_march_series(h: *mut hb_handle, series: *const hb_series) -> hb_status;
    pub fn hb_upload(h: *mut hb_handle, host: *const c_double, count: i64, dev_out: *mut *mut c_double) -> hb_status;
    pub fn hb_free_dev(h: *mut hb_handle, dev: *mut c_double) -> hb_status;
    pub fn hb_march_series_dev(h: *mut hb_handle, series: *const hb_series, elapsed_ms: *mut c_float) -> hb_status;
    pub fn hb_launch_count(h: *const hb_handle) -> i64;
    pub fn hb_reduce(h: *mut hb_handle, out: *mut hb_ensemble_stats) -> hb_status;
    pub fn hb_reduce_noalloc(h: *mut hb_handle, out_host: *mut hb_ensemble_stats, out_dev: *mut hb_ensemble_stats) -> hb_status;
    pub fn hb_reset_energy(h: *mut hb_handle) -> hb_status;
    // the path's only collective (NCCL inside the library)
    pub fn hb_comm_unique_id(id_out: *mut c_void) -> hb_status;
    pub fn hb_comm_init(h: *mut hb_handle, id: *const c_void, rank: i32, world: i32) -> hb_status;
    pub fn hb_comm_destroy(h: *mut hb_handle) -> hb_status;
    pub fn hb_reduce_all(h: *mut hb_handle, out: *mut hb_ensemble_stats) -> hb_status;
    // EPW records in hb_series layout (host only)
    pub fn hb_read_epw(path: *const c_char, weather: *mut c_double, inputs: *mut c_double, capacity: i64,
                       n_rows: *mut i64) -> hb_status;
    // device probes used by the benchmark
    pub fn hb_measure_fp64_peak(device: i32, tflops: *mut c_double) -> hb_status;
    pub fn hb_flush_l2(h: *mut hb_handle, bytes: i64) -> hb_status;
}

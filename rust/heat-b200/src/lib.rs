// UNCOMPILED IN THIS ENVIRONMENT (no cargo / rustc in the image; the reference's git dependencies are not vendored).
// heat-b200/src/lib.rs — safe wrapper over heat-b200-sys.
//
//   ThermalBatch      the batched entry point the north star adds: many independent models in one handle, marched over
//                     whole series in one persistent launch per geometry (executable twin: heat_b200/model.py).
//   ThermalModelGpu   (feature "simple") drop-in for heat::model::ThermalModel behind the reference's own trait
//                     communication_protocols::SimulationModel (reference src/model.rs:188-428).
use heat_b200_sys as sys;
use std::ffi::CStr;
use std::ptr;

fn check(rc: sys::hb_status) -> Result<(), String> {
    if rc == sys::HB_OK {
        return Ok(());
    }
    Err(unsafe { CStr::from_ptr(sys::hb_last_error()) }.to_string_lossy().into_owned())
}

/// Flat description of a batch of models (mirrors `hb_batch_desc`); the vectors own what the raw struct borrows.
pub struct BatchDesc {
    pub layers: Vec<sys::hb_layer>,
    pub construction_offset: Vec<i32>,
    pub model_zone_offset: Vec<i32>,
    pub model_surface_offset: Vec<i32>,
    pub zone_volume: Vec<f64>,
    pub surfaces: Vec<sys::hb_surface>,
    pub model_terrain: Option<Vec<i32>>,
    pub steps_per_hour: i32,
}

impl BatchDesc {
    fn raw(&self) -> sys::hb_batch_desc {
        sys::hb_batch_desc {
            n_layers: self.layers.len() as i32,
            layers: self.layers.as_ptr(),
            n_constructions: self.construction_offset.len() as i32 - 1,
            construction_offset: self.construction_offset.as_ptr(),
            n_models: self.model_zone_offset.len() as i32 - 1,
            model_zone_offset: self.model_zone_offset.as_ptr(),
            model_surface_offset: self.model_surface_offset.as_ptr(),
            zone_volume: self.zone_volume.as_ptr(),
            surfaces: self.surfaces.as_ptr(),
            model_terrain: self.model_terrain.as_ref().map_or(ptr::null(), |v| v.as_ptr()),
            steps_per_hour: self.steps_per_hour,
        }
    }
}

/// One `hb_handle`: a batch of independent models resident on one GPU.  Not re-entrant (`&mut self`).
pub struct ThermalBatch {
    h: *mut sys::hb_handle,
}
unsafe impl Send for ThermalBatch {}

/// Weather and irradiance series of one `march_series` call (host memory).
pub struct Series<'a> {
    pub n_steps: i32,
    pub weather_mode: i32,
    pub weather: &'a [f64], // [site][row][3]
    pub n_sites: i32,
    pub input_mode: i32,
    pub ir_mode: i32,
    pub inputs: Option<&'a [f64]>,
    pub model_site: Option<&'a [i32]>,
    pub zone_gain: Option<&'a [f64]>,
    pub zone_infiltration_volume: Option<&'a [f64]>,
    pub zone_infiltration_temperature: Option<&'a [f64]>,
    pub zone_ventilation_volume: Option<&'a [f64]>,
    pub zone_ventilation_temperature: Option<&'a [f64]>,
}

impl ThermalBatch {
    pub fn new(desc: &BatchDesc, device: i32) -> Result<Self, String> {
        let raw = desc.raw();
        let mut h = ptr::null_mut();
        check(unsafe { sys::hb_create(&raw, device, &mut h) })?;
        Ok(Self { h })
    }
    pub fn state_len(&self) -> usize {
        unsafe { sys::hb_state_len_total(self.h) as usize }
    }
    pub fn updates_per_step(&self) -> i64 {
        unsafe { sys::hb_updates_per_step(self.h) }
    }
    pub fn model_info(&self, model: i32) -> Result<sys::hb_model_info, String> {
        let mut out = sys::hb_model_info::default();
        check(unsafe { sys::hb_get_model_info(self.h, model, &mut out) })?;
        Ok(out)
    }
    pub fn surface_info(&self, surface: i64) -> Result<sys::hb_surface_info, String> {
        let mut out = sys::hb_surface_info::default();
        check(unsafe { sys::hb_get_surface_info(self.h, surface, &mut out) })?;
        Ok(out)
    }
    /// dt·λmax of the wall's explicit march (bounded below 2.7853); see `hb_surface_rk4_number`.
    pub fn surface_rk4_number(&self, surface: i64, film: f64) -> Result<f64, String> {
        let mut out = 0.0f64;
        check(unsafe { sys::hb_surface_rk4_number(self.h, surface, film, &mut out) })?;
        Ok(out)
    }
    pub fn get_state(&mut self, state: &mut [f64]) -> Result<(), String> {
        assert_eq!(state.len(), self.state_len());
        check(unsafe { sys::hb_get_state(self.h, state.as_mut_ptr()) })
    }
    pub fn set_state(&mut self, state: &[f64]) -> Result<(), String> {
        assert_eq!(state.len(), self.state_len());
        check(unsafe { sys::hb_set_state(self.h, state.as_ptr()) })
    }
    pub fn set_zone_inputs(&mut self, q: Option<&[f64]>, inf_v: Option<&[f64]>, inf_t: Option<&[f64]>,
                           vent_v: Option<&[f64]>, vent_t: Option<&[f64]>) -> Result<(), String> {
        let p = |o: Option<&[f64]>| o.map_or(ptr::null(), |s| s.as_ptr());
        check(unsafe { sys::hb_set_zone_inputs(self.h, p(q), p(inf_v), p(inf_t), p(vent_v), p(vent_t)) })
    }
    /// Advance every model `n_steps` main steps (model.rs:359-427 for each of them, `dt_subdivisions` sub-steps each).
    pub fn march_series(&mut self, s: &Series) -> Result<(), String> {
        let p = |o: Option<&[f64]>| o.map_or(ptr::null(), |v| v.as_ptr());
        let raw = sys::hb_series {
            n_steps: s.n_steps,
            weather_mode: s.weather_mode,
            n_sites: s.n_sites,
            model_site: s.model_site.map_or(ptr::null(), |v| v.as_ptr()),
            weather: s.weather.as_ptr(),
            weather_rows: (s.weather.len() / (3 * s.n_sites as usize)) as i64,
            input_mode: s.input_mode,
            ir_mode: s.ir_mode,
            inputs: p(s.inputs),
            record_every: 0,
            zone_t_record: ptr::null_mut(),
            state_record: ptr::null_mut(),
            zone_gain: p(s.zone_gain),
            zone_infiltration_volume: p(s.zone_infiltration_volume),
            zone_infiltration_temperature: p(s.zone_infiltration_temperature),
            zone_ventilation_volume: p(s.zone_ventilation_volume),
            zone_ventilation_temperature: p(s.zone_ventilation_temperature),
        };
        check(unsafe { sys::hb_march_series(self.h, &raw) })
    }
    /// Statistics of this GPU's shard.
    pub fn reduce(&mut self) -> Result<sys::hb_ensemble_stats, String> {
        let mut out = sys::hb_ensemble_stats::default();
        check(unsafe { sys::hb_reduce(self.h, &mut out) })?;
        Ok(out)
    }
    /// NCCL id made on rank 0; the host ships the 128 bytes to the other ranks (file, socket, MPI ...).
    pub fn comm_unique_id() -> Result<[u8; sys::HB_COMM_ID_BYTES], String> {
        let mut id = [0u8; sys::HB_COMM_ID_BYTES];
        check(unsafe { sys::hb_comm_unique_id(id.as_mut_ptr() as *mut _) })?;
        Ok(id)
    }
    pub fn comm_init(&mut self, id: &[u8; sys::HB_COMM_ID_BYTES], rank: i32, world: i32) -> Result<(), String> {
        check(unsafe { sys::hb_comm_init(self.h, id.as_ptr() as *const _, rank, world) })
    }
    /// Statistics of the whole ensemble over all GPUs of the job (NCCL all-reduce inside the library).
    pub fn reduce_all(&mut self) -> Result<sys::hb_ensemble_stats, String> {
        let mut out = sys::hb_ensemble_stats::default();
        check(unsafe { sys::hb_reduce_all(self.h, &mut out) })?;
        Ok(out)
    }
}

impl Drop for ThermalBatch {
    fn drop(&mut self) {
        unsafe { sys::hb_destroy(self.h) }
    }
}

#[cfg(feature = "simple")]
pub mod simple;

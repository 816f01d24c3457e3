// UNCOMPILED IN THIS ENVIRONMENT (no cargo / rustc; simple_model, communication_protocols, weather, calendar are
// un-vendored git dependencies of the reference, Cargo.toml:10-20).
// Drop-in for heat::model::ThermalModel behind the same SimulationModel trait (reference src/model.rs:188-428).
// heat_b200/model.py:ThermalModel is the executable twin (same new / allocate_memory / march, same weather sampling)
// exercised by tests/test_parity_gpu.py::test_drop_in_march_matches_series_and_oracle.
use crate::{check, sys};
use communication_protocols::{ErrorHandling, MetaOptions, SimulationModel};
use simple_model::{SimpleModel, SimulationState, SimulationStateHeader};
use std::{borrow::Borrow, ptr};
use weather::Weather;

pub struct ThermalModelGpu {
    h: *mut sys::hb_handle,
    info: sys::hb_model_info,
    first_slot: usize,
}
pub struct GpuMemory; // scratch lives on the device (cf. ThermalModelMemory, model.rs:46-49)
unsafe impl Send for ThermalModelGpu {}

impl ErrorHandling for ThermalModelGpu {
    fn module_name() -> &'static str {
        "Thermal model" // model.rs:182-186
    }
}

impl SimulationModel for ThermalModelGpu {
    type OutputType = Self;
    type OptionType = ();
    type AllocType = GpuMemory; // model.rs:189-191

    fn allocate_memory(&self) -> Result<GpuMemory, String> {
        Ok(GpuMemory) // model.rs:193
    }

    fn new<M: Borrow<SimpleModel>>(_meta: &MetaOptions, _o: (), model: M, state: &mut SimulationStateHeader,
                                   n: usize) -> Result<Self, String> {
        let model = model.borrow();
        // 1. flatten what ThermalModel::new reads (model.rs:224-323): substances + materials -> hb_layer,
        //    constructions -> offsets, spaces -> zone volumes, surfaces then fenestrations -> hb_surface
        //    {area, perimeter, centroid.z, normal, boundaries}
        let desc = flatten::batch_of_one(model, n)?;
        let raw = desc.raw();
        let mut h = ptr::null_mut();
        check(unsafe { sys::hb_create(&raw, 0, &mut h) })?;
        // 2. register the same state elements in the same order as the reference (zone.rs:43-50,
        //    surface.rs:428-442) with the values of hb_get_state right after create
        let first_slot = state.len();
        flatten::register_state_elements(model, state, h)?;
        let mut info = sys::hb_model_info::default();
        check(unsafe { sys::hb_get_model_info(h, 0, &mut info) })?;
        Ok(Self { h, info, first_slot })
    }

    fn march<W: Weather, M: Borrow<SimpleModel>>(&self, mut date: calendar::Date, weather: &W, _model: M,
                                                 state: &mut SimulationState, _a: &mut GpuMemory)
                                                 -> Result<(), String> {
        // weather sampled at the END of each sub-step, as model.rs:369-382 does
        let s = self.info.dt_subdivisions as usize;
        let mut rows = Vec::with_capacity(3 * s);
        for _ in 0..s {
            date.add_seconds(self.info.dt);
            let w = weather.get_weather_data(date);
            rows.extend([w.dry_bulb_temperature, w.wind_direction.ok_or("wind direction")?, w.wind_speed]);
        }
        let slice = &mut state.as_mut_slice()[self.first_slot..self.first_slot + self.info.state_len as usize];
        check(unsafe { sys::hb_set_state(self.h, slice.as_ptr()) })?;
        let series = sys::hb_series {
            n_steps: 1,
            weather_mode: sys::HB_WEATHER_PER_SUBSTEP,
            n_sites: 1,
            model_site: ptr::null(),
            weather: rows.as_ptr(),
            weather_rows: s as i64,
            input_mode: sys::HB_INPUT_STATE, // irradiances are already in the state
            ir_mode: sys::HB_IR_IRRADIANCE,
            inputs: ptr::null(),
            record_every: 0,
            zone_t_record: ptr::null_mut(),
            state_record: ptr::null_mut(),
            zone_gain: ptr::null(), // one step: gains / infiltration / ventilation were passed with hb_set_zone_inputs
            zone_infiltration_volume: ptr::null(),
            zone_infiltration_temperature: ptr::null(),
            zone_ventilation_volume: ptr::null(),
            zone_ventilation_temperature: ptr::null(),
        };
        check(unsafe { sys::hb_march_series(self.h, &series) })?;
        check(unsafe { sys::hb_get_state(self.h, slice.as_mut_ptr()) })
    }
}

impl Drop for ThermalModelGpu {
    fn drop(&mut self) {
        unsafe { sys::hb_destroy(self.h) }
    }
}

/// SimpleModel -> flat descriptor.  What each field is read from is listed in INTEGRATION.md §3; the bodies depend on
/// the un-vendored `simple_model` API and are left to the maintainer who has it.
mod flatten {
    use crate::{sys, BatchDesc};
    use simple_model::{SimpleModel, SimulationStateHeader};
    pub fn batch_of_one(_model: &SimpleModel, _n: usize) -> Result<BatchDesc, String> {
        unimplemented!("needs the simple_model crate: see INTEGRATION.md §3 for the field-by-field mapping")
    }
    pub fn register_state_elements(_model: &SimpleModel, _state: &mut SimulationStateHeader,
                                   _h: *mut sys::hb_handle) -> Result<(), String> {
        unimplemented!("push SpaceDryBulbTemperature / Surface* / Fenestration* elements in the reference's order")
    }
}

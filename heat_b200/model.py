"""Host mirror of the reference interface over the C ABI.

* `ThermalModel` mirrors `heat::model::ThermalModel` as a `SimulationModel` (reference src/model.rs:188-428):
  `new(meta_options, options, model, state_header, n)`, `allocate_memory()`, `march(date, weather, model, state,
  alloc)`, `dt_subdivisions`, `dt`.  The caller owns the flat `SimulationState`; `march` advances it in place.
* `ThermalBatch` is the batched entry point: many independent models in one handle, marched over whole series
  in one persistent kernel launch.

Everything numerical happens in libheat_b200.so (CUDA, sm_100a).  There is no CPU path here.
"""
from __future__ import annotations

import ctypes as C
import dataclasses
import math
import os
from typing import List, Optional, Sequence

import numpy as np

from . import _lib
from .desc import BatchDesc

MODULE_NAME = "Thermal model"  # src/model.rs:42


# --------------------------------------------------------------------------------------------------------------
# Small stand-ins for the un-vendored crates the reference's API names (calendar, weather, simple_model state).
# --------------------------------------------------------------------------------------------------------------
_DAYS = (31, 28, 31, 30, 31, 30, 31, 31, 30, 31, 30, 31)


@dataclasses.dataclass
class Date:
    """calendar::Date {month, day, hour} (1-based month/day, fractional hour, non-leap year)."""

    month: int
    day: int
    hour: float

    def add_seconds(self, seconds: float) -> None:
        self.add_hours(seconds / 3600.0)

    def add_hours(self, hours: float) -> None:
        self.hour += hours
        while self.hour >= 24.0:
            self.hour -= 24.0
            self.day += 1
            if self.day > _DAYS[self.month - 1]:
                self.day = 1
                self.month = self.month % 12 + 1

    def copy(self) -> "Date":
        return Date(self.month, self.day, self.hour)

    def hour_of_year(self) -> float:
        return 24.0 * (sum(_DAYS[: self.month - 1]) + self.day - 1) + self.hour


@dataclasses.dataclass
class CurrentWeather:
    """weather::CurrentWeather — the three fields `march` reads (src/model.rs:372-382)."""

    dry_bulb_temperature: Optional[float]
    wind_speed: Optional[float]
    wind_direction: Optional[float]  # degrees


class ConstantWeather:
    """weather::SyntheticWeather with ScheduleConstant members (validate_wall_heat_transfer.rs:144-147)."""

    def __init__(self, dry_bulb_temperature: float, wind_direction: float = 0.0, wind_speed: float = 0.0):
        self._w = CurrentWeather(dry_bulb_temperature, wind_speed, wind_direction)

    def get_weather_data(self, date: Date) -> CurrentWeather:
        return self._w


class SimulationStateHeader:
    """simple_model::SimulationStateHeader: ordered (element, initial value) pairs."""

    def __init__(self):
        self.elements: List[str] = []
        self.values: List[float] = []

    def push(self, element: str, value: float) -> int:
        self.elements.append(element)
        self.values.append(float(value))
        return len(self.values) - 1

    def __len__(self) -> int:
        return len(self.values)

    def take_values(self) -> np.ndarray:
        return np.array(self.values, dtype=np.float64)


class ThermalModelMemory:
    """`AllocType` of the reference (src/model.rs:46-49). Scratch lives on the device; this is a token."""

    def __init__(self, owner: "ThermalModel"):
        self.owner = owner


# --------------------------------------------------------------------------------------------------------------
# Batched entry point
# --------------------------------------------------------------------------------------------------------------
class ThermalBatch:
    """One `hb_handle`: a batch of independent models resident on one GPU."""

    def __init__(self, desc: BatchDesc, device: int = 0):
        self._lib = _lib.load()
        self.desc = desc
        d = _lib.BatchDescC()
        d.n_layers = len(desc.layers)
        d.layers = _lib.ptr(desc.layers)
        d.n_constructions = desc.n_constructions
        d.construction_offset = _lib.ptr(desc.construction_offset)
        d.n_models = desc.n_models
        d.model_zone_offset = _lib.ptr(desc.model_zone_offset)
        d.model_surface_offset = _lib.ptr(desc.model_surface_offset)
        d.zone_volume = _lib.ptr(desc.zone_volume)
        d.surfaces = _lib.ptr(desc.surfaces)
        d.model_terrain = _lib.ptr(desc.model_terrain)
        d.steps_per_hour = int(desc.steps_per_hour)
        h = C.c_void_p()
        _lib.check(self._lib.hb_create(C.byref(d), int(device), C.byref(h)))
        self._h = h
        self.device = device
        self.n_models = self._lib.hb_n_models(h)
        self.n_surfaces = self._lib.hb_n_surfaces_total(h)
        self.n_zones = self._lib.hb_n_zones_total(h)
        self.state_len = self._lib.hb_state_len_total(h)
        self.nodes = self._lib.hb_nodes_total(h)
        self.updates_per_step = self._lib.hb_updates_per_step(h)

    def close(self) -> None:
        if getattr(self, "_h", None):
            self._lib.hb_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- layout ----
    def model_info(self, m: int) -> _lib.ModelInfoC:
        out = _lib.ModelInfoC()
        _lib.check(self._lib.hb_get_model_info(self._h, m, C.byref(out)))
        return out

    def surface_info(self, s: int) -> _lib.SurfaceInfoC:
        out = _lib.SurfaceInfoC()
        _lib.check(self._lib.hb_get_surface_info(self._h, s, C.byref(out)))
        return out

    def surface_arrays(self, s: int) -> dict:
        info = self.surface_info(s)
        out = {
            "n_elements": np.zeros(info.n_layers, np.int32),
            "mass": np.zeros(info.n_nodes), "u": np.zeros(info.n_nodes),
            "front_alpha": np.zeros(info.n_nodes), "back_alpha": np.zeros(info.n_nodes),
            "massive_chunks": np.zeros((info.n_massive_chunks, 2), np.int32),
            "nomass_chunks": np.zeros((info.n_nomass_chunks, 2), np.int32),
        }
        _lib.check(self._lib.hb_get_surface_arrays(
            self._h, s, _lib.ptr(out["n_elements"]), _lib.ptr(out["mass"]), _lib.ptr(out["u"]),
            _lib.ptr(out["front_alpha"]), _lib.ptr(out["back_alpha"]), _lib.ptr(out["massive_chunks"]),
            _lib.ptr(out["nomass_chunks"])))
        out["tstep_subdivision"] = info.tstep_subdivision
        out["n_nodes"] = info.n_nodes
        return out

    def surface_rk4_number(self, s: int, film: float = 30.0) -> float:
        """dt·λmax of the wall's explicit march (bounded below 2.7853), hb_surface_rk4_number."""
        out = C.c_double(0.0)
        _lib.check(self._lib.hb_surface_rk4_number(self._h, s, float(film), C.byref(out)))
        return out.value

    # ---- state ----
    def get_state(self, out: Optional[np.ndarray] = None) -> np.ndarray:
        if out is None:
            out = np.empty(self.state_len, np.float64)
        assert out.dtype == np.float64 and out.size == self.state_len and out.flags["C_CONTIGUOUS"]
        _lib.check(self._lib.hb_get_state(self._h, _lib.ptr(out)))
        return out

    def set_state(self, state: np.ndarray) -> None:
        state = _lib.f8(state)
        assert state.size == self.state_len
        _lib.check(self._lib.hb_set_state(self._h, _lib.ptr(state)))

    def set_zone_inputs(self, q_internal=None, infiltration_volume=None, infiltration_temperature=None,
                        ventilation_volume=None, ventilation_temperature=None) -> None:
        arrs = [None if a is None else _lib.f8(a) for a in
                (q_internal, infiltration_volume, infiltration_temperature, ventilation_volume, ventilation_temperature)]
        for a in arrs:
            assert a is None or a.size == self.n_zones
        _lib.check(self._lib.hb_set_zone_inputs(self._h, *[_lib.ptr(a) for a in arrs]))

    # ---- marching ----
    def _series(self, n_steps, weather, weather_mode, inputs, input_mode, ir_mode, model_site, keep):
        s = _lib.SeriesC()
        weather = _lib.f8(weather)
        if weather.ndim == 2:
            weather = weather[None]
        assert weather.ndim == 3 and weather.shape[2] == 3, "weather must be [site][row][3]"
        s.n_steps = int(n_steps)
        s.weather_mode = int(weather_mode)
        s.n_sites = weather.shape[0]
        s.weather_rows = weather.shape[1]
        s.weather = _lib.ptr(weather)
        keep.append(weather)
        if model_site is not None:
            model_site = np.ascontiguousarray(model_site, np.int32)
            assert model_site.size == self.n_models
            s.model_site = _lib.ptr(model_site)
            keep.append(model_site)
        s.input_mode = int(input_mode)
        s.ir_mode = int(ir_mode)
        if inputs is not None:
            inputs = _lib.f8(inputs)
            want = (s.n_sites * n_steps * 4) if input_mode == _lib.INPUT_SITE else n_steps * self.n_surfaces * 4
            assert inputs.size == want, f"inputs has {inputs.size} values, expected {want}"
            s.inputs = _lib.ptr(inputs)
            keep.append(inputs)
        return s

    def march_series(self, n_steps: int, weather, *, weather_mode=_lib.WEATHER_STEP_CONSTANT, inputs=None,
                     input_mode=_lib.INPUT_STATE, ir_mode=_lib.IR_IRRADIANCE, model_site=None, record_every: int = 0,
                     record_state: bool = False, zone_gain=None, zone_infiltration_volume=None,
                     zone_infiltration_temperature=None, zone_ventilation_volume=None,
                     zone_ventilation_temperature=None):
        """Advance every model `n_steps` main steps. Returns (zone_T[rec][zones], state[rec][state_len]) or None.

        `zone_gain` [n_steps][zones]: schedule of the zones' HVAC + luminaire power [W] (model.rs:502-516); row t
        replaces `q_internal` of `set_zone_inputs` during main step t.  `zone_infiltration_*` / `zone_ventilation_*`
        [n_steps][zones]: schedules of the infiltration / ventilation volume flow [m3/s] and temperature [C]
        (model.rs:524-544; NaN temperature = absent in that step)."""
        keep: list = []
        s = self._series(n_steps, weather, weather_mode, inputs, input_mode, ir_mode, model_site, keep)
        for name, table in zip(_lib.ZONE_SCHEDULES, (zone_gain, zone_infiltration_volume, zone_infiltration_temperature,
                                                     zone_ventilation_volume, zone_ventilation_temperature)):
            if table is not None:
                table = _lib.f8(table)
                assert table.size == n_steps * self.n_zones, f"{name} has {table.size} values"
                setattr(s, name, _lib.ptr(table))
                keep.append(table)
        zone_rec = state_rec = None
        if record_every > 0:
            n_rec = n_steps // record_every
            s.record_every = record_every
            zone_rec = np.zeros((n_rec, self.n_zones))
            s.zone_t_record = _lib.ptr(zone_rec)
            if record_state:
                state_rec = np.zeros((n_rec, self.state_len))
                s.state_record = _lib.ptr(state_rec)
        _lib.check(self._lib.hb_march_series(self._h, C.byref(s)))
        return zone_rec, state_rec

    def upload(self, host) -> int:
        host = _lib.f8(host)
        dev = C.c_void_p()
        _lib.check(self._lib.hb_upload(self._h, _lib.ptr(host), host.size, C.byref(dev)))
        return dev.value

    def free_dev(self, dev: int) -> None:
        _lib.check(self._lib.hb_free_dev(self._h, C.c_void_p(dev)))

    def march_series_dev(self, n_steps: int, weather_dev: int, weather_rows: int, n_sites: int = 1, *,
                         weather_mode=_lib.WEATHER_STEP_CONSTANT, inputs_dev: Optional[int] = None,
                         input_mode=_lib.INPUT_STATE, ir_mode=_lib.IR_IRRADIANCE, model_site=None) -> float:
        """Inputs already resident on the device. Returns the CUDA-event time of the launch(es) in ms."""
        s = _lib.SeriesC()
        s.n_steps, s.weather_mode, s.n_sites = int(n_steps), int(weather_mode), int(n_sites)
        s.weather, s.weather_rows = weather_dev, int(weather_rows)
        s.input_mode, s.ir_mode = int(input_mode), int(ir_mode)
        s.inputs = inputs_dev
        keep = None
        if model_site is not None:
            keep = np.ascontiguousarray(model_site, np.int32)
            s.model_site = _lib.ptr(keep)
        ms = C.c_float()
        _lib.check(self._lib.hb_march_series_dev(self._h, C.byref(s), C.byref(ms)))
        return float(ms.value)

    def reduce(self) -> _lib.EnsembleStatsC:
        out = _lib.EnsembleStatsC()
        _lib.check(self._lib.hb_reduce(self._h, C.byref(out)))
        return out

    def reduce_noalloc(self, out_dev: int = 0, host: bool = True) -> Optional[_lib.EnsembleStatsC]:
        """hb_reduce_noalloc: the same statistics without any allocation (static device scratch, at most
        HB_REDUCE_NOALLOC_MAX_TILES tiles).  `out_dev`: device address of 96 bytes the kernel also writes (no
        synchronize); `host=False`: device output only, returns None."""
        out = _lib.EnsembleStatsC() if host else None
        _lib.check(self._lib.hb_reduce_noalloc(self._h, C.byref(out) if host else None, C.c_void_p(out_dev or None)))
        return out

    # ---- the ensemble statistics over all GPUs of a job: NCCL inside the library ----
    @staticmethod
    def comm_unique_id() -> bytes:
        """ncclGetUniqueId (rank 0); ship the 128 bytes to the other ranks by any means."""
        buf = C.create_string_buffer(128)
        _lib.check(_lib.load().hb_comm_unique_id(buf))
        return buf.raw

    def comm_init(self, unique_id: bytes, rank: int, world: int) -> None:
        assert len(unique_id) == 128
        _lib.check(self._lib.hb_comm_init(self._h, C.c_char_p(unique_id), int(rank), int(world)))

    def reduce_all(self) -> _lib.EnsembleStatsC:
        """hb_reduce + ncclAllReduce (sum / min / max) — every rank gets the statistics of the whole ensemble."""
        out = _lib.EnsembleStatsC()
        _lib.check(self._lib.hb_reduce_all(self._h, C.byref(out)))
        return out

    def reset_energy(self) -> None:
        _lib.check(self._lib.hb_reset_energy(self._h))

    def flush_l2(self, nbytes: int = 256 << 20) -> None:
        _lib.check(self._lib.hb_flush_l2(self._h, nbytes))

    @property
    def launch_count(self) -> int:
        return self._lib.hb_launch_count(self._h)


def read_epw(path: str, periodic: bool = True):
    """Hourly records of an EPW file → (weather[rows(+1)][3], inputs[rows][4]) as `march_series` takes them
    (`weather_mode=WEATHER_LINEAR, input_mode=INPUT_SITE`); `periodic` appends the first record after the last."""
    lib = _lib.load()
    n = C.c_int64()
    _lib.check(lib.hb_read_epw(os.fsencode(path), None, None, 0, C.byref(n)))
    weather = np.zeros((n.value + (1 if periodic else 0), 3))
    inputs = np.zeros((n.value, 4))
    _lib.check(lib.hb_read_epw(os.fsencode(path), _lib.ptr(weather), _lib.ptr(inputs), n.value, C.byref(n)))
    if periodic and n.value:
        weather[-1] = weather[0]
    return weather, inputs


def measure_fp64_peak(device: int = 0) -> float:
    out = C.c_double()
    _lib.check(_lib.load().hb_measure_fp64_peak(device, C.byref(out)))
    return float(out.value)


# --------------------------------------------------------------------------------------------------------------
# Drop-in mirror of heat::model::ThermalModel
# --------------------------------------------------------------------------------------------------------------
class ThermalModel:
    """`impl SimulationModel for ThermalModel` (reference src/model.rs:188-428) on the GPU."""

    def __init__(self):
        raise TypeError("use ThermalModel.new(...)")

    @staticmethod
    def module_name() -> str:  # ErrorHandling::module_name, src/model.rs:182-186
        return MODULE_NAME

    @classmethod
    def new(cls, meta_options, options, model: BatchDesc, state: SimulationStateHeader, n: int,
            device: int = 0) -> "ThermalModel":
        """src/model.rs:215-354. `model` is a single-model BatchDesc; `state` receives the thermal state slots."""
        if model.n_models != 1:
            raise ValueError("ThermalModel.new takes exactly one model (use ThermalBatch for ensembles)")
        if int(n) != int(model.steps_per_hour):
            model = dataclasses.replace(model, steps_per_hour=int(n))
        self = object.__new__(cls)
        self.batch = ThermalBatch(model, device)
        info = self.batch.model_info(0)
        self.dt_subdivisions = info.dt_subdivisions  # pub field, src/model.rs:73
        self.dt = info.dt                            # pub field, src/model.rs:76
        self.n_zones = info.n_zones
        self.n_surfaces = info.n_surfaces
        self.state_base = len(state)
        initial = self.batch.get_state()
        # registration order: zones (zone.rs:43-50), then per surface 8 scalars + nodes (surface.rs:428-442)
        names = [f"SpaceDryBulbTemperature({z})" for z in range(info.n_zones)]
        for s in range(info.n_surfaces):
            si = self.batch.surface_info(s)
            kind = "Fenestration" if model.surfaces[s]["is_fenestration"] else "Surface"
            names += [f"{kind}{e}({s})" for e in (
                "FrontConvectionCoefficient", "BackConvectionCoefficient", "FrontConvectiveHeatFlow",
                "BackConvectiveHeatFlow", "FrontSolarIrradiance", "BackSolarIrradiance", "FrontIRIrradiance",
                "BackIRIrradiance")]
            names += [f"{kind}NodeTemperature({s},{i})" for i in range(si.n_nodes)]
        assert len(names) == initial.size
        for name, v in zip(names, initial):
            state.push(name, v)
        self.state_len = initial.size
        self._zone_inputs = None
        return self

    def allocate_memory(self) -> ThermalModelMemory:  # src/model.rs:193-207
        return ThermalModelMemory(self)

    # state slot helpers (what SurfaceTrait / Space accessors do in the reference)
    def zone_index(self, z: int) -> int:
        return self.state_base + z

    def surface_slot(self, s: int, slot: int) -> int:
        """slot: 0 front h, 1 back h, 2 front flow, 3 back flow, 4 front solar, 5 back solar, 6 front IR, 7 back IR."""
        return self.state_base + self.batch.surface_info(s).state_offset + slot

    def node_index(self, s: int, node: int) -> int:
        info = self.batch.surface_info(s)
        if node < 0:
            node += info.n_nodes
        return self.state_base + info.state_offset + 8 + node

    def set_zone_inputs(self, **kw) -> None:
        """HVAC/luminaire power, infiltration and ventilation (SimulationState elements in the reference)."""
        self.batch.set_zone_inputs(**kw)

    def march(self, date: Date, weather, model, state: np.ndarray, alloc: ThermalModelMemory) -> None:
        """src/model.rs:359-427: one main time step = dt_subdivisions sub-steps; `state` is advanced in place."""
        if alloc.owner is not self:
            raise ValueError("ThermalModelMemory belongs to another model")
        date = date.copy()  # `mut date: Date` is taken by value
        rows = np.empty((self.dt_subdivisions, 3))
        for s in range(self.dt_subdivisions):
            date.add_seconds(self.dt)
            w = weather.get_weather_data(date)
            if w.dry_bulb_temperature is None:
                raise ValueError("Trying to march on Thermal Model, but dry bulb temperature was not provided")
            rows[s] = (w.dry_bulb_temperature, w.wind_direction, w.wind_speed)
        view = state[self.state_base:self.state_base + self.state_len]
        self.batch.set_state(view)
        self.batch.march_series(1, rows, weather_mode=_lib.WEATHER_PER_SUBSTEP)
        view[:] = self.batch.get_state()


def radians(deg: float) -> float:
    return deg * (math.pi / 180.0)

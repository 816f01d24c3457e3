"""ctypes wrapper of the CPU ORACLE (oracle/heat_oracle.{h,cc}) — TEST INFRASTRUCTURE ONLY.

May be imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs, never by
the product package `heat_b200`.  It consumes the same flat descriptions (`heat_b200.desc.BatchDesc`) so that the
checker and the CUDA path are driven by identical inputs.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libheat_oracle.so")
_P = C.c_void_p


class ModelDescC(C.Structure):
    _fields_ = [
        ("n_layers", C.c_int32), ("layers", _P),
        ("n_constructions", C.c_int32), ("construction_offset", _P),
        ("n_zones", C.c_int32), ("zone_volume", _P),
        ("n_surfaces", C.c_int32), ("surfaces", _P),
        ("terrain", C.c_int32), ("steps_per_hour", C.c_int32),
    ]


class ZoneInputsC(C.Structure):
    _fields_ = [("q_internal", _P), ("infiltration_volume", _P), ("infiltration_temperature", _P),
                ("ventilation_volume", _P), ("ventilation_temperature", _P)]


_lib = None


def build() -> None:
    subprocess.run(["make", "-C", _HERE, "libheat_oracle.so"], check=True, capture_output=True)


def load() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        lib = C.CDLL(LIB_PATH)
        lib.ho_model_new.restype = _P
        lib.ho_model_new.argtypes = [C.POINTER(ModelDescC)]
        lib.ho_model_new_ex.restype = _P
        lib.ho_model_new_ex.argtypes = [C.POINTER(ModelDescC), C.c_double, C.c_double, C.c_double]
        lib.ho_model_free.argtypes = [_P]
        lib.ho_last_error.restype = C.c_char_p
        lib.ho_dt.restype = C.c_double
        lib.ho_dt.argtypes = [_P]
        for name in ("ho_dt_subdivisions", "ho_state_len"):
            getattr(lib, name).restype = C.c_int32
            getattr(lib, name).argtypes = [_P]
        for name in ("ho_surface_n_nodes", "ho_surface_state_offset", "ho_surface_tstep_subdivision"):
            getattr(lib, name).restype = C.c_int32
            getattr(lib, name).argtypes = [_P, C.c_int32]
        lib.ho_surface_n_elements.restype = C.c_int32
        lib.ho_surface_n_elements.argtypes = [_P, C.c_int32, _P, C.c_int32]
        lib.ho_surface_segments.restype = C.c_int32
        lib.ho_surface_segments.argtypes = [_P, C.c_int32, _P, _P]
        lib.ho_surface_chunks.restype = C.c_int32
        lib.ho_surface_chunks.argtypes = [_P, C.c_int32, C.c_int32, _P, C.c_int32]
        lib.ho_surface_alphas.restype = C.c_int32
        lib.ho_surface_alphas.argtypes = [_P, C.c_int32, _P, _P]
        lib.ho_surface_wind_modifier.restype = C.c_double
        lib.ho_surface_wind_modifier.argtypes = [_P, C.c_int32]
        lib.ho_surface_emissivity.restype = C.c_double
        lib.ho_surface_emissivity.argtypes = [_P, C.c_int32, C.c_int32]
        lib.ho_initial_state.argtypes = [_P, _P]
        lib.ho_march.restype = C.c_int32
        lib.ho_march.argtypes = [_P, _P, _P, C.POINTER(ZoneInputsC)]
        lib.ho_iterate_surfaces.restype = C.c_int32
        lib.ho_iterate_surfaces.argtypes = [_P, _P, C.c_double, C.c_double, C.c_double, C.c_double]
        lib.ho_zones_abc.restype = C.c_int32
        lib.ho_zones_abc.argtypes = [_P, _P, C.POINTER(ZoneInputsC), _P, _P, _P]
        lib.ho_nomass_iterations.restype = C.c_int64
        lib.ho_nomass_iterations.argtypes = [_P]
        lib.ho_nusselt.restype = C.c_double
        lib.ho_nusselt.argtypes = [C.c_double] * 3
        lib.ho_gas_property.restype = C.c_double
        lib.ho_gas_property.argtypes = [C.c_int32, C.c_int32, C.c_double]
        lib.ho_cavity_u_value.restype = C.c_double
        lib.ho_cavity_u_value.argtypes = [C.c_int32] + [C.c_double] * 7
        lib.ho_tarp_natural.restype = C.c_double
        lib.ho_tarp_natural.argtypes = [C.c_double] * 3
        lib.ho_tarp_forced_plus_natural.restype = C.c_double
        lib.ho_tarp_forced_plus_natural.argtypes = [C.c_double] * 6 + [C.c_int32]
        lib.ho_is_windward.restype = C.c_int32
        lib.ho_is_windward.argtypes = [C.c_double, C.c_double, _P]
        lib.ho_wind_speed_modifier.restype = C.c_double
        lib.ho_wind_speed_modifier.argtypes = [C.c_double, C.c_int32]
        lib.ho_rk4_steps.argtypes = [C.c_int32, _P, _P, _P, _P, C.c_double, C.c_int32, C.c_int32]
        lib.ho_discretize.restype = C.c_int32
        lib.ho_discretize.argtypes = [_P, C.c_int32, C.c_double, C.c_double, C.c_double, C.c_double, C.c_double, _P, _P,
                                      _P, _P, C.c_int32]
        lib.ho_build.restype = C.c_int32
        lib.ho_build.argtypes = [_P, C.c_int32, C.c_int32, _P, C.c_double, C.c_double, _P, _P, C.c_int32]
        lib.ho_chunks.restype = C.c_int32
        lib.ho_chunks.argtypes = [_P, C.c_int32, C.c_int32, _P, C.c_int32]
        lib.ho_get_k_q.restype = C.c_int32
        lib.ho_get_k_q.argtypes = [C.c_int32, _P, _P, C.c_int32, C.c_int32, _P] + [C.c_double] * 8 + [_P, _P]
        lib.ho_glazing_combine.restype = C.c_int32
        lib.ho_glazing_combine.argtypes = [_P, _P, _P, C.c_int32, _P]
        lib.ho_glazing_alphas.restype = C.c_int32
        lib.ho_glazing_alphas.argtypes = [_P, _P, _P, C.c_int32, _P]
        lib.ho_run_ensemble.restype = C.c_int64
        lib.ho_run_ensemble.argtypes = [_P, C.c_int32, C.c_int32, _P, _P, C.c_int32, C.c_int32, _P]
        lib.ho_run_ensemble_q.restype = C.c_int64
        lib.ho_run_ensemble_q.argtypes = [_P, C.c_int32, C.c_int32, _P, _P, C.c_int32, C.c_int32, _P, _P]
        _lib = lib
    return _lib


class OracleError(RuntimeError):
    pass


def _f8(a):
    return np.ascontiguousarray(a, np.float64)


class OracleModel:
    """One `ThermalModel` of the reference, restated on the CPU."""

    def __init__(self, desc, m: int = 0, discretization=None):
        """`desc`: heat_b200.desc.BatchDesc; `m`: which model of the batch; `discretization` = (main_dt, max_dx,
        min_dt) as the reference's unit tests pass to Discretization::new (default: model.rs:236-240)."""
        lib = load()
        self._lib = lib
        z0, z1 = int(desc.model_zone_offset[m]), int(desc.model_zone_offset[m + 1])
        s0, s1 = int(desc.model_surface_offset[m]), int(desc.model_surface_offset[m + 1])
        self._keep = (desc.layers, desc.construction_offset, np.ascontiguousarray(desc.zone_volume[z0:z1]),
                      np.ascontiguousarray(desc.surfaces[s0:s1]))
        d = ModelDescC()
        d.n_layers = len(desc.layers)
        d.layers = self._keep[0].ctypes.data
        d.n_constructions = len(desc.construction_offset) - 1
        d.construction_offset = self._keep[1].ctypes.data
        d.n_zones = z1 - z0
        d.zone_volume = self._keep[2].ctypes.data
        d.n_surfaces = s1 - s0
        d.surfaces = self._keep[3].ctypes.data
        d.terrain = -1 if desc.model_terrain is None else int(desc.model_terrain[m])
        d.steps_per_hour = int(desc.steps_per_hour)
        if discretization is None:
            self._m = lib.ho_model_new(C.byref(d))
        else:
            self._m = lib.ho_model_new_ex(C.byref(d), *[float(v) for v in discretization])
        if not self._m:
            raise OracleError(lib.ho_last_error().decode())
        self.n_zones = d.n_zones
        self.n_surfaces = d.n_surfaces
        self.dt_subdivisions = lib.ho_dt_subdivisions(self._m)
        self.dt = lib.ho_dt(self._m)
        self.state_len = lib.ho_state_len(self._m)
        self.zone_inputs = None
        self._zin_keep = None

    def __del__(self):
        if getattr(self, "_m", None):
            self._lib.ho_model_free(self._m)
            self._m = None

    @property
    def handle(self):
        return self._m

    def initial_state(self) -> np.ndarray:
        s = np.zeros(self.state_len)
        self._lib.ho_initial_state(self._m, s.ctypes.data)
        return s

    def n_nodes(self, s: int) -> int:
        return self._lib.ho_surface_n_nodes(self._m, s)

    def state_offset(self, s: int) -> int:
        return self._lib.ho_surface_state_offset(self._m, s)

    def surface_arrays(self, s: int, n_layers: int) -> dict:
        lib = self._lib
        n = self.n_nodes(s)
        el = np.zeros(max(n_layers, 1), np.int32)
        nl = lib.ho_surface_n_elements(self._m, s, el.ctypes.data, el.size)
        mass, u, fa, ba = np.zeros(n), np.zeros(n), np.zeros(n), np.zeros(n)
        lib.ho_surface_segments(self._m, s, mass.ctypes.data, u.ctypes.data)
        lib.ho_surface_alphas(self._m, s, fa.ctypes.data, ba.ctypes.data)
        buf = np.zeros(2 * (n + 1), np.int32)
        nm = lib.ho_surface_chunks(self._m, s, 1, buf.ctypes.data, buf.size)
        massive = buf[:2 * nm].reshape(-1, 2).copy()
        nn = lib.ho_surface_chunks(self._m, s, 0, buf.ctypes.data, buf.size)
        nomass = buf[:2 * nn].reshape(-1, 2).copy()
        return {"n_elements": el[:nl].copy(), "mass": mass, "u": u, "front_alpha": fa, "back_alpha": ba,
                "massive_chunks": massive, "nomass_chunks": nomass, "n_nodes": n,
                "tstep_subdivision": lib.ho_surface_tstep_subdivision(self._m, s),
                "wind_speed_modifier": lib.ho_surface_wind_modifier(self._m, s),
                "front_emissivity": lib.ho_surface_emissivity(self._m, s, 0),
                "back_emissivity": lib.ho_surface_emissivity(self._m, s, 1)}

    def set_zone_inputs(self, q_internal=None, infiltration_volume=None, infiltration_temperature=None,
                        ventilation_volume=None, ventilation_temperature=None):
        arrs = [None if a is None else _f8(a) for a in (q_internal, infiltration_volume, infiltration_temperature,
                                                        ventilation_volume, ventilation_temperature)]
        z = ZoneInputsC(*[None if a is None else a.ctypes.data for a in arrs])
        self._zin_keep = arrs
        self.zone_inputs = z

    def _zin(self):
        return C.byref(self.zone_inputs) if self.zone_inputs is not None else None

    def march(self, state: np.ndarray, weather_rows) -> None:
        """One main step; weather_rows [dt_subdivisions][3] = dry bulb, wind direction [deg], wind speed."""
        w = _f8(weather_rows)
        assert w.shape == (self.dt_subdivisions, 3), w.shape
        assert state.dtype == np.float64 and state.flags["C_CONTIGUOUS"] and state.size == self.state_len
        if self._lib.ho_march(self._m, state.ctypes.data, w.ctypes.data, self._zin()):
            raise OracleError(self._lib.ho_last_error().decode())

    def iterate_surfaces(self, state, wind_direction_rad, wind_speed, t_out, dt) -> None:
        if self._lib.ho_iterate_surfaces(self._m, state.ctypes.data, wind_direction_rad, wind_speed, t_out, dt):
            raise OracleError(self._lib.ho_last_error().decode())

    def zones_abc(self, state):
        a, b, c = np.zeros(self.n_zones), np.zeros(self.n_zones), np.zeros(self.n_zones)
        if self._lib.ho_zones_abc(self._m, state.ctypes.data, self._zin(), a.ctypes.data, b.ctypes.data, c.ctypes.data):
            raise OracleError(self._lib.ho_last_error().decode())
        return a, b, c

    @property
    def nomass_iterations(self) -> int:
        return self._lib.ho_nomass_iterations(self._m)

    # ---- series drivers (the recipe of tests/validate_wall_heat_transfer.rs:615-711, generalised) ----
    def march_series(self, state, n_steps, weather, *, weather_mode=0, inputs=None, input_mode=0, ir_mode=0,
                     record_every=0, record_state=False, surface_base=0, n_surfaces_total=None):
        """Same semantics as hb_march_series for ONE model. weather [rows][3]; inputs per hb_series for this model:
        INPUT_SITE [n_steps][4]; INPUT_PER_SURFACE [n_steps][n_surfaces_total][4] (this model's surfaces start at
        surface_base)."""
        S = self.dt_subdivisions
        weather = _f8(weather)
        zone_rec, state_rec = [], []
        offs = [self.state_offset(s) for s in range(self.n_surfaces)]
        nn = [self.n_nodes(s) for s in range(self.n_surfaces)]
        eps = [self._lib.ho_surface_emissivity(self._m, s, 0) for s in range(self.n_surfaces)]
        areas = self._keep[3]["area"]
        sigma = 5.670374419e-8
        ir_state0 = [state[o + 6] for o in offs]
        rows = np.empty((S, 3))
        for step in range(n_steps):
            if input_mode != 0:
                for s in range(self.n_surfaces):
                    row = inputs[step] if input_mode == 1 else inputs[step, surface_base + s]
                    o = offs[s]
                    state[o + 4], state[o + 5], state[o + 6], state[o + 7] = row[0], row[1], row[2], row[3]
                    if ir_mode == 1:
                        if eps[s] > 1e-3:
                            tk = state[o + 8] + 273.15
                            state[o + 6] = row[2] / areas[s] / eps[s] + sigma * ((tk * tk) * (tk * tk))
                        else:
                            state[o + 6] = ir_state0[s]
            if weather_mode == 0:
                rows[:] = weather[step]
            elif weather_mode == 1:
                for s in range(S):
                    f = float(s + 1) / float(S)
                    rows[s] = weather[step] + (weather[step + 1] - weather[step]) * f
            else:
                rows[:] = weather[step * S:(step + 1) * S]
            self.march(state, rows)
            if record_every and (step + 1) % record_every == 0:
                zone_rec.append(state[:self.n_zones].copy())
                if record_state:
                    state_rec.append(state.copy())
        return (np.array(zone_rec) if record_every else None), (np.array(state_rec) if record_state else None)


# ---- pure-function hooks ----
def nusselt(ra, gamma, a_gi):
    return load().ho_nusselt(ra, gamma, a_gi)


def gas_property(gas, which, t):
    return load().ho_gas_property(gas, which, t)


def cavity_u_value(gas, thickness, height, angle, eout, ein, t_front, t_back):
    return load().ho_cavity_u_value(gas, thickness, height, angle, eout, ein, t_front, t_back)


def tarp_natural(t_air, t_surf, cos_tilt):
    return load().ho_tarp_natural(t_air, t_surf, cos_tilt)


def tarp(t_air, t_surf, cos_tilt, air_speed, area, perimeter, windward):
    return load().ho_tarp_forced_plus_natural(t_air, t_surf, cos_tilt, air_speed, area, perimeter, int(windward))


def is_windward(wind_direction_rad, cos_tilt, normal):
    n = _f8(normal)
    return bool(load().ho_is_windward(wind_direction_rad, cos_tilt, n.ctypes.data))


def wind_speed_modifier(height, terrain=-1):
    return load().ho_wind_speed_modifier(height, terrain)


def rk4_steps(k, c, q, temps, dt, nsteps, rearrange=True):
    k = _f8(k).copy()
    c = _f8(c).copy()
    q = _f8(q).copy()
    t = _f8(temps).copy()
    n = t.size
    load().ho_rk4_steps(n, k.ctypes.data, c.ctypes.data, q.ctypes.data, t.ctypes.data, dt, nsteps, int(rearrange))
    return t, k, q


def discretize(layers, model_dt, max_dx, min_dt, height=1.0, angle=0.0, cap=512):
    layers = np.ascontiguousarray(layers)
    tsub = C.c_int32()
    el = np.zeros(len(layers), np.int32)
    mass, u = np.zeros(cap), np.zeros(cap)
    n = load().ho_discretize(layers.ctypes.data, len(layers), model_dt, max_dx, min_dt, height, angle, C.byref(tsub),
                             el.ctypes.data, mass.ctypes.data, u.ctypes.data, cap)
    if n < 0:
        raise OracleError(load().ho_last_error().decode())
    return tsub.value, el, mass[:n].copy(), u[:n].copy()


def build_segments(layers, tstep_subdivision, n_elements, height=1.0, angle=0.0, cap=512):
    """Discretization::build on given element counts → (mass[n_nodes], u[n_nodes]) (u: solid k/dx, cavity -1, back 0)."""
    layers = np.ascontiguousarray(layers)
    el = np.ascontiguousarray(n_elements, np.int32)
    mass, u = np.zeros(cap), np.zeros(cap)
    n = load().ho_build(layers.ctypes.data, len(layers), tstep_subdivision, el.ctypes.data, height, angle,
                        mass.ctypes.data, u.ctypes.data, cap)
    if n < 0:
        raise OracleError(load().ho_last_error().decode())
    return mass[:n].copy(), u[:n].copy()


def chunks(mass, massive: bool):
    mass = _f8(mass)
    out = np.zeros(2 * mass.size + 2, np.int32)
    n = load().ho_chunks(mass.ctypes.data, mass.size, int(massive), out.ctypes.data, out.size)
    if n < 0:
        raise OracleError(load().ho_last_error().decode())
    return [(int(out[2 * i]), int(out[2 * i + 1])) for i in range(n)]


def get_k_q(mass, u, ini, fin, temps, front, back):
    """front/back = (air_temperature, rad_temperature, hs, rad_hs). Returns dense K[(fin-ini)²], q[fin-ini]."""
    mass, u, temps = _f8(mass), _f8(u), _f8(temps)
    m = fin - ini
    k, q = np.zeros((m, m)), np.zeros(m)
    rc = load().ho_get_k_q(mass.size, mass.ctypes.data, u.ctypes.data, ini, fin, temps.ctypes.data, *map(float, front),
                           *map(float, back), k.ctypes.data, q.ctypes.data)
    if rc:
        raise OracleError(load().ho_last_error().decode())
    return k, q


def glazing_combine(tau, rho_front, rho_back):
    """ISO 9050 combination → (tau, rho_front, rho_back, alpha_front, alpha_back)."""
    tau, rf, rb = _f8(tau), _f8(rho_front), _f8(rho_back)
    out = np.zeros(5)
    if load().ho_glazing_combine(tau.ctypes.data, rf.ctypes.data, rb.ctypes.data, tau.size, out.ctypes.data):
        raise OracleError(load().ho_last_error().decode())
    return out


def glazing_alphas(tau, rho_front, rho_back):
    tau, rf, rb = _f8(tau), _f8(rho_front), _f8(rho_back)
    out = np.zeros(tau.size)
    n = load().ho_glazing_alphas(tau.ctypes.data, rf.ctypes.data, rb.ctypes.data, tau.size, out.ctypes.data)
    if n < 0:
        raise OracleError(load().ho_last_error().decode())
    return out[:n]


def run_ensemble(models, n_steps, weather, surf_inputs=None, weather_linear=False, threads=1, want_state=False,
                 q_internal=None):
    """Time-able CPU baseline: each model marched on one of `threads` host threads. Returns (updates, states).
    `q_internal`: HVAC + luminaire power per zone, models concatenated."""
    lib = load()
    arr = (C.c_void_p * len(models))(*[m.handle for m in models])
    weather = _f8(weather)
    si = None if surf_inputs is None else _f8(surf_inputs)
    total = sum(m.state_len for m in models)
    out = np.zeros(total) if want_state else None
    q = None if q_internal is None else _f8(q_internal)
    n = lib.ho_run_ensemble_q(arr, len(models), n_steps, weather.ctypes.data, None if si is None else si.ctypes.data,
                              int(weather_linear), threads, None if out is None else out.ctypes.data,
                              None if q is None else q.ctypes.data)
    if n < 0:
        raise OracleError(lib.ho_last_error().decode())
    return n, out

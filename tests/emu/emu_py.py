"""ctypes wrapper of tests/emu/libhb_emu.so — TEST INFRASTRUCTURE (see emu.cc).

`EmuBatch` offers the subset of `heat_b200.model.ThermalBatch` that tests/parity_cases.py uses, so the same cases can be
replayed through the host re-enactment of the CUDA kernel's data flow on a machine without a GPU.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

from heat_b200 import _lib

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libhb_emu.so")
_emu = None


def load():
    global _emu
    if _emu is None:
        subprocess.run(["make", "-C", _HERE], check=True, capture_output=True)
        lib = C.CDLL(LIB_PATH)
        lib.emu_last_error.restype = C.c_char_p
        lib.emu_state_len.restype = C.c_longlong
        lib.emu_state_len.argtypes = [C.POINTER(_lib.BatchDescC)]
        lib.emu_n_tiles.restype = C.c_longlong
        lib.emu_n_tiles.argtypes = [C.POINTER(_lib.BatchDescC)]
        lib.emu_tile_table.restype = C.c_longlong
        lib.emu_tile_table.argtypes = [C.POINTER(_lib.BatchDescC), C.c_void_p, C.c_longlong]
        lib.emu_run.restype = C.c_int
        lib.emu_run.argtypes = [C.POINTER(_lib.BatchDescC), C.c_int] + [C.c_void_p] * 6 + [
            C.POINTER(_lib.SeriesC), C.POINTER(C.c_longlong), C.POINTER(C.c_longlong)]
        _emu = lib
    return _emu


class EmuBatch:
    def __init__(self, desc):
        self.lib = load()
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
        self._d = d
        self.state_len = self.lib.emu_state_len(C.byref(d))
        if self.state_len < 0:
            raise RuntimeError(self.lib.emu_last_error().decode())
        self.n_zones = int(desc.model_zone_offset[-1])
        self.n_surfaces = int(desc.model_surface_offset[-1])
        self.n_models = desc.n_models
        self._state = None
        self._zone_inputs = [None] * 5
        self.nomass_iterations = 0
        self.status = 0

    @property
    def n_tiles(self) -> int:
        return int(self.lib.emu_n_tiles(C.byref(self._d)))

    def tile_table(self) -> np.ndarray:
        """[n_tiles][5] = (G, NL, W, flags, S) of the plan the library builds for this batch."""
        n = self.n_tiles
        out = np.zeros((max(n, 1), 5), np.int32)
        got = self.lib.emu_tile_table(C.byref(self._d), out.ctypes.data, n)
        assert got == n
        return out[:n]

    def get_state(self):
        if self._state is None:
            # run zero steps to obtain the initial values
            self._run(0, np.zeros((1, 1, 3)), 0, None, 0, 0, None, 0, False)
        return self._state.copy()

    def set_state(self, s):
        self._state = np.ascontiguousarray(s, np.float64).copy()

    def set_zone_inputs(self, q_internal=None, infiltration_volume=None, infiltration_temperature=None,
                        ventilation_volume=None, ventilation_temperature=None):
        self._zone_inputs = [None if a is None else _lib.f8(a) for a in
                             (q_internal, infiltration_volume, infiltration_temperature, ventilation_volume,
                              ventilation_temperature)]

    def _run(self, n_steps, weather, weather_mode, inputs, input_mode, ir_mode, model_site, record_every, record_state,
             zone_gain=None, schedules=None):
        s = _lib.SeriesC()
        weather = _lib.f8(weather)
        if weather.ndim == 2:
            weather = weather[None]
        s.n_steps, s.weather_mode, s.n_sites = int(n_steps), int(weather_mode), weather.shape[0]
        s.weather_rows, s.weather = weather.shape[1], _lib.ptr(weather)
        if model_site is not None:
            model_site = np.ascontiguousarray(model_site, np.int32)
            s.model_site = _lib.ptr(model_site)
        s.input_mode, s.ir_mode = int(input_mode), int(ir_mode)
        if inputs is not None:
            inputs = _lib.f8(inputs)
            s.inputs = _lib.ptr(inputs)
        if zone_gain is not None:
            zone_gain = _lib.f8(zone_gain)
            assert zone_gain.size == n_steps * self.n_zones
            s.zone_gain = _lib.ptr(zone_gain)
        keep = []
        for name, table in (schedules or {}).items():
            if table is not None:
                table = _lib.f8(table)
                assert table.size == n_steps * self.n_zones
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
        use_initial = self._state is None
        state = np.zeros(self.state_len) if use_initial else self._state
        it, st = C.c_longlong(0), C.c_longlong(0)
        rc = self.lib.emu_run(C.byref(self._d), int(use_initial), _lib.ptr(state),
                              *[_lib.ptr(a) for a in self._zone_inputs], C.byref(s), C.byref(it), C.byref(st))
        if rc != 0:
            raise RuntimeError(self.lib.emu_last_error().decode())
        self._state = state
        self.nomass_iterations += it.value
        self.status |= st.value
        return zone_rec, state_rec

    def march_series(self, n_steps, weather, *, weather_mode=0, inputs=None, input_mode=0, ir_mode=0, model_site=None,
                     record_every=0, record_state=False, zone_gain=None, **schedules):
        return self._run(n_steps, weather, weather_mode, inputs, input_mode, ir_mode, model_site, record_every,
                         record_state, zone_gain, schedules)


def run_emu(name, max_steps=None, record_state=False):
    """Same contract as parity_cases.run_gpu, through the emulator."""
    import parity_cases as pc
    c = pc.make_case(name, max_steps)
    b = EmuBatch(c.desc)
    if c.zone_t0 is not None:
        s = b.get_state()
        s[0] = c.zone_t0
        b.set_state(s)
    if c.zone_inputs:
        b.set_zone_inputs(**c.zone_inputs)
    zone, states = b.march_series(c.n_steps, c.weather, weather_mode=c.weather_mode, inputs=c.inputs,
                                  input_mode=c.input_mode, ir_mode=c.ir_mode, record_every=1, record_state=record_state)
    final = b.get_state()
    if record_state:
        return zone, final, states, b
    return zone, final

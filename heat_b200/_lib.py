"""ctypes binding of libheat_b200.so (C ABI: include/heat_b200.h). No CPU fallback: a missing library raises."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("HEAT_B200_LIB") or os.path.join(_HERE, "libheat_b200.so")

HB_OK, HB_ERR_INVALID, HB_ERR_CUDA, HB_ERR_PHYSICS, HB_ERR_UNSUPPORTED = 0, 1, 2, 3, 4
WEATHER_STEP_CONSTANT, WEATHER_LINEAR, WEATHER_PER_SUBSTEP = 0, 1, 2
INPUT_STATE, INPUT_SITE, INPUT_PER_SURFACE = 0, 1, 2
IR_IRRADIANCE, IR_NET_GAIN = 0, 1


class HeatB200Error(RuntimeError):
    def __init__(self, status: int, message: str):
        super().__init__(f"[hb_status {status}] {message}")
        self.status = status
        self.message = message


class BatchDescC(C.Structure):
    _fields_ = [
        ("n_layers", C.c_int32), ("layers", C.c_void_p),
        ("n_constructions", C.c_int32), ("construction_offset", C.c_void_p),
        ("n_models", C.c_int32), ("model_zone_offset", C.c_void_p), ("model_surface_offset", C.c_void_p),
        ("zone_volume", C.c_void_p), ("surfaces", C.c_void_p), ("model_terrain", C.c_void_p),
        ("steps_per_hour", C.c_int32),
    ]


class ModelInfoC(C.Structure):
    _fields_ = [
        ("n_zones", C.c_int32), ("n_surfaces", C.c_int32), ("zone_begin", C.c_int32), ("surface_begin", C.c_int32),
        ("dt_subdivisions", C.c_int32), ("dt", C.c_double), ("state_offset", C.c_int64), ("state_len", C.c_int32),
        ("nodes", C.c_int64),
    ]


class SurfaceInfoC(C.Structure):
    _fields_ = [
        ("n_nodes", C.c_int32), ("n_layers", C.c_int32), ("tstep_subdivision", C.c_int32),
        ("n_massive_chunks", C.c_int32), ("n_nomass_chunks", C.c_int32), ("state_offset", C.c_int64),
        ("front_emissivity", C.c_double), ("back_emissivity", C.c_double), ("wind_speed_modifier", C.c_double),
        ("cos_tilt", C.c_double),
    ]


class SeriesC(C.Structure):
    _fields_ = [
        ("n_steps", C.c_int32), ("weather_mode", C.c_int32), ("n_sites", C.c_int32), ("model_site", C.c_void_p),
        ("weather", C.c_void_p), ("weather_rows", C.c_int64), ("input_mode", C.c_int32), ("ir_mode", C.c_int32),
        ("inputs", C.c_void_p), ("record_every", C.c_int32), ("zone_t_record", C.c_void_p),
        ("state_record", C.c_void_p), ("zone_gain", C.c_void_p),
        ("zone_infiltration_volume", C.c_void_p), ("zone_infiltration_temperature", C.c_void_p),
        ("zone_ventilation_volume", C.c_void_p), ("zone_ventilation_temperature", C.c_void_p),
    ]


ZONE_SCHEDULES = ("zone_gain", "zone_infiltration_volume", "zone_infiltration_temperature", "zone_ventilation_volume",
                  "zone_ventilation_temperature")


class EnsembleStatsC(C.Structure):
    _fields_ = [
        ("sum_zone_t", C.c_double), ("min_zone_t", C.c_double), ("max_zone_t", C.c_double),
        ("sum_node_t", C.c_double), ("min_node_t", C.c_double), ("max_node_t", C.c_double),
        ("front_conv_energy_j", C.c_double), ("back_conv_energy_j", C.c_double),
        ("n_zones", C.c_int64), ("n_nodes", C.c_int64), ("nomass_iterations", C.c_int64),
        ("status_flags", C.c_int64),
    ]


# every symbol include/heat_b200.h declares: (restype, argtypes)
_P = C.c_void_p
SYMBOLS = {
    "hb_last_error": (C.c_char_p, []),
    "hb_version": (C.c_char_p, []),
    "hb_create": (C.c_int32, [C.POINTER(BatchDescC), C.c_int32, C.POINTER(_P)]),
    "hb_destroy": (None, [_P]),
    "hb_n_models": (C.c_int32, [_P]),
    "hb_n_surfaces_total": (C.c_int64, [_P]),
    "hb_n_zones_total": (C.c_int64, [_P]),
    "hb_state_len_total": (C.c_int64, [_P]),
    "hb_nodes_total": (C.c_int64, [_P]),
    "hb_updates_per_step": (C.c_int64, [_P]),
    "hb_get_model_info": (C.c_int32, [_P, C.c_int32, C.POINTER(ModelInfoC)]),
    "hb_get_surface_info": (C.c_int32, [_P, C.c_int64, C.POINTER(SurfaceInfoC)]),
    "hb_get_surface_arrays": (C.c_int32, [_P, C.c_int64, _P, _P, _P, _P, _P, _P, _P]),
    "hb_surface_rk4_number": (C.c_int32, [_P, C.c_int64, C.c_double, _P]),
    "hb_set_state": (C.c_int32, [_P, _P]),
    "hb_get_state": (C.c_int32, [_P, _P]),
    "hb_set_zone_inputs": (C.c_int32, [_P, _P, _P, _P, _P, _P]),
    "hb_march_series": (C.c_int32, [_P, C.POINTER(SeriesC)]),
    "hb_upload": (C.c_int32, [_P, _P, C.c_int64, C.POINTER(_P)]),
    "hb_free_dev": (C.c_int32, [_P, _P]),
    "hb_march_series_dev": (C.c_int32, [_P, C.POINTER(SeriesC), C.POINTER(C.c_float)]),
    "hb_launch_count": (C.c_int64, [_P]),
    "hb_reduce": (C.c_int32, [_P, C.POINTER(EnsembleStatsC)]),
    "hb_reduce_noalloc": (C.c_int32, [_P, _P, _P]),
    "hb_reset_energy": (C.c_int32, [_P]),
    "hb_comm_unique_id": (C.c_int32, [_P]),
    "hb_comm_init": (C.c_int32, [_P, _P, C.c_int32, C.c_int32]),
    "hb_comm_destroy": (C.c_int32, [_P]),
    "hb_reduce_all": (C.c_int32, [_P, C.POINTER(EnsembleStatsC)]),
    "hb_measure_fp64_peak": (C.c_int32, [C.c_int32, C.POINTER(C.c_double)]),
    "hb_flush_l2": (C.c_int32, [_P, C.c_int64]),
    "hb_read_epw": (C.c_int32, [C.c_char_p, _P, _P, C.c_int64, C.POINTER(C.c_int64)]),
}

_lib = None


def _point_at_bundled_nccl() -> None:
    """HEAT_B200_NCCL (read by hb_comm_*): the libnccl.so.2 of the `nvidia-nccl` wheel torch is built against, when there
    is one — a process holds one NCCL, and if this library loaded the system's older copy first, a later `import torch`
    would miss symbols.  Nothing is imported or loaded here; a value set by the user wins."""
    if os.environ.get("HEAT_B200_NCCL"):
        return
    try:
        import importlib.util
        spec = importlib.util.find_spec("nvidia.nccl")
        for root in (spec.submodule_search_locations if spec else ()):
            path = os.path.join(root, "lib", "libnccl.so.2")
            if os.path.exists(path):
                os.environ["HEAT_B200_NCCL"] = path
                return
    except (ImportError, ValueError, AttributeError):
        pass


def load() -> C.CDLL:
    """Load the CUDA library. Raises if it has not been built (`python -c 'import __graft_entry__ as g; g.build()'`)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `make -C heat_b200/csrc` (there is no CPU fallback)")
        _point_at_bundled_nccl()
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def check(status: int) -> None:
    if status != HB_OK:
        raise HeatB200Error(status, load().hb_last_error().decode("utf-8", "replace"))


def ptr(a) -> int:
    """Host pointer of a C-contiguous numpy array (None → NULL)."""
    if a is None:
        return None
    assert a.flags["C_CONTIGUOUS"]
    return a.ctypes.data


def f8(a) -> np.ndarray:
    return np.ascontiguousarray(a, dtype=np.float64)

"""The C-ABI library loads and exports every symbol include/heat_b200.h declares (no compute calls here)."""
import ctypes
import os
import re

from heat_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "heat_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(hb_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_are_exported_and_bound():
    lib = ctypes.CDLL(_lib.LIB_PATH)
    names = declared_symbols()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), f"{n} declared in heat_b200.h but not exported"
        assert n in _lib.SYMBOLS, f"{n} has no ctypes binding"
    assert sorted(_lib.SYMBOLS) == names


def test_struct_layouts_match_header():
    from heat_b200 import desc
    assert desc.LAYER_DTYPE.itemsize == 80      # hb_layer: 2×i32 + 9×f64
    assert desc.SURFACE_DTYPE.itemsize == 120   # hb_surface: 6×i32 + 12×f64
    assert ctypes.sizeof(_lib.BatchDescC) == 88
    assert ctypes.sizeof(_lib.SeriesC) == 120     # hb_series incl. zone_gain (offset 80) + 4 infiltration / ventilation tables
    assert _lib.SeriesC.zone_gain.offset == 80 and _lib.SeriesC.zone_ventilation_temperature.offset == 112
    assert ctypes.sizeof(_lib.ModelInfoC) == 56
    assert ctypes.sizeof(_lib.SurfaceInfoC) == 64
    assert ctypes.sizeof(_lib.EnsembleStatsC) == 96


def test_version_and_error_string():
    lib = _lib.load()
    assert b"sm_100a" in lib.hb_version()
    assert isinstance(lib.hb_last_error(), bytes)


def test_product_does_not_touch_the_oracle():
    """Nothing under heat_b200/ may import, link or execute oracle/ (it is test infrastructure)."""
    pkg = os.path.join(ROOT, "heat_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cc", ".cu", ".cuh", ".hpp", ".h")) or f == "Makefile":
                text = open(os.path.join(dirpath, f), errors="replace").read()
                assert "oracle" not in text.lower(), os.path.join(dirpath, f)

"""Scratch: a handful of long lean walls (S >= 100, 16 < N <= 24 -> G = 2, NL = 12) alone, one series; for ncu --set full."""
import sys, os, ctypes as C
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT]
from heat_b200 import cases, _lib, desc as D
from heat_b200.model import ThermalBatch
n, hours = 30000, int(sys.argv[1]) if len(sys.argv) > 1 else 48
lo, hi = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (16, 24)
d = cases.wall_ensemble(n, seed=42)
plan = ThermalBatch(d, device=-1)
info = _lib.SurfaceInfoC(); mi = _lib.ModelInfoC()
N = np.zeros(n, int); S = np.zeros(n, int); NM = np.zeros(n, int)
for s in range(n):
    plan._lib.hb_get_surface_info(plan._h, s, C.byref(info)); N[s] = info.n_nodes; NM[s] = info.n_nomass_chunks
    plan._lib.hb_get_model_info(plan._h, s, C.byref(mi)); S[s] = mi.dt_subdivisions
plan.close()
idx = np.nonzero((S >= 100) & (NM == 0) & (N > lo) & (N <= hi))[0][:16]
k = len(idx)
ar = np.arange(k + 1, dtype=np.int32)
sub = D.BatchDesc(d.layers, d.construction_offset, ar, ar, d.zone_volume[idx], d.surfaces[idx], d.steps_per_hour)
weather, inputs = cases.synthetic_weather(hours + 1)
b = ThermalBatch(sub)
w = b.upload(weather); i = b.upload(np.ascontiguousarray(inputs[:hours]))
for rep in range(2):
    ms = b.march_series_dev(hours, w, hours + 1, 1, weather_mode=_lib.WEATHER_LINEAR, inputs_dev=i, input_mode=_lib.INPUT_SITE)
print("walls", k, "S", S[idx], "N", N[idx], "ms", ms)

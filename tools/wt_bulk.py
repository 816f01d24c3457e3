"""Scratch: throughput of the bulk of the walls ensemble — only walls with smin <= S < smax (no long tiles)."""
import sys, os, ctypes as C
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT]
from heat_b200 import cases, _lib, desc as D
from heat_b200.model import ThermalBatch

n = int(sys.argv[1]); hours = int(sys.argv[2]); smin = int(sys.argv[3]); smax = int(sys.argv[4])
d = cases.wall_ensemble(n, seed=42)
plan = ThermalBatch(d, device=-1)
info = _lib.SurfaceInfoC(); mi = _lib.ModelInfoC()
N = np.zeros(n, int); S = np.zeros(n, int); NM = np.zeros(n, int)
for s in range(n):
    plan._lib.hb_get_surface_info(plan._h, s, C.byref(info)); N[s] = info.n_nodes; NM[s] = info.n_nomass_chunks
    plan._lib.hb_get_model_info(plan._h, s, C.byref(mi)); S[s] = mi.dt_subdivisions
plan.close()
idx = np.nonzero((S >= smin) & (S < smax))[0]
k = len(idx)
ar = np.arange(k + 1, dtype=np.int32)
sub = D.BatchDesc(d.layers, d.construction_offset, ar, ar, d.zone_volume[idx], d.surfaces[idx], d.steps_per_hour)
weather, inputs = cases.synthetic_weather(hours + 1)
b = ThermalBatch(sub)
w = b.upload(weather); i = b.upload(np.ascontiguousarray(inputs[:hours]))
for rep in range(3):
    ms = b.march_series_dev(hours, w, hours + 1, 1, weather_mode=_lib.WEATHER_LINEAR, inputs_dev=i, input_mode=_lib.INPUT_SITE)
    print(f"S in [{smin},{smax}) walls {k} updates/step {b.updates_per_step}: rep {rep} {ms:.2f} ms -> {b.updates_per_step*hours/(ms*1e-3):.3e} updates/s", flush=True)

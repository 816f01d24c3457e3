"""Scratch: latency probe of the warp-tile kernel — a few long tiles (S >= smin) alone on the GPU, lean and no-mass-end
walls separately; with the phase-clock build (HEAT_B200_LIB=heat_b200/libheat_b200_clk.so) the cycles per phase print."""
import sys, os, ctypes as C
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT]
from heat_b200 import cases, _lib, desc as D
from heat_b200.model import ThermalBatch

n = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
hours = int(sys.argv[2]) if len(sys.argv) > 2 else 240
smin = int(sys.argv[3]) if len(sys.argv) > 3 else 100
keep = int(sys.argv[4]) if len(sys.argv) > 4 else 16
d = cases.wall_ensemble(n, seed=42)
plan = ThermalBatch(d, device=-1)
info = _lib.SurfaceInfoC(); mi = _lib.ModelInfoC()
N = np.zeros(n, int); S = np.zeros(n, int); NM = np.zeros(n, int)
for s in range(n):
    plan._lib.hb_get_surface_info(plan._h, s, C.byref(info)); N[s] = info.n_nodes; NM[s] = info.n_nomass_chunks
    plan._lib.hb_get_model_info(plan._h, s, C.byref(mi)); S[s] = mi.dt_subdivisions
plan.close()


def pick(idx):
    idx = np.asarray(idx)
    k = len(idx)
    ar = np.arange(k + 1, dtype=np.int32)
    return D.BatchDesc(d.layers, d.construction_offset, ar, ar, d.zone_volume[idx], d.surfaces[idx], d.steps_per_hour)


weather, inputs = cases.synthetic_weather(hours + 1)
for label, mask in (("lean", (S >= smin) & (NM == 0)), ("end", (S >= smin) & (NM > 0))):
    for lo, hi in ((0, 8), (8, 16), (16, 32), (32, 64), (64, 128)):
        idx = np.nonzero(mask & (N > lo) & (N <= hi))[0][:keep]
        if idx.size == 0:
            continue
        b = ThermalBatch(pick(idx))
        w = b.upload(weather); i = b.upload(np.ascontiguousarray(inputs[:hours]))
        for rep in range(2):
            ms = b.march_series_dev(hours, w, hours + 1, 1, weather_mode=_lib.WEATHER_LINEAR, inputs_dev=i, input_mode=_lib.INPUT_SITE)
        smax = int(S[idx].max())
        print(f"{label} N in ({lo},{hi}] walls {idx.size} Smax {smax}: {ms:.2f} ms -> {ms*1e-3*1.965e9/(hours*smax):.0f} cycles per sub-step of the longest tile", flush=True)
        b.close()

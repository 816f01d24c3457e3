"""Scratch: one short series launch for ncu (walls workload)."""
import sys, os
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT]
from heat_b200 import cases, _lib
from heat_b200.model import ThermalBatch
units = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
hours = int(sys.argv[2]) if len(sys.argv) > 2 else 24
kind = sys.argv[3] if len(sys.argv) > 3 else "walls"
if kind == "walls":
    d = cases.wall_ensemble(units, seed=42)
    gains = None
elif kind == "buildings2g":
    d, gains = cases.shoebox_ensemble(units, seed=43, glazing=cases.double_glazing())
else:
    d, gains = cases.shoebox_ensemble(units, seed=43)
weather, inputs = cases.synthetic_weather(hours + 1)
b = ThermalBatch(d)
if gains is not None:
    b.set_zone_inputs(q_internal=gains)
w = b.upload(weather); i = b.upload(np.ascontiguousarray(inputs[:hours]))
for rep in range(2):
    ms = b.march_series_dev(hours, w, hours + 1, 1, weather_mode=_lib.WEATHER_LINEAR, inputs_dev=i, input_mode=_lib.INPUT_SITE)
    print("rep", rep, "ms", ms, "updates/s %.3e" % (b.updates_per_step * hours / (ms * 1e-3)))

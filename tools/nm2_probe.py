"""Scratch: phase clocks of a wall with two-node no-mass end chunks (HEAT_B200_LIB = clock build, HB_CLK_MIN_S=1)."""
import sys, os
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT]
from heat_b200 import cases, _lib
from heat_b200.model import ThermalBatch
poly2 = [("polyurethane", 0.01), ("polyurethane", 0.012)]
which = sys.argv[1] if len(sys.argv) > 1 else "back"
cons = {"back": [("concrete", 0.15)] + poly2, "both": poly2 + [("concrete", 0.2)] + poly2,
        "one": [("concrete", 0.15), ("polyurethane", 0.01)], "none": [("concrete", 0.15)]}[which]
d = cases.single_zone_test_building(cons, zone_volume=600.0, surface_height=3.0, surface_width=20.0, emissivity=0.9,
                                    solar_absorbtance=0.7, steps_per_hour=1, back_emissivity_override=0.0)
weather, inputs = cases.synthetic_weather(241)
b = ThermalBatch(d)
b.march_series(240, weather, weather_mode=_lib.WEATHER_LINEAR, inputs=inputs[:240], input_mode=_lib.INPUT_SITE)
print(which, "iters", b.reduce().nomass_iterations, "substeps", b.updates_per_step)

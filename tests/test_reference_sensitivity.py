"""Where parity with the reference is well defined — and where it is not.

The reference's no-mass fixed point (surface.rs:790-898: solve K·T* = −q with the radiative term explicit in q, then
T ← (T + T*)/2) iterates the map T → (T + T*(T))/2 whose slope at a one-node chunk is (1 − r)/2 with
r = 4εσT³ / (u + h).  For r > 3 the map is expanding: the loop leaves through `err > old_err`, and the value it leaves
with depends on the last bits of its inputs.  A zone-facing no-mass skin (u = 1.26 W/m²K of 2 cm polyurethane, natural
convection only, ε ≈ 0.9 ⇒ r ≈ 3…4) is in that regime.  The reference's validation driver sidesteps it by zeroing the
zone-facing emissivity (tests/validate_wall_heat_transfer.rs:630); heat_b200.cases.wall_ensemble does the same.
This test pins both facts on the CPU restatement itself, so that nobody "fixes" a parity failure that is the
reference's own chaos.
"""
import numpy as np

import oracle_py as O
from heat_b200 import cases, desc as D


def _mixed_wall(back_eps_override):
    layers = [cases.polyurethane(0.02, 0.92, 0.5),
              D.solid_layer(0.17, conductivity=0.66, density=1770.0, specific_heat=1710.0, front_emissivity=0.92,
                            back_emissivity=0.92, front_solar_abs=0.5, back_solar_abs=0.5),
              cases.polyurethane(0.02, 0.92, 0.5)]
    s = D.surface(0, 60.0, 46.0, (0.6, 0.8, 0.0), 1.5, front=(D.BOUNDARY_OUTDOOR, 0), back=(D.BOUNDARY_SPACE, 0),
                  back_emissivity=back_eps_override)
    return D.single_model([layers], [300.0], [s], 1)


def _divergence(desc, hours=6, eps=1e-12):
    weather, inputs = cases.synthetic_weather(hours + 1)
    out = []
    for bump in (0.0, eps):
        m = O.OracleModel(desc)
        s = m.initial_state()
        s[m.state_offset(0) + 8 + 10] += bump
        m.march_series(s, hours, weather, weather_mode=1, inputs=inputs[:hours], input_mode=1)
        out.append(s)
    return float(np.abs(out[0] - out[1]).max())


def test_zone_facing_nomass_skin_with_emissivity_is_chaotic_in_the_reference():
    assert _divergence(_mixed_wall(D.NAN)) > 1e-6      # 1e-12 in → far more than the parity tolerance out


def test_same_wall_with_zero_zone_facing_emissivity_is_well_conditioned():
    assert _divergence(_mixed_wall(0.0)) < 1e-9

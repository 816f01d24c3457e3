"""The oracle against EnergyPlus: the one whole-path external check the reference's fixtures offer.

tests/validate_wall_heat_transfer.rs:615-711 (`march_model`) drives `ThermalModel::march` with the boundary series of
an EnergyPlus run and records, for rows i > 5000, (expected = EnergyPlus' zone mean air temperature, found = the zone
temperature in the state BEFORE marching row i) — `:665-679`.  The reference only charts the pair (its SeriesValidator
sets no tolerance); here the RMSE ceilings are asserted.  They pin the pieces no reference unit test asserts:
`discretize_construction`'s integers, `calc_border_conditions`, TARP, `wind_speed_modifier` — an oracle that
mis-discretized or mis-signed any of them would leave these bands.

Ceilings = what this oracle measures today, rounded up (°C): massive ≤ 0.08, mixed ≤ 0.17, no-mass / tilted /
horizontal ≤ 0.25; max |error| ≤ 0.40.
"""
import numpy as np
import pytest

import parity_cases as pc

CEILING = {"massive": 0.08, "mixed": 0.17, "nomass": 0.25, "tilted": 0.25, "horizontal": 0.25}


@pytest.mark.parametrize("case", pc.EPLUS_CASES)
def test_oracle_zone_temperature_tracks_energyplus(case):
    expected = pc._npz("eplus_inputs.npz")[case][:, 5]          # eplusout.csv column 12: zone mean air temperature
    zone, _ = pc.run_oracle(f"eplus/{case}")                     # live oracle, all 10 080 rows
    zone = np.asarray(zone).reshape(-1)
    assert zone.size == expected.size == 10080
    i = np.arange(5001, expected.size)                           # `if i > 5000` (:665)
    err = zone[i - 1] - expected[i]                              # found = state before the march of row i
    rmse = float(np.sqrt(np.mean(err * err)))
    assert rmse <= CEILING[case.split("_")[0]], (case, rmse)
    assert float(np.abs(err).max()) <= 0.40, (case, float(np.abs(err).max()))
    # the committed golden output is this oracle's output
    zone_g, _ = pc.golden(f"eplus/{case}")
    assert np.abs(np.asarray(zone_g).reshape(-1) - zone).max() <= 1e-9

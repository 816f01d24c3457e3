"""Extract the boundary-condition columns the validation driver reads from the reference's EnergyPlus outputs and
weather file into small fixtures, and record oracle outputs for them.

Run HERE (the container with /root/reference mounted); the GPU box has no /root/reference, so tests read only the
.npz files this script writes.

    python tests/golden/make_fixtures.py

Inputs (reference tree):
  tests/<case>/eplusout.csv  — columns used by march_model (tests/validate_wall_heat_transfer.rs:636-650):
      1 wind speed, 2 wind direction, 3 incident solar, 8 outdoor dry bulb, 11 outside net LW gain, 12 zone T
  tests/epw/CHL_Santiago.855740_IWEC.epw — hourly records; fields 7 dry bulb (idx 6), 13 horizontal IR (idx 12),
      14 global horizontal (idx 13), 21 wind direction (idx 20), 22 wind speed (idx 21)   (1-based field numbers)
Outputs:
  eplus_inputs.npz   {case: float64[10080][6]}  (wind_speed, wind_dir, solar, t_out, lw_gain, zone_T)
  epw_santiago.npz   {hourly: float64[8760][5]} (dry_bulb, ir_horizontal, global_horizontal, wind_dir, wind_speed)
  oracle_golden.npz  oracle results for the parity cases (see tests/test_parity_gpu.py)
"""
import csv
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference/tests"
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))

CASES = [
    "massive_full", "massive_no_ir_no_solar", "massive_no_ir_yes_solar", "massive_yes_ir_no_solar",
    "mixed_full", "mixed_no_ir_no_solar", "mixed_no_ir_yes_solar", "mixed_yes_ir_no_solar",
    "nomass_full", "nomass_no_ir_no_solar", "nomass_no_ir_yes_solar", "nomass_yes_ir_no_solar",
    "tilted", "horizontal",
]
COLS = [1, 2, 3, 8, 11, 12]


def read_eplus(case):
    rows = []
    with open(os.path.join(REF, case, "eplusout.csv")) as f:
        r = csv.reader(f)
        next(r)
        for line in r:
            rows.append([float(line[c]) for c in COLS])
    return np.array(rows)


def read_epw():
    rows = []
    with open(os.path.join(REF, "epw", "CHL_Santiago.855740_IWEC.epw")) as f:
        for i, line in enumerate(f):
            if i < 8:
                continue
            p = line.strip().split(",")
            rows.append([float(p[6]), float(p[12]), float(p[13]), float(p[20]), float(p[21])])
    return np.array(rows)


def main():
    eplus = {c: read_eplus(c) for c in CASES}
    for c, a in eplus.items():
        assert a.shape == (10080, 6), (c, a.shape)
    np.savez_compressed(os.path.join(HERE, "eplus_inputs.npz"), **eplus)
    epw = read_epw()
    assert epw.shape == (8760, 5), epw.shape
    np.savez_compressed(os.path.join(HERE, "epw_santiago.npz"), hourly=epw)

    # oracle outputs (golden): built by the same drivers the tests use
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import parity_cases as pc
    gold = {}
    for name in pc.GOLDEN_CASES:
        zone, final = pc.run_oracle(name)
        gold[name + "/zone"] = zone
        gold[name + "/final"] = final
        print(name, zone.shape, final.shape, float(zone[-1, 0]))
    np.savez_compressed(os.path.join(HERE, "oracle_golden.npz"), **gold)


if __name__ == "__main__":
    main()

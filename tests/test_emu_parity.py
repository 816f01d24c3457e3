"""CPU check of the DEVICE arithmetic and packing: tests/emu re-enacts hb_march_kernel's data flow on the host with the
same HB_HD functions (heat_b200/csrc/hb_math.cuh) and the same packed buffers, and must agree with the oracle.

This is test infrastructure (no GPU needed); the parity tests proper are in test_parity_gpu.py.
"""
import os
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "emu"))

import parity_cases as pc  # noqa: E402
import emu_py  # noqa: E402

TOL = 1e-6
FAST = [n for n in pc.GOLDEN_CASES if n not in ("epw/mixed_full", "epw/nomass_full")]


@pytest.mark.parametrize("name", FAST)
def test_emulated_kernel_against_golden(name):
    zone_g, final_g = pc.golden(name)
    zone, final = emu_py.run_emu(name)
    assert np.abs(zone - zone_g).max() <= TOL
    assert np.abs(final - final_g).max() <= 1e-5


@pytest.mark.parametrize("name", ["epw/mixed_full", "epw/nomass_full"])
def test_emulated_kernel_long_cases_prefix(name):
    """Full-year cases with 118 sub-steps per hour: first 600 hours against the live oracle, incl. iteration counts."""
    zone_o, final_o, states_o, m = pc.run_oracle(name, 600, record_state=True)
    zone, final, states, b = emu_py.run_emu(name, 600, record_state=True)
    off = m.state_offset(0)
    nodes = slice(off + 8, off + 8 + m.n_nodes(0))
    assert np.abs(states[:, nodes] - states_o[:, nodes]).max() <= TOL
    assert np.abs(zone - zone_o).max() <= TOL
    assert b.nomass_iterations == m.nomass_iterations


def test_emulated_heterogeneous_batch():
    """Every lanes-per-wall tier (1..32 lanes), plain and no-mass tiles, several sub-step counts."""
    import oracle_py as O
    from heat_b200 import cases

    d = cases.wall_ensemble(160, seed=5, mixed_fraction=0.5, steps_per_hour=1, max_nodes=250)
    weather, inputs = cases.synthetic_weather(13)
    n_steps = 12
    b = emu_py.EmuBatch(d)
    b.march_series(n_steps, weather, weather_mode=1, inputs=inputs[:n_steps], input_mode=1)
    got = b.get_state()
    assert b.status == 0
    off = 0
    worst = 0.0
    lens = []
    for m in range(d.n_models):
        om = O.OracleModel(d, m)
        so = om.initial_state()
        lens.append(len(so))
        if m % 3 == 0:
            om.march_series(so, n_steps, weather, weather_mode=1, inputs=inputs[:n_steps], input_mode=1)
            mine = got[off:off + len(so)]
            worst = max(worst, np.abs(mine[9:] - so[9:]).max(), abs(mine[0] - so[0]))
        off += len(so)
    assert off == b.state_len
    assert worst <= TOL, worst


def test_emulated_shoebox_buildings():
    import oracle_py as O
    from heat_b200 import cases

    d, gains = cases.shoebox_ensemble(12, seed=9, steps_per_hour=2)
    weather, inputs = cases.synthetic_weather(7)
    n_steps = 6
    b = emu_py.EmuBatch(d)
    b.set_zone_inputs(q_internal=gains)
    b.march_series(n_steps, weather, weather_mode=1, inputs=inputs[:n_steps], input_mode=1)
    got = b.get_state()
    off = 0
    for m in range(d.n_models):
        om = O.OracleModel(d, m)
        om.set_zone_inputs(q_internal=gains[m:m + 1])
        so = om.initial_state()
        om.march_series(so, n_steps, weather, weather_mode=1, inputs=inputs[:n_steps], input_mode=1)
        mine = got[off:off + len(so)]
        # irradiance slots (4..7 of every surface) are inputs; compare temperatures
        assert abs(mine[0] - so[0]) <= TOL
        for s in range(7):
            o = om.state_offset(s)
            n = om.n_nodes(s)
            assert np.abs(mine[o + 8:o + 8 + n] - so[o + 8:o + 8 + n]).max() <= TOL
            assert np.abs(mine[o:o + 4] - so[o:o + 4]).max() <= 1e-5
        off += len(so)


def test_emulated_general_face_mode_buildings():
    pc.check_general_mode_buildings(emu_py.EmuBatch)


def test_emulated_double_glazed_buildings():
    pc.check_double_glazed_buildings(emu_py.EmuBatch)


def test_emulated_epw_file_year(tmp_path):
    pc.check_epw_file_year(emu_py.EmuBatch, str(tmp_path / "santiago.epw"))


def test_emulated_scheduled_zone_gains():
    pc.check_scheduled_zone_gains(emu_py.EmuBatch)


def test_emulated_multi_site_walls():
    pc.check_multi_site_walls(emu_py.EmuBatch)


def test_emulated_two_zone_buildings():
    pc.check_two_zone_buildings(emu_py.EmuBatch)


def test_emulated_two_node_end_chunks():
    pc.check_two_node_end_chunks(emu_py.EmuBatch)


def test_emulated_walls_beyond_254_nodes():
    pc.check_walls_beyond_254_nodes(emu_py.EmuBatch)


# ---- round 2 cases (the same checks the GPU tests run) ----
@pytest.mark.parametrize("kind,n", [("walls", 48), ("mixed", 48), ("buildings", 8)])
def test_emulated_benchmarked_ensembles_prefix(kind, n):
    """The benchmarked generators over 5 days (the GPU test runs the whole year)."""
    pc.check_full_horizon_ensemble(emu_py.EmuBatch, kind, n, hours=120)


def test_emulated_negative_front_solar():
    pc.check_negative_front_solar(emu_py.EmuBatch)


def test_emulated_back_ambient():
    pc.check_back_ambient(emu_py.EmuBatch)


def test_emulated_infiltration_and_ventilation():
    pc.check_infiltration_and_ventilation(emu_py.EmuBatch)


def test_emulated_per_surface_inputs():
    pc.check_per_surface_inputs(emu_py.EmuBatch)


def test_emulated_terrain_classes():
    pc.check_terrain_classes(emu_py.EmuBatch)


def test_emulated_tilted_glazing():
    pc.check_tilted_glazing(emu_py.EmuBatch)


def test_emulated_scheduled_infiltration_ventilation():
    pc.check_scheduled_infiltration_ventilation(emu_py.EmuBatch)


def test_emulated_unstable_wall_fails_like_reference():
    pc.check_unstable_wall_fails_like_reference(emu_py.EmuBatch)


def test_emulated_slow_nomass_convergence():
    pc.check_slow_nomass_convergence(emu_py.EmuBatch)


def test_emulated_long_nomass_chunks():
    pc.check_long_nomass_chunks(emu_py.EmuBatch)

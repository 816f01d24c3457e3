"""Host side of ThermalModel::new (product code in heat_b200/csrc/hb_host.cc) against the oracle:
0 mismatches allowed in tstep_subdivision, n_elements, node counts, chunk ranges and state offsets
(BASELINE.json north_star); masses / conductances / alphas bit-exact."""
import ctypes as C
import os

import numpy as np
import pytest

import oracle_py as O
from heat_b200 import _lib, cases, desc as D
from heat_b200.model import ThermalBatch

HB_DEVICE_NONE = -1


def plan(desc):
    return ThermalBatch(desc, device=HB_DEVICE_NONE)


def compare_model(b, d, m, surface_base):
    om = O.OracleModel(d, m)
    info = b.model_info(m)
    assert info.dt_subdivisions == om.dt_subdivisions
    assert info.dt == om.dt
    assert info.state_len == om.state_len
    for s in range(om.n_surfaces):
        g = surface_base + s
        mine = b.surface_arrays(g)
        c0 = d.construction_offset[d.surfaces[g]["construction"]]
        c1 = d.construction_offset[d.surfaces[g]["construction"] + 1]
        ref = om.surface_arrays(s, c1 - c0)
        assert mine["tstep_subdivision"] == ref["tstep_subdivision"]
        assert mine["n_nodes"] == ref["n_nodes"]
        assert np.array_equal(mine["n_elements"], ref["n_elements"])
        assert np.array_equal(mine["massive_chunks"], ref["massive_chunks"])
        assert np.array_equal(mine["nomass_chunks"], ref["nomass_chunks"])
        assert np.array_equal(mine["mass"], ref["mass"])
        assert np.array_equal(mine["u"], ref["u"])
        assert np.array_equal(mine["front_alpha"], ref["front_alpha"])
        assert np.array_equal(mine["back_alpha"], ref["back_alpha"])
        si = b.surface_info(g)
        assert si.state_offset - info.state_offset == om.state_offset(s)
        assert si.wind_speed_modifier == ref["wind_speed_modifier"]
        assert si.front_emissivity == ref["front_emissivity"] and si.back_emissivity == ref["back_emissivity"]
    return om.n_surfaces


@pytest.mark.parametrize("name", ["massive_full", "mixed_full", "nomass_full", "trombe_wall_full", "tilted",
                                  "horizontal", "massive_no_ir_no_solar"])
@pytest.mark.parametrize("n", [1, 2, 4, 6, 20, 60])
def test_validation_cases(name, n):
    d = cases.validation_case(name, n)
    compare_model(plan(d), d, 0, 0)


def test_reference_sizes_table():
    """BASELINE.md §2 table (rule at discretization.rs:410-544, model.rs:236-240,326-331)."""
    expect = {("massive_full", 20): (2, 90.0, [12], 13), ("nomass_full", 20): (4, 45.0, [0], 2),
              ("mixed_full", 20): (4, 45.0, [0, 18, 0], 21), ("massive_full", 1): (10, 360.0, [5], 6),
              ("mixed_full", 1): (118, 3600.0 / 118, [0, 22, 0], 25)}
    for (name, n), (S, dt, el, N) in expect.items():
        b = plan(cases.validation_case(name, n))
        info = b.model_info(0)
        a = b.surface_arrays(0)
        assert info.dt_subdivisions == S and abs(info.dt - dt) < 1e-12
        assert a["n_elements"].tolist() == el and a["n_nodes"] == N


def test_random_wall_ensemble_zero_mismatches():
    d = cases.wall_ensemble(1500, seed=123, mixed_fraction=0.4)
    b = plan(d)
    for m in range(d.n_models):
        compare_model(b, d, m, m)
    assert b.updates_per_step == sum(b.model_info(m).nodes * b.model_info(m).dt_subdivisions for m in range(d.n_models))


def test_rk4_number_matches_numpy_statement():
    """hb_surface_rk4_number (Sturm bisection in hb_host.cc) against the dense-eigenvalue statement in cases.py, on walls
    with massive chunks at the ends, between no-mass layers, and next to a gas cavity."""
    descs = [cases.wall_ensemble(300, seed=321, mixed_fraction=0.4), cases.validation_case("trombe_wall_full"),
             cases.validation_case("mixed_full"), cases.unstable_wall()]
    checked = 0
    for d in descs:
        b = plan(d)
        for g in range(b.n_surfaces):
            a = b.surface_arrays(g)
            dt = b.model_info(int(np.searchsorted(d.model_surface_offset, g, side="right")) - 1).dt
            for film in (0.0, 30.0):
                want = cases.rk4_number_np(a["mass"], a["u"], a["massive_chunks"], dt, film)
                got = b.surface_rk4_number(g, film)
                assert abs(got - want) <= 1e-6 * max(want, 1e-12), (g, film, got, want)
            checked += 1
    assert checked > 300
    assert plan(cases.unstable_wall()).surface_rk4_number(0) > cases.RK4_STABILITY_LIMIT


def test_wall_ensemble_redraws_unstable_walls():
    """The generator's third re-draw test: a planner that reports wall 5 unstable once gets wall 5 re-drawn, the others
    stay as drawn; the benchmark ensemble of one GPU (seed 42) has no such wall, so it is what round 1 measured."""
    raw = cases._wall_ensemble_draw(40, 11, 0.0, 1, 4)
    calls = []

    def planner(desc):
        calls.append(1)
        rk4 = np.full(desc.n_models, 1.0)
        if len(calls) == 1:
            rk4[5] = 3.0
        return np.full(desc.n_models, 20), np.ones(desc.n_models, np.int64), rk4

    d = cases.wall_ensemble(40, seed=11, planner=planner)
    assert len(calls) == 2
    lo, hi = raw.construction_offset[5], raw.construction_offset[6]
    same = raw.layers["thickness"] == d.layers["thickness"]
    assert not same[lo:hi].any() and same[:lo].all() and same[hi:].all()
    b = plan(cases.wall_ensemble(4000, seed=42))
    assert max(b.surface_rk4_number(g) for g in range(4000)) < cases.RK4_STABILITY_LIMIT


def test_shoebox_ensemble_zero_mismatches():
    d, _ = cases.shoebox_ensemble(60, seed=77, steps_per_hour=2)
    b = plan(d)
    for m in range(d.n_models):
        assert compare_model(b, d, m, 7 * m) == 7


def test_glazing_stack_alphas():
    """glass | argon | glass fenestration: ISO 9050 absorptances distributed over the panes' nodes."""
    g1 = cases.glass(0.006, 0.7, 0.84, 0.1)
    g2 = cases.glass(0.004, 0.8, 0.84, 0.05)
    s = D.surface(0, 2.0, 6.0, (0, -1, 0), 1.0, front=(D.BOUNDARY_OUTDOOR, 0), back=(D.BOUNDARY_SPACE, 0),
                  is_fenestration=True)
    d = D.single_model([[g1, D.gas_layer(0.012, D.GAS_ARGON), g2]], [30.0], [s], 4)
    compare_model(plan(d), d, 0, 0)


def test_errors_mirror_reference():
    lay = cases.concrete(0.2)
    gas = D.gas_layer(0.05)
    s = D.surface(0, 1.0, 4.0, (0, -1, 0), 1.0)
    for layers, msg in (([gas, lay], "first layer"), ([lay, gas], "last layer"), ([lay, gas, gas, lay], "two gases")):
        with pytest.raises(Exception) as e:
            plan(D.single_model([layers], [10.0], [s], 4))
        assert msg in str(e.value)
    ground = D.surface(0, 1.0, 4.0, (0, -1, 0), 1.0, front=(D.BOUNDARY_GROUND, 0))
    with pytest.raises(Exception) as e:
        plan(D.single_model([[lay]], [10.0], [ground], 4))
    assert e.value.status == 4  # HB_ERR_UNSUPPORTED (reference: unimplemented!(), model.rs:92)
    bad_zone = D.surface(0, 1.0, 4.0, (0, -1, 0), 1.0, back=(D.BOUNDARY_SPACE, 3))
    with pytest.raises(Exception) as e:
        plan(D.single_model([[lay]], [10.0], [bad_zone], 4))
    assert "Space not found" in str(e.value)


def test_planning_queries_reject_bad_arguments():
    """hb_surface_rk4_number / hb_get_surface_info: index and argument errors are HB_ERR_INVALID with a message, and a
    chunk beyond the 32-node limit of the no-mass solver is refused at hb_create, not mis-solved."""
    from heat_b200 import HeatB200Error

    b = plan(cases.validation_case("massive_full"))
    assert b.surface_rk4_number(0) > 0.0
    for call in (lambda: b.surface_rk4_number(5), lambda: b.surface_rk4_number(0, film=-1.0), lambda: b.surface_info(-1)):
        with pytest.raises(HeatB200Error) as e:
            call()
        assert e.value.status == 1 and str(e.value)
    thin = [("polyurethane", 0.004)] * 40   # one no-mass chunk of 41 nodes
    with pytest.raises(HeatB200Error, match="no-mass chunk longer"):
        plan(cases.single_zone_test_building(thin, steps_per_hour=4))


def test_no_cpu_fallback():
    """A planning handle cannot march; there is no CPU path behind the C ABI."""
    from heat_b200 import HeatB200Error
    b = plan(cases.validation_case("massive_full"))
    with pytest.raises(HeatB200Error) as e:
        b.march_series(1, np.array([[20.0, 0.0, 1.0]]))
    assert e.value.status == 2  # HB_ERR_CUDA
    with pytest.raises(HeatB200Error):
        b.get_state()


# ---- EPW ingestion (hb_read_epw) ------------------------------------------------------------------------------
_EPW_HEADER = ["LOCATION,TESTVILLE,-,XXX,Synthetic,000000,-33.38,-70.78,-4.0,476.0", "DESIGN CONDITIONS,0",
               "TYPICAL/EXTREME PERIODS,0", "GROUND TEMPERATURES,0", "HOLIDAYS/DAYLIGHT SAVINGS,No,0,0,0",
               'COMMENTS 1,"a, quoted, comment"', "COMMENTS 2,none", "DATA PERIODS,1,1,Data,Sunday, 1/ 1,12/31"]


def _epw_record(hour, t, ir, ghi, wdir, wspd):
    f = ["1987", "1", "1", str(hour), "60", "C9C9*0?9", f"{t}", "9.6", "63", "95600", "0", "1415", f"{ir}", f"{ghi}",
         "0", "0", "0", "0", "0", "0", f"{wdir}", f"{wspd}", "0", "0", "9.9", "77777", "9", "999999999", "0", "0.2680",
         "0", "88", "0.000", "0.0", "0"]
    return ",".join(f)


def test_read_epw_synthetic_file(tmp_path):
    from heat_b200.model import read_epw

    rng = np.random.default_rng(3)
    vals = np.round(rng.uniform(-5, 400, (30, 5)), 1)
    p = tmp_path / "synthetic.epw"
    body = [_epw_record(i % 24 + 1, *vals[i]) for i in range(30)]
    p.write_text("\r\n".join(_EPW_HEADER + body) + "\r\n")      # CRLF like the files in the wild, trailing newline
    weather, inputs = read_epw(str(p))
    assert weather.shape == (31, 3) and inputs.shape == (30, 4)
    assert np.array_equal(weather[:30], vals[:, [0, 3, 4]])       # dry bulb, wind direction, wind speed
    assert np.array_equal(weather[30], weather[0])                # periodic closure for HB_WEATHER_LINEAR
    assert np.array_equal(inputs[:, 0], vals[:, 2]) and np.array_equal(inputs[:, 2], vals[:, 1])
    assert not inputs[:, [1, 3]].any()
    w2, _ = read_epw(str(p), periodic=False)
    assert w2.shape == (30, 3)


def test_read_epw_errors(tmp_path):
    from heat_b200.model import read_epw

    with pytest.raises(Exception) as e:
        read_epw(str(tmp_path / "missing.epw"))
    assert "cannot open" in str(e.value)
    p = tmp_path / "short.epw"
    p.write_text("\n".join(_EPW_HEADER + ["1987,1,1,1,60,flags,16.7,9.6"]) + "\n")
    with pytest.raises(Exception) as e:
        read_epw(str(p))
    assert "fields" in str(e.value)
    p = tmp_path / "nan.epw"
    p.write_text("\n".join(_EPW_HEADER + [_epw_record(1, "abc", 1, 2, 3, 4)]) + "\n")
    with pytest.raises(Exception) as e:
        read_epw(str(p))
    assert "not a number" in str(e.value)


@pytest.mark.skipif(not os.path.exists("/root/reference/tests/epw/CHL_Santiago.855740_IWEC.epw"),
                    reason="reference tree only exists in the build container")
def test_read_epw_reference_file_matches_fixture():
    """The reference's own EPW through hb_read_epw == tests/golden/epw_santiago.npz (parsed by make_fixtures.py)."""
    from heat_b200.model import read_epw

    weather, inputs = read_epw("/root/reference/tests/epw/CHL_Santiago.855740_IWEC.epw")
    h = np.load(os.path.join(os.path.dirname(__file__), "golden", "epw_santiago.npz"))["hourly"]
    assert weather.shape == (8761, 3)
    assert np.array_equal(weather[:8760], h[:, [0, 3, 4]])
    assert np.array_equal(inputs[:, 0], h[:, 2]) and np.array_equal(inputs[:, 2], h[:, 1])


def test_every_planned_warp_tile_has_a_kernel():
    """The geometries the planner hands out against the instantiations of heat_b200/csrc/hb_wtile_tier.cu (a miss is a
    launch error that only a GPU run would show): all-massive walls NL in {8, 12, 16, 24, 32} (24: G <= 4, 32: G = 1),
    walls with no-mass ends and buildings NL in {8, 12, 16}, buildings G <= 8; G·W = 32 always."""
    import re
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "emu"))
    import emu_py

    src = open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "heat_b200", "csrc", "hb_wtile_tier.cu")).read()
    compiled = {int(m) for m in re.findall(r"case (\d+): return launch_wt_mode<G, \d+>", src)}
    TILE_END, TILE_SIMPLE, TILE_WARP = 8, 64, 128
    batches = [cases.wall_ensemble(6000, seed=42), cases.wall_ensemble(6000, seed=44, mixed_fraction=0.5),
               cases.wall_ensemble(300, seed=5, mixed_fraction=0.5, max_nodes=500),
               cases.shoebox_ensemble(400, seed=43)[0], cases.shoebox_ensemble(50, seed=43, glazing=cases.double_glazing())[0]]
    seen = set()
    for d in batches:
        for G, NL, W, flags, S in emu_py.EmuBatch(d).tile_table():
            if not flags & TILE_WARP:
                continue
            assert G * W == 32 and NL in compiled, (G, NL, W)
            lean = (flags & TILE_SIMPLE) and not (flags & TILE_END)
            if NL > 16:
                assert lean and G <= 4 and (NL < 32 or G == 1), (G, NL, flags)
            if not flags & TILE_SIMPLE:
                assert G <= 8, (G, NL, flags)
            seen.add((int(G), int(NL), bool(lean)))
    assert {nl for _, nl, lean in seen if lean} >= {8, 12, 16, 24, 32}
    assert {nl for _, nl, lean in seen if not lean} >= {8, 12, 16}

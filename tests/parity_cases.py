"""Parity cases shared by the oracle tests, the GPU tests, the golden-fixture generator and smoke().

Every case is (model description, weather table, per-step inputs, modes, initial state tweak); `run_oracle` drives the
CPU oracle and `run_gpu` the CUDA path through the C ABI with the SAME arrays.
"""
from __future__ import annotations

import math
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

from heat_b200 import cases, desc as D  # noqa: E402

GOLDEN_DIR = os.path.join(HERE, "golden")

EPLUS_CASES = [
    "massive_full", "massive_no_ir_no_solar", "massive_no_ir_yes_solar", "massive_yes_ir_no_solar",
    "mixed_full", "mixed_no_ir_no_solar", "mixed_no_ir_yes_solar", "mixed_yes_ir_no_solar",
    "nomass_full", "nomass_no_ir_no_solar", "nomass_no_ir_yes_solar", "nomass_yes_ir_no_solar",
    "tilted", "horizontal",
]
EPW_CASES = ["massive_full", "mixed_full", "nomass_full", "trombe_wall_full"]
THEORETICAL = ["very_simple_march", "march_with_window", "march_with_window_and_luminaire",
               "march_with_window_and_heater", "march_with_window_heater_and_infiltration"]

GOLDEN_CASES = ([f"eplus/{c}" for c in EPLUS_CASES] + ["eplus/trombe_wall_full"] + [f"epw/{c}" for c in EPW_CASES]
                + [f"theory/{c}" for c in THEORETICAL])

_cache = {}


def _npz(name):
    if name not in _cache:
        _cache[name] = np.load(os.path.join(GOLDEN_DIR, name))
    return _cache[name]


def golden(name):
    g = _npz("oracle_golden.npz")
    return g[name + "/zone"], g[name + "/final"]


class Case:
    def __init__(self, desc, n_steps, weather, weather_mode, inputs=None, input_mode=0, ir_mode=0, zone_t0=None,
                 zone_inputs=None):
        self.desc, self.n_steps = desc, n_steps
        self.weather, self.weather_mode = np.ascontiguousarray(weather, np.float64), weather_mode
        self.inputs = None if inputs is None else np.ascontiguousarray(inputs, np.float64)
        self.input_mode, self.ir_mode = input_mode, ir_mode
        self.zone_t0 = zone_t0
        self.zone_inputs = zone_inputs or {}


def make_case(name: str, max_steps: int | None = None) -> Case:
    group, _, case = name.partition("/")
    if group == "eplus":
        # tests/validate_wall_heat_transfer.rs:615-711 (march_model): n = 20/h, one CSV row per main step
        src = "massive_full" if case == "trombe_wall_full" else case  # Trombe: its own CSV is a missing blob
        cols = _npz("eplus_inputs.npz")[src]
        n = len(cols) if max_steps is None else min(max_steps, len(cols))
        weather = np.stack([cols[:n, 3], cols[:n, 1], cols[:n, 0]], axis=1)        # t_out, wind dir, wind speed
        inputs = np.stack([cols[:n, 2], np.zeros(n), cols[:n, 4], np.zeros(n)], 1)  # solar, -, LW gain [W], -
        return Case(cases.validation_case(case, 20), n, weather, 0, inputs, 1, 1, zone_t0=cols[0, 5])
    if group == "epw":
        # 1-year hourly run, n = 1/h, linear interpolation at sub-step end times (builder-defined)
        h = _npz("epw_santiago.npz")["hourly"]
        n = len(h) if max_steps is None else min(max_steps, len(h))
        rows = np.concatenate([h, h[:1]], axis=0)
        weather = np.stack([rows[:n + 1, 0], rows[:n + 1, 3], rows[:n + 1, 4]], axis=1)
        inputs = np.stack([h[:n, 2], np.zeros(n), h[:n, 1], np.zeros(n)], axis=1)    # GHI → front solar, IR → front IR
        return Case(cases.validation_case(case, 1), n, weather, 1, inputs, 1, 0)
    if group == "theory":
        # tests/validate_wall_heat_transfer.rs:95-613: polyurethane wall, h overrides 10/10, T_out = 30 constant
        opts = {
            "very_simple_march": dict(n=60, steps=1000),
            "march_with_window": dict(n=6, steps=80, window=True),
            "march_with_window_and_luminaire": dict(n=20, steps=800, q=100.0),
            "march_with_window_and_heater": dict(n=20, steps=800, q=100.0),
            "march_with_window_heater_and_infiltration": dict(n=20, steps=22, q=10.0, inf=0.1),
        }[case]
        d = cases.single_zone_test_building([("polyurethane", 0.02)], zone_volume=40.0, surface_height=2.0,
                                            surface_width=2.0, window_height=1.0 if opts.get("window") else 0.0,
                                            window_width=1.0 if opts.get("window") else 0.0, emissivity=0.0,
                                            steps_per_hour=opts["n"], front_hs=10.0, back_hs=10.0)
        n = opts["steps"] if max_steps is None else min(max_steps, opts["steps"])
        weather = np.tile(np.array([[30.0, 0.0, 0.0]]), (n, 1))
        zin = {}
        if "q" in opts:
            zin["q_internal"] = np.array([opts["q"]])
        if "inf" in opts:
            zin["infiltration_volume"] = np.array([opts["inf"]])
            zin["infiltration_temperature"] = np.array([30.0])
        return Case(d, n, weather, 0, zone_inputs=zin)
    raise KeyError(name)


def run_oracle(name: str, max_steps: int | None = None, record_state: bool = False):
    import oracle_py as O
    c = make_case(name, max_steps)
    m = O.OracleModel(c.desc)
    state = m.initial_state()
    if c.zone_t0 is not None:
        state[0] = c.zone_t0
    if c.zone_inputs:
        m.set_zone_inputs(**c.zone_inputs)
    zone, states = m.march_series(state, c.n_steps, c.weather, weather_mode=c.weather_mode, inputs=c.inputs,
                                  input_mode=c.input_mode, ir_mode=c.ir_mode, record_every=1,
                                  record_state=record_state)
    if record_state:
        return zone, state, states, m
    return zone, state


def run_gpu(name: str, max_steps: int | None = None, record_state: bool = False):
    from heat_b200.model import ThermalBatch
    c = make_case(name, max_steps)
    b = ThermalBatch(c.desc)
    if c.zone_t0 is not None:
        s = b.get_state()
        s[0] = c.zone_t0
        b.set_state(s)
    if c.zone_inputs:
        b.set_zone_inputs(**c.zone_inputs)
    zone, states = b.march_series(c.n_steps, c.weather, weather_mode=c.weather_mode, inputs=c.inputs,
                                  input_mode=c.input_mode, ir_mode=c.ir_mode, record_every=1,
                                  record_state=record_state)
    final = b.get_state()
    if record_state:
        return zone, final, states, b
    return zone, final


def _general_mode_buildings(n):
    """Buildings that need the general face mode: per-model weather sites and a film-coefficient override on the roof
    of every third building (windows stay single panes)."""
    from heat_b200 import cases

    d, gains = cases.shoebox_ensemble(n, seed=11, steps_per_hour=2)
    surf = d.surfaces.reshape(n, 7)
    surf["back_hs_override"][::3, 4] = 6.5
    w0, i0 = cases.synthetic_weather(7)
    w1 = w0.copy()
    w1[:, 0] += 4.0
    w1[:, 1] = np.mod(w1[:, 1] + 90.0, 360.0)
    w1[:, 2] *= 0.5
    i1 = i0.copy()
    i1[:, 0] *= 0.7
    weather = np.ascontiguousarray(np.stack([w0, w1]))
    inputs = np.ascontiguousarray(np.stack([i0[:6], i1[:6]]))
    site = (np.arange(n) % 2).astype(np.int32)
    return d, gains, weather, inputs, site


def check_general_mode_buildings(batch_cls, n=8):
    import oracle_py as O

    d, gains, weather, inputs, site = _general_mode_buildings(n)
    n_steps = 6
    b = batch_cls(d)
    b.set_zone_inputs(q_internal=gains)
    b.march_series(n_steps, weather, weather_mode=1, inputs=inputs, input_mode=1, model_site=site)
    got = b.get_state()
    off = 0
    for m in range(d.n_models):
        om = O.OracleModel(d, m)
        om.set_zone_inputs(q_internal=gains[m:m + 1])
        so = om.initial_state()
        om.march_series(so, n_steps, weather[site[m]], weather_mode=1, inputs=inputs[site[m]], input_mode=1)
        mine = got[off:off + len(so)]
        assert abs(mine[0] - so[0]) <= 1e-6
        for s in range(7):
            o = om.state_offset(s)
            nn = om.n_nodes(s)
            assert np.abs(mine[o + 8:o + 8 + nn] - so[o + 8:o + 8 + nn]).max() <= 1e-6
        off += len(so)


def write_epw_from_fixture(path, hours=8760):
    """EPW text file (8 header lines + hourly records, CRLF) holding the Santiago fixture's five columns — the reference's
    own file does not travel to the GPU box.  `repr` of a double round-trips through strtod exactly."""
    h = _npz("epw_santiago.npz")["hourly"][:hours]
    head = ["LOCATION,SANTIAGO,-,CHL,IWEC Data,855740,-33.38,-70.78,-4.0,476.0", "DESIGN CONDITIONS,0",
            "TYPICAL/EXTREME PERIODS,0", "GROUND TEMPERATURES,0", "HOLIDAYS/DAYLIGHT SAVINGS,No,0,0,0", "COMMENTS 1,fixture",
            "COMMENTS 2,fixture", "DATA PERIODS,1,1,Data,Sunday, 1/ 1,12/31"]
    with open(path, "w", newline="") as f:
        f.write("\r\n".join(head) + "\r\n")
        for i, (t, ir, ghi, wd, ws) in enumerate(h):
            rec = ["1987", str(1 + i // 730), str(1 + (i // 24) % 30), str(1 + i % 24), "60", "?9?9", repr(float(t)), "9.6",
                   "63", "95600", "0", "1415", repr(float(ir)), repr(float(ghi)), "0", "0", "0", "0", "0", "0",
                   repr(float(wd)), repr(float(ws)), "0", "0", "9.9", "77777", "9", "999999999", "0", "0.2680", "0", "88",
                   "0.000", "0.0", "0"]
            f.write(",".join(rec) + "\r\n")


def check_epw_file_year(batch_cls, path, case="massive_full"):
    """EPW file → hb_read_epw → one series call over the whole year (n = 1/h, linear weather between records) against
    the committed oracle output of the same case (north_star: full-year hourly EPW run within 1e-6 °C)."""
    from heat_b200.model import read_epw

    write_epw_from_fixture(path)
    weather, inputs = read_epw(path)
    c = make_case(f"epw/{case}")
    assert np.array_equal(weather, c.weather) and np.array_equal(inputs, c.inputs)
    b = batch_cls(cases.validation_case(case, 1))
    zone, _ = b.march_series(len(inputs), weather, weather_mode=1, inputs=inputs, input_mode=1, ir_mode=0, record_every=1)
    zone_g, final_g = golden(f"epw/{case}")
    assert np.abs(zone[:, 0] - np.asarray(zone_g).reshape(len(inputs), -1)[:, 0]).max() <= 1e-6
    assert np.abs(b.get_state() - final_g).max() <= 1e-5


def check_double_glazed_buildings(batch_cls, n=6, n_steps=6):
    """Shoeboxes whose window is pane | gas | pane: ISO 9050 absorptances per pane (glazing.rs:180-286), the cavity's
    temperature-dependent U (cavity.rs:59-69, gas.rs:82-315) and the no-mass solve of a 4-node window, next to the zone
    coupling.  One ensemble per gas / gap so that both Nusselt regimes of a vertical cavity are met."""
    import oracle_py as O

    weather, inputs = cases.synthetic_weather(n_steps + 1)
    worst = 0.0
    for gas, gap in ((D.GAS_ARGON, 0.012), (D.GAS_AIR, 0.03)):
        d, gains = cases.shoebox_ensemble(n, seed=17, steps_per_hour=2, glazing=cases.double_glazing(gas, gap))
        b = batch_cls(d)
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
            assert om.n_nodes(6) == 4, om.n_nodes(6)
            assert abs(mine[0] - so[0]) <= 1e-6, (gas, m, mine[0], so[0])
            for s in range(7):
                o = om.state_offset(s)
                nn = om.n_nodes(s)
                err = np.abs(mine[o + 8:o + 8 + nn] - so[o + 8:o + 8 + nn]).max()
                worst = max(worst, err)
                assert err <= 1e-6, (gas, m, s, err)
            off += len(so)
        assert off == len(got)
    return worst


def check_scheduled_zone_gains(batch_cls, n=10, n_steps=8):
    """HVAC / luminaire power as a per-main-step schedule (model.rs:502-516: the reference reads the state elements the
    caller rewrites before each march): one series call with `zone_gain` against the oracle marched step by step with
    the gains reset in between.  Shoeboxes (face kernel, zone scan), two-zone models (general zone sums) and
    double-glazed shoeboxes (general kernel)."""
    import oracle_py as O

    weather, inputs = cases.synthetic_weather(n_steps + 1)
    builders = (lambda: cases.shoebox_ensemble(n, seed=23, steps_per_hour=2),
                lambda: cases.two_zone_ensemble(max(2, n // 2), seed=29, steps_per_hour=2),
                lambda: cases.shoebox_ensemble(max(2, n // 2), seed=31, steps_per_hour=2, glazing=cases.double_glazing()))
    for build in builders:
        d, gains = build()
        nz = len(gains)
        rng = cases.SplitMix64(1234)
        sched = rng.uniform(n_steps * nz, -800.0, 1500.0).reshape(n_steps, nz)   # heating and cooling
        sched[1] = 0.0
        b = batch_cls(d)
        b.set_zone_inputs(q_internal=gains)   # overridden by the schedule while it runs
        b.march_series(n_steps, weather, weather_mode=1, inputs=inputs[:n_steps], input_mode=1, zone_gain=sched)
        got = b.get_state()
        off = 0
        zoff = d.model_zone_offset
        for m in range(d.n_models):
            z0, z1 = int(zoff[m]), int(zoff[m + 1])
            om = O.OracleModel(d, m)
            so = om.initial_state()
            for t in range(n_steps):
                om.set_zone_inputs(q_internal=sched[t, z0:z1])
                om.march_series(so, 1, weather[t:t + 2], weather_mode=1, inputs=inputs[t:t + 1], input_mode=1)
            mine = got[off:off + len(so)]
            nzm = z1 - z0
            assert np.abs(mine[:nzm] - so[:nzm]).max() <= 1e-6, (m, mine[:nzm], so[:nzm])
            for s in range(om.n_surfaces):
                o = om.state_offset(s)
                nn = om.n_nodes(s)
                assert np.abs(mine[o + 8:o + 8 + nn] - so[o + 8:o + 8 + nn]).max() <= 1e-6, (m, s)
            off += len(so)
        # the schedule must matter: the same run with constant gains ends elsewhere
        b2 = batch_cls(d)
        b2.set_zone_inputs(q_internal=gains)
        b2.march_series(n_steps, weather, weather_mode=1, inputs=inputs[:n_steps], input_mode=1)
        assert np.abs(b2.get_state() - got).max() > 1e-3


def check_multi_site_walls(batch_cls, n=48):
    """Walls (massive and with no-mass layers) under per-model weather sites: general face mode / general kernel."""
    import oracle_py as O

    d = cases.wall_ensemble(n, seed=21, mixed_fraction=0.5, steps_per_hour=1, max_nodes=120)
    w0, i0 = cases.synthetic_weather(9)
    w1 = w0.copy()
    w1[:, 0] -= 6.0
    w1[:, 1] = np.mod(w1[:, 1] + 135.0, 360.0)
    weather = np.ascontiguousarray(np.stack([w0, w1]))
    n_steps = 8
    inputs = np.ascontiguousarray(np.stack([i0[:n_steps], 0.5 * i0[:n_steps]]))
    site = ((np.arange(n) // 3) % 2).astype(np.int32)
    b = batch_cls(d)
    b.march_series(n_steps, weather, weather_mode=1, inputs=inputs, input_mode=1, model_site=site)
    got = b.get_state()
    off = 0
    for m in range(d.n_models):
        om = O.OracleModel(d, m)
        so = om.initial_state()
        om.march_series(so, n_steps, weather[site[m]], weather_mode=1, inputs=inputs[site[m]], input_mode=1)
        mine = got[off:off + len(so)]
        assert max(np.abs(mine[9:] - so[9:]).max(), abs(mine[0] - so[0])) <= 1e-6, m
        off += len(so)


def check_two_zone_buildings(batch_cls, n=6, n_steps=6):
    """Two zones per model coupled by a partition (front face -> zone 0, back face -> zone 1) + a pane on zone 1:
    the zone sums take the general (non-segmented) path, with more than 8 surfaces per model."""
    import oracle_py as O

    d, gains = cases.two_zone_ensemble(n, seed=44, steps_per_hour=2)
    weather, inputs = cases.synthetic_weather(n_steps + 1)
    b = batch_cls(d)
    b.set_zone_inputs(q_internal=gains)
    b.march_series(n_steps, weather, weather_mode=1, inputs=inputs[:n_steps], input_mode=1)
    got = b.get_state()
    off = 0
    for m in range(d.n_models):
        om = O.OracleModel(d, m)
        om.set_zone_inputs(q_internal=gains[2 * m:2 * m + 2])
        so = om.initial_state()
        om.march_series(so, n_steps, weather, weather_mode=1, inputs=inputs[:n_steps], input_mode=1)
        mine = got[off:off + len(so)]
        assert np.abs(mine[0:2] - so[0:2]).max() <= 1e-6, (m, mine[0:2], so[0:2])
        for s in range(om.n_surfaces):
            o = om.state_offset(s)
            nn = om.n_nodes(s)
            assert np.abs(mine[o + 8:o + 8 + nn] - so[o + 8:o + 8 + nn]).max() <= 1e-6, (m, s)
        off += len(so)


def check_two_node_end_chunks(batch_cls, n_steps=48):
    """Two thin no-mass layers at one or both ends of a wall: chunks (0,2) / (n-2,n), solved by the face lanes
    (solve_end_pair) — such walls must not need the general kernel."""
    import oracle_py as O

    poly2 = [("polyurethane", 0.01), ("polyurethane", 0.012)]
    stacks = [poly2 + [("concrete", 0.2)] + poly2, [("concrete", 0.15)] + poly2, poly2 + [("concrete", 0.15)],
              [("polyurethane", 0.01), ("concrete", 0.15)] + poly2]
    weather, inputs = cases.synthetic_weather(n_steps + 1)
    for cons in stacks:
        d = cases.single_zone_test_building(cons, zone_volume=600.0, surface_height=3.0, surface_width=20.0,
                                            emissivity=0.9, solar_absorbtance=0.7, steps_per_hour=4,
                                            back_emissivity_override=0.0)
        b = batch_cls(d)
        b.march_series(n_steps, weather, weather_mode=1, inputs=inputs[:n_steps], input_mode=1)
        got = b.get_state()
        om = O.OracleModel(d, 0)
        so = om.initial_state()
        om.march_series(so, n_steps, weather, weather_mode=1, inputs=inputs[:n_steps], input_mode=1)
        assert abs(got[0] - so[0]) <= 1e-6, cons
        assert np.abs(got[9:] - so[9:]).max() <= 1e-6, cons
        iters = b.nomass_iterations if hasattr(b, "nomass_iterations") else b.reduce().nomass_iterations
        assert iters == om.nomass_iterations, cons


def check_long_nomass_chunks(batch_cls, n_steps=24):
    """No-mass chunks of 9 … 31 nodes (surface.rs:790-898 solves any length): inside a wall between massive layers, at
    an end, and a wall that is one long no-mass chunk — the general kernel's tridiagonal elimination."""
    import oracle_py as O

    thin = lambda k: [("polyurethane", 0.004 + 0.0005 * (i % 3)) for i in range(k)]
    stacks = [[("concrete", 0.12)] + thin(10) + [("concrete", 0.1)],          # chunk of 9 inner nodes
              thin(14) + [("concrete", 0.15)],                                  # 14-node chunk at the front
              [("concrete", 0.15)] + thin(22),                                  # 22-node chunk at the back
              thin(30)]                                                         # the whole wall: 31 no-mass nodes
    weather, inputs = cases.synthetic_weather(n_steps + 1)
    longest = 0
    for cons in stacks:
        d = cases.single_zone_test_building(cons, zone_volume=600.0, surface_height=3.0, surface_width=20.0,
                                            emissivity=0.9, solar_absorbtance=0.7, steps_per_hour=4,
                                            back_emissivity_override=0.0)
        om = O.OracleModel(d, 0)
        chunks = om.surface_arrays(0, len(cons))["nomass_chunks"]
        longest = max(longest, int((chunks[:, 1] - chunks[:, 0]).max()))
        b = batch_cls(d)
        b.march_series(n_steps, weather, weather_mode=1, inputs=inputs[:n_steps], input_mode=1)
        got = b.get_state()
        so = om.initial_state()
        om.march_series(so, n_steps, weather, weather_mode=1, inputs=inputs[:n_steps], input_mode=1)
        assert abs(got[0] - so[0]) <= 1e-6, cons
        assert np.abs(got[9:] - so[9:]).max() <= 1e-6, (len(cons), np.abs(got[9:] - so[9:]).max())
        iters = b.nomass_iterations if hasattr(b, "nomass_iterations") else b.reduce().nomass_iterations
        assert iters == om.nomass_iterations, cons
    assert longest >= 31, longest


def _dense_layer(th, eps=0.9, alpha=0.7):
    """Low-diffusivity solid (k = 0.05, rho = 2500, cp = 2000): ~1200 nodes per metre under a 30 s sub-step."""
    return D.solid_layer(th, conductivity=0.05, density=2500.0, specific_heat=2000.0, front_emissivity=eps,
                         back_emissivity=eps, front_solar_abs=alpha, back_solar_abs=alpha)


def check_walls_beyond_254_nodes(batch_cls, n_steps=3):
    """Walls of more than 254 nodes (32 lanes x 8 slots): the 16- and 32-slot tiers of the face kernel.  Hourly steps;
    a thin polyurethane skin (or the building's pane) imposes the 30.5 s sub-step under which the dense layers
    discretize finely (discretization.rs:410-544).  (a) single walls with their own zone, (b) a shoebox with a pane."""
    import oracle_py as O

    weather, inputs = cases.synthetic_weather(n_steps + 1)
    normal = (0.0, -1.0, 0.0)
    seen = []
    for th in (0.40, 0.70):
        layers = [cases.polyurethane(0.02, 0.9, 0.7), _dense_layer(th), cases.polyurethane(0.02, 0.9, 0.7)]
        surf = [D.surface(0, 60.0, 46.0, normal, 1.5, front=(D.BOUNDARY_OUTDOOR, 0), back=(D.BOUNDARY_SPACE, 0))]
        d = D.single_model([layers], [600.0], surf, 1)
        b = batch_cls(d)
        b.march_series(n_steps, weather, weather_mode=1, inputs=inputs[:n_steps], input_mode=1)
        got = b.get_state()
        om = O.OracleModel(d, 0)
        so = om.initial_state()
        om.march_series(so, n_steps, weather, weather_mode=1, inputs=inputs[:n_steps], input_mode=1)
        nn = om.n_nodes(0)
        seen.append(nn)
        assert len(got) == len(so)
        assert np.abs(got - so).max() <= 1e-6, (th, nn, np.abs(got - so).max())
    assert 254 < seen[0] <= 510 < seen[1] <= 1022, seen

    # (b) two shoeboxes; the first one's opaque construction (two layers) becomes a polyurethane skin on a dense layer
    d, gains = cases.shoebox_ensemble(2, seed=43)
    lo = int(d.construction_offset[0])
    assert int(d.construction_offset[1]) - lo == 2
    d.layers[lo] = cases.polyurethane(0.02, 0.8, 0.5)
    d.layers[lo + 1] = _dense_layer(0.36, 0.8, 0.5)
    b = batch_cls(d)
    b.set_zone_inputs(q_internal=gains)
    b.march_series(n_steps, weather, weather_mode=1, inputs=inputs[:n_steps], input_mode=1)
    got = b.get_state()
    off = 0
    big = 0
    for m in range(d.n_models):
        om = O.OracleModel(d, m)
        om.set_zone_inputs(q_internal=gains[m:m + 1])
        so = om.initial_state()
        om.march_series(so, n_steps, weather, weather_mode=1, inputs=inputs[:n_steps], input_mode=1)
        big = max(big, max(om.n_nodes(s) for s in range(om.n_surfaces)))
        assert np.abs(got[off:off + len(so)] - so).max() <= 1e-6, (m, np.abs(got[off:off + len(so)] - so).max())
        off += len(so)
    assert big > 254, big


# ------------------------------------------------------------------------------------------------------------------
# Round 2: the benchmarked ensembles over the full horizon, and the device branches round 1 never reached
# ------------------------------------------------------------------------------------------------------------------
def ensemble_of(kind: str, n: int):
    """The generators bench.py runs (configs[2], [3], [4]) with their seeds, at `n` units."""
    if kind == "walls":
        return cases.wall_ensemble(n, seed=42), None
    if kind == "mixed":
        return cases.wall_ensemble(n, seed=44, mixed_fraction=0.5), None
    if kind == "buildings":
        return cases.shoebox_ensemble(n, seed=43)
    raise KeyError(kind)


def check_full_horizon_ensemble(batch_cls, kind: str, n: int, hours: int = 8760, threads: int | None = None,
                                tol: float = 1e-6):
    """A seeded sample of a benchmarked ensemble marched over the whole horizon in ONE series call (as bench.py does)
    against the live oracle: every node and zone temperature of every model at the end of the horizon, the zone
    temperatures at every 24th main step, and equal no-mass iteration counts (branch parity).  north_star: "within
    1e-6 °C absolute over a full-year hourly EPW run"."""
    import oracle_py as O

    d, gains = ensemble_of(kind, n)
    weather, inputs = cases.synthetic_weather(hours + 1)
    inputs = np.ascontiguousarray(inputs[:hours])
    b = batch_cls(d)
    if gains is not None:
        b.set_zone_inputs(q_internal=gains)
    b.march_series(hours, weather, weather_mode=1, inputs=inputs, input_mode=1)
    got = b.get_state()
    models = [O.OracleModel(d, m) for m in range(d.n_models)]
    threads = threads or os.cpu_count() or 1
    _, ref = O.run_ensemble(models, hours, weather, inputs, weather_linear=True, threads=threads, want_state=True,
                            q_internal=gains)
    assert ref.shape == got.shape
    off = 0
    worst_node = worst_zone = 0.0
    for om in models:
        nz = om.n_zones
        worst_zone = max(worst_zone, float(np.abs(got[off:off + nz] - ref[off:off + nz]).max()))
        for s in range(om.n_surfaces):
            o = off + om.state_offset(s)
            nn = om.n_nodes(s)
            worst_node = max(worst_node, float(np.abs(got[o + 8:o + 8 + nn] - ref[o + 8:o + 8 + nn]).max()))
            # film coefficients and convective flows of the last sub-step
            assert np.abs(got[o:o + 4] - ref[o:o + 4]).max() <= 1e-5 * max(1.0, np.abs(ref[o:o + 4]).max())
        off += om.state_len
    assert off == got.size
    assert worst_node <= tol and worst_zone <= tol, (kind, worst_node, worst_zone)
    # Branch parity of the no-mass fixed point: identical iteration counts wherever no exit test is decided by the last
    # bits.  Over a whole year (tens of millions of iterations; exit tests `err > old_err`, `err/n < 0.01` on values that
    # differ from the oracle's in the last ulp) a handful of exits may flip without moving any temperature past `tol`;
    # short runs (every other parity case) demand equality.
    iters = b.nomass_iterations if hasattr(b, "nomass_iterations") else b.reduce().nomass_iterations
    want = sum(om.nomass_iterations for om in models)
    slack = 0 if hours <= 240 else max(1, int(1e-5 * want))
    assert abs(iters - want) <= slack, (kind, iters, want)
    return worst_node, worst_zone


def _march_and_compare(batch_cls, d, n_steps, weather, *, weather_mode=1, inputs=None, input_mode=1, zone_inputs=None,
                       tol=1e-6, label=""):
    """Batch against one oracle per model on the same arrays: all zones, all nodes, h and flows, iteration counts."""
    import oracle_py as O

    b = batch_cls(d)
    if zone_inputs:
        b.set_zone_inputs(**zone_inputs)
    b.march_series(n_steps, weather, weather_mode=weather_mode, inputs=inputs, input_mode=input_mode)
    got = b.get_state()
    off = 0
    iters = 0
    n_surf_total = int(d.model_surface_offset[-1])
    for m in range(d.n_models):
        om = O.OracleModel(d, m)
        z0, z1 = int(d.model_zone_offset[m]), int(d.model_zone_offset[m + 1])
        if zone_inputs:
            om.set_zone_inputs(**{k: np.asarray(v, float)[z0:z1] for k, v in zone_inputs.items()})
        so = om.initial_state()
        om.march_series(so, n_steps, weather, weather_mode=weather_mode, inputs=inputs, input_mode=input_mode,
                        surface_base=int(d.model_surface_offset[m]), n_surfaces_total=n_surf_total)
        mine = got[off:off + len(so)]
        nz = z1 - z0
        assert np.abs(mine[:nz] - so[:nz]).max() <= tol, (label, m, "zone", mine[:nz], so[:nz])
        for s in range(om.n_surfaces):
            o = om.state_offset(s)
            nn = om.n_nodes(s)
            err = np.abs(mine[o + 8:o + 8 + nn] - so[o + 8:o + 8 + nn]).max()
            assert err <= tol, (label, m, s, "nodes", err)
            assert np.abs(mine[o:o + 4] - so[o:o + 4]).max() <= 1e-5 * max(1.0, np.abs(so[o:o + 4]).max()), (label, m, s)
        iters += om.nomass_iterations
        off += len(so)
    assert off == got.size
    mine_it = b.nomass_iterations if hasattr(b, "nomass_iterations") else b.reduce().nomass_iterations
    assert mine_it == iters, (label, mine_it, iters)
    return got


def check_negative_front_solar(batch_cls, n=24, n_steps=12):
    """surface.rs:916-923: the back clamp tests `solar_front < 0.0` AFTER solar_front was clamped to >= 0, so a negative
    front irradiance zeroes the front only.  Massive walls (lean face path), mixed walls (no-mass ends) and a Trombe
    wall (general kernel), front = -5, back = 200 W/m2 on some steps."""
    weather, inputs = cases.synthetic_weather(n_steps + 1)
    inputs = inputs[:n_steps].copy()
    inputs[:, 0] = np.where(np.arange(n_steps) % 3 == 0, -5.0, inputs[:, 0])
    inputs[:, 1] = 200.0
    d = cases.wall_ensemble(n, seed=5, mixed_fraction=0.5)
    _march_and_compare(batch_cls, d, n_steps, weather, inputs=inputs, label="neg-solar walls")
    _march_and_compare(batch_cls, cases.validation_case("trombe_wall_full", 2), n_steps, weather, inputs=inputs,
                       label="neg-solar trombe")


def back_ambient_models(n=12, seed=61):
    """Models whose walls have an AmbientTemperature BACK boundary — the arm the reference builds from FRONT values
    (surface.rs:672-685): (a) Outdoor | wall | Ambient next to an ordinary wall of the same zone, (b) a floor whose
    front looks at the zone and whose back is the ground at a fixed temperature."""
    rng = cases.SplitMix64(seed)
    k, rho, cp, th = cases._random_layers(rng, n)
    layers = np.zeros(n, D.LAYER_DTYPE)
    layers["kind"] = D.LAYER_SOLID
    layers["thickness"], layers["conductivity"], layers["density"], layers["specific_heat"] = th, k, rho, cp
    for f, v in (("front_emissivity", 0.9), ("back_emissivity", 0.8), ("front_solar_abs", 0.6), ("back_solar_abs", 0.5)):
        layers[f] = v
    layers["solar_transmittance"] = D.NAN
    surfaces = np.zeros((n, 3), D.SURFACE_DTYPE)
    surfaces["construction"] = np.arange(n)[:, None]
    surfaces["area"], surfaces["perimeter"], surfaces["centroid_z"] = 12.0, 14.0, 1.5
    surfaces["normal"][:, 0] = (0.0, -1.0, 0.0)
    surfaces["normal"][:, 1] = (1.0, 0.0, 0.0)
    surfaces["normal"][:, 2] = (0.0, 0.0, 1.0)       # floor seen from the zone
    surfaces["front_zone"], surfaces["back_zone"] = -1, -1
    # 0: ordinary exterior wall
    surfaces["front_kind"][:, 0], surfaces["back_kind"][:, 0], surfaces["back_zone"][:, 0] = D.BOUNDARY_OUTDOOR, D.BOUNDARY_SPACE, 0
    # 1: Outdoor front, AmbientTemperature back
    surfaces["front_kind"][:, 1], surfaces["back_kind"][:, 1] = D.BOUNDARY_OUTDOOR, D.BOUNDARY_AMBIENT
    surfaces["back_ambient_t"][:, 1] = 18.0
    # 2: zone in front, ground (AmbientTemperature) behind
    surfaces["front_kind"][:, 2], surfaces["front_zone"][:, 2] = D.BOUNDARY_SPACE, 0
    surfaces["back_kind"][:, 2] = D.BOUNDARY_AMBIENT
    surfaces["back_ambient_t"][:, 2] = 12.0
    for f in ("front_hs_override", "back_hs_override", "front_emissivity_override", "back_emissivity_override"):
        surfaces[f] = D.NAN
    volume = rng.uniform(n, 50.0, 300.0)
    return D.BatchDesc(layers, np.arange(n + 1, dtype=np.int32), np.arange(n + 1, dtype=np.int32),
                       (np.arange(n + 1) * 3).astype(np.int32), volume, surfaces.reshape(-1), 2)


def check_back_ambient(batch_cls, n=12, n_steps=12):
    weather, inputs = cases.synthetic_weather(n_steps + 1)
    _march_and_compare(batch_cls, back_ambient_models(n), n_steps, weather, inputs=inputs[:n_steps], label="back-ambient")


def check_infiltration_and_ventilation(batch_cls, n=16, n_steps=12):
    """model.rs:524-544: infiltration AND ventilation together (ρ·V̇·cp at each stream's own temperature), on shoeboxes
    (shared zone sums), walls with their own zone (register zone) and double-glazed shoeboxes (general kernel)."""
    weather, inputs = cases.synthetic_weather(n_steps + 1)
    rng = cases.SplitMix64(77)
    for label, (d, gains) in (("inf+vent shoebox", cases.shoebox_ensemble(n, seed=71, steps_per_hour=2)),
                              ("inf+vent walls", (cases.wall_ensemble(n, seed=72, mixed_fraction=0.3), None)),
                              ("inf+vent 2g", cases.shoebox_ensemble(max(2, n // 4), seed=73, steps_per_hour=2,
                                                                     glazing=cases.double_glazing()))):
        nz = int(d.model_zone_offset[-1])
        zin = dict(q_internal=np.zeros(nz) if gains is None else gains,
                   infiltration_volume=rng.uniform(nz, 0.0, 0.08), infiltration_temperature=rng.uniform(nz, -5.0, 30.0),
                   ventilation_volume=rng.uniform(nz, 0.0, 0.2), ventilation_temperature=rng.uniform(nz, 14.0, 26.0))
        zin["ventilation_temperature"][::5] = np.nan      # element absent on some zones
        zin["infiltration_temperature"][1::7] = np.nan
        _march_and_compare(batch_cls, d, n_steps, weather, inputs=inputs[:n_steps], zone_inputs=zin, label=label)


def check_per_surface_inputs(batch_cls, n=10, n_steps=10):
    """HB_INPUT_PER_SURFACE: rows [t][surface][4] — each surface of each building gets its own irradiances."""
    weather, inputs = cases.synthetic_weather(n_steps + 1)
    for label, (d, gains) in (("per-surface shoebox", cases.shoebox_ensemble(n, seed=81, steps_per_hour=2)),
                              ("per-surface walls", (cases.wall_ensemble(3 * n, seed=82, mixed_fraction=0.4), None))):
        ns = int(d.model_surface_offset[-1])
        rng = cases.SplitMix64(83)
        scale = rng.uniform(n_steps * ns, 0.2, 1.3).reshape(n_steps, ns)
        per = np.zeros((n_steps, ns, 4))
        per[:, :, 0] = inputs[:n_steps, 0][:, None] * scale
        per[:, :, 1] = 30.0 * scale
        per[:, :, 2] = inputs[:n_steps, 2][:, None] * (0.9 + 0.1 * scale)
        per[:, :, 3] = inputs[:n_steps, 3][:, None]
        zin = None if gains is None else dict(q_internal=gains)
        _march_and_compare(batch_cls, d, n_steps, weather, inputs=per, input_mode=2, zone_inputs=zin, label=label)


def check_terrain_classes(batch_cls, n_steps=12):
    """surface.rs:135-166: every TerrainClass (and 'details without terrain') changes the wind-speed modifier of the
    forced-convection term; one wall + zone per class, same wall otherwise."""
    base = cases.validation_case("massive_full", 2)
    terr = [D.TERRAIN_NONE, D.TERRAIN_COUNTRY, D.TERRAIN_SUBURBS, D.TERRAIN_CITY, D.TERRAIN_OCEAN, D.TERRAIN_URBAN]
    n = len(terr)
    surfaces = np.repeat(base.surfaces, n)
    d = D.BatchDesc(base.layers, base.construction_offset, np.arange(n + 1), np.arange(n + 1),
                    np.repeat(base.zone_volume, n), surfaces, 2, np.array(terr, np.int32))
    weather, inputs = cases.synthetic_weather(n_steps + 1)
    weather[:, 2] += 2.0   # keep some wind
    got = _march_and_compare(batch_cls, d, n_steps, weather, inputs=inputs[:n_steps], label="terrain")
    per = got.reshape(n, -1)
    # NONE == URBAN == SUBURBS (0.22, 370); the others differ
    assert np.array_equal(per[0], per[5]) and np.array_equal(per[0], per[2])
    for k in (1, 3, 4):
        assert np.abs(per[k] - per[0]).max() > 1e-4, k


def tilted_glazing_models(steps_per_hour=2):
    """Double-glazed fenestrations at tilts 0° (roof), 45°, 60°, 75°, 90° and 135° next to an opaque wall, one model per
    tilt: with the tilt complement taken when t_front > t_back (gas.rs:137) every regime of `nusselt`
    (gas.rs:197-221: 0-60°, 60°, 60-90°, 90°, 90-180°) runs on the device."""
    tilts = [0.0, 45.0, 60.0, 75.0, 90.0, 135.0]
    n = len(tilts)
    glazing = cases.double_glazing(D.GAS_AIR, 0.02)
    layers = [cases.concrete(0.15, 0.9, 0.6)] + glazing
    surfaces = []
    for k, t in enumerate(tilts):
        a = math.radians(t)
        normal = (0.0, -math.sin(a), math.cos(a))
        surfaces.append(D.surface(0, 10.0, 13.0, (0.0, -1.0, 0.0), 1.5, front=(D.BOUNDARY_OUTDOOR, 0),
                                  back=(D.BOUNDARY_SPACE, 0)))
        surfaces.append(D.surface(1, 4.0, 8.0, normal, 2.0, front=(D.BOUNDARY_OUTDOOR, 0), back=(D.BOUNDARY_SPACE, 0),
                                  is_fenestration=True))
    return D.BatchDesc(np.array(layers, D.LAYER_DTYPE), [0, 1, 4], np.arange(n + 1), np.arange(n + 1) * 2,
                       np.full(n, 60.0), np.array(surfaces, D.SURFACE_DTYPE), steps_per_hour)


def check_tilted_glazing(batch_cls, n_steps=24):
    weather, inputs = cases.synthetic_weather(n_steps + 1)
    d = tilted_glazing_models()
    nz = d.n_models
    # heat some zones above and cool others below the outdoor temperature: both signs of (t_front - t_back)
    q = np.where(np.arange(nz) % 2 == 0, 1500.0, -1500.0)
    _march_and_compare(batch_cls, d, n_steps, weather, inputs=inputs[:n_steps], zone_inputs=dict(q_internal=q),
                       label="tilted glazing")
    q2 = -q
    _march_and_compare(batch_cls, d, n_steps, weather, inputs=inputs[:n_steps], zone_inputs=dict(q_internal=q2),
                       label="tilted glazing (signs swapped)")


def check_marches_again_after_failure(batch_cls):
    """The reference's Err / panic belongs to ONE march call: after a reported physics failure, set_state with a sane
    state and march again must succeed (status words are per call)."""
    from heat_b200 import HeatB200Error

    b = batch_cls(cases.validation_case("massive_full"))
    s0 = b.get_state()
    try:
        b.march_series(1, np.array([[float("nan"), 0.0, 1.0]]))
        raise AssertionError("NaN weather must be reported")
    except HeatB200Error as e:
        assert e.status == 3
    b.set_state(s0)
    b.march_series(3, np.array([[10.0, 90.0, 2.0]] * 3))
    s1 = b.get_state()
    assert np.all(np.isfinite(s1)) and b.reduce().status_flags == 0
    b2 = batch_cls(cases.validation_case("massive_full"))
    b2.march_series(3, np.array([[10.0, 90.0, 2.0]] * 3))
    assert np.array_equal(b2.get_state(), s1)


def check_unstable_wall_fails_like_reference(batch_cls, hours=36):
    """A wall on which the reference's explicit RK4 is unstable at the sub-step its own rule picks (cases.unstable_wall:
    dt·λmax = 2.807 > 2.7853) diverges and ends in the reference's panic — here in HB_ERR_PHYSICS with the same words,
    not in silently wrong numbers; while it is still bounded (the first hours) the two agree."""
    import oracle_py as O
    from heat_b200 import HeatB200Error

    d = cases.unstable_wall()
    weather, inputs = cases.synthetic_weather(hours + 1)
    om = O.OracleModel(d, 0)
    arr = om.surface_arrays(0, 4)
    number = cases.rk4_number_np(arr["mass"], arr["u"], arr["massive_chunks"], om.dt)
    assert 2.80 < number < 2.81 and number > cases.RK4_STABILITY_LIMIT, number
    so = om.initial_state()
    om.march_series(so, 3, weather, weather_mode=1, inputs=inputs[:3], input_mode=1)
    b = batch_cls(d)
    b.march_series(3, weather, weather_mode=1, inputs=inputs[:3], input_mode=1)
    assert np.abs(b.get_state() - so).max() <= 1e-6
    with np.errstate(all="ignore"):
        try:
            om.march_series(so, hours - 3, weather[3:], weather_mode=1, inputs=inputs[3:hours], input_mode=1)
            raise AssertionError("the oracle must panic on the unstable wall")
        except O.OracleError as e:
            assert "NaN" in str(e), e
    try:
        b.march_series(hours - 3, weather[3:], weather_mode=1, inputs=inputs[3:hours], input_mode=1)
        failed = bool(getattr(b, "status", 0))   # the emulator reports through its status word
    except HeatB200Error as e:
        failed = e.status == 3 and "NaN" in str(e)
    assert failed, "the unstable wall must be reported as a physics failure"


def check_slow_nomass_convergence(batch_cls, hours=68):
    """cases.slow_nomass_shoebox: a main step in which the reference's no-mass fixed point takes > 10^5 passes in one
    solve.  The reference's only limit is 99 199 000 passes (surface.rs:869-873); the library must carry on like it —
    same pass count, same temperatures — instead of reporting a failure of its own making."""
    import oracle_py as O

    d, gains = cases.slow_nomass_shoebox()
    weather, inputs = cases.synthetic_weather(hours + 1)
    om = O.OracleModel(d, 0)
    om.set_zone_inputs(q_internal=gains)
    so = om.initial_state()
    om.march_series(so, hours - 3, weather, weather_mode=1, inputs=inputs[:hours - 3], input_mode=1)
    before = om.nomass_iterations
    om.march_series(so, 3, weather[hours - 3:], weather_mode=1, inputs=inputs[hours - 3:hours], input_mode=1)
    assert om.nomass_iterations - before > 120_000, (before, om.nomass_iterations)   # hour 65 alone: ~1.4·10^5
    b = batch_cls(d)
    b.set_zone_inputs(q_internal=gains)
    b.march_series(hours, weather, weather_mode=1, inputs=inputs[:hours], input_mode=1)
    got = b.get_state()
    assert np.all(np.isfinite(got)) and np.abs(got - so).max() <= 1e-6, np.abs(got - so).max()
    its = b.nomass_iterations if hasattr(b, "nomass_iterations") else b.reduce().nomass_iterations
    assert its == om.nomass_iterations, (its, om.nomass_iterations)


def check_scheduled_infiltration_ventilation(batch_cls, n=10, n_steps=8):
    """Infiltration / ventilation volume flow and temperature as per-main-step schedules (model.rs:524-544: the reference
    reads these state elements every sub-step and the caller rewrites them between march calls): one series call with
    the four tables (+ a gain schedule) against the oracle marched step by step with the inputs reset in between.
    Walls with their own zone (register zone), shoeboxes (shared zone) and double-glazed shoeboxes (general kernel)."""
    import oracle_py as O

    weather, inputs = cases.synthetic_weather(n_steps + 1)
    builders = (lambda: (cases.wall_ensemble(2 * n, seed=91, mixed_fraction=0.4), None),
                lambda: cases.shoebox_ensemble(n, seed=92, steps_per_hour=2),
                lambda: cases.shoebox_ensemble(max(2, n // 3), seed=93, steps_per_hour=2, glazing=cases.double_glazing()))
    for build in builders:
        d, gains = build()
        nz = int(d.model_zone_offset[-1])
        rng = cases.SplitMix64(4321)
        tab = dict(zone_gain=rng.uniform(n_steps * nz, -300.0, 900.0).reshape(n_steps, nz),
                   zone_infiltration_volume=rng.uniform(n_steps * nz, 0.0, 0.1).reshape(n_steps, nz),
                   zone_infiltration_temperature=rng.uniform(n_steps * nz, -5.0, 32.0).reshape(n_steps, nz),
                   zone_ventilation_volume=rng.uniform(n_steps * nz, 0.0, 0.25).reshape(n_steps, nz),
                   zone_ventilation_temperature=rng.uniform(n_steps * nz, 12.0, 28.0).reshape(n_steps, nz))
        tab["zone_ventilation_temperature"][2] = np.nan          # ventilation off during step 2
        tab["zone_infiltration_temperature"][::3, ::2] = np.nan   # and infiltration absent here and there
        b = batch_cls(d)
        if gains is not None:
            b.set_zone_inputs(q_internal=gains)   # overridden by the schedule while it runs
        b.march_series(n_steps, weather, weather_mode=1, inputs=inputs[:n_steps], input_mode=1, **tab)
        got = b.get_state()
        off = 0
        zoff = d.model_zone_offset
        for m in range(d.n_models):
            z0, z1 = int(zoff[m]), int(zoff[m + 1])
            om = O.OracleModel(d, m)
            so = om.initial_state()
            for t in range(n_steps):
                om.set_zone_inputs(q_internal=tab["zone_gain"][t, z0:z1],
                                   infiltration_volume=tab["zone_infiltration_volume"][t, z0:z1],
                                   infiltration_temperature=tab["zone_infiltration_temperature"][t, z0:z1],
                                   ventilation_volume=tab["zone_ventilation_volume"][t, z0:z1],
                                   ventilation_temperature=tab["zone_ventilation_temperature"][t, z0:z1])
                om.march_series(so, 1, weather[t:t + 2], weather_mode=1, inputs=inputs[t:t + 1], input_mode=1)
            mine = got[off:off + len(so)]
            nzm = z1 - z0
            assert np.abs(mine[:nzm] - so[:nzm]).max() <= 1e-6, (m, mine[:nzm], so[:nzm])
            for s in range(om.n_surfaces):
                o = om.state_offset(s)
                nn = om.n_nodes(s)
                assert np.abs(mine[o + 8:o + 8 + nn] - so[o + 8:o + 8 + nn]).max() <= 1e-6, (m, s)
            off += len(so)
        # the schedules must matter
        b2 = batch_cls(d)
        if gains is not None:
            b2.set_zone_inputs(q_internal=gains)
        b2.march_series(n_steps, weather, weather_mode=1, inputs=inputs[:n_steps], input_mode=1)
        assert np.abs(b2.get_state() - got).max() > 1e-3

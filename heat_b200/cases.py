"""Model builders for the reference's validation cases and the synthetic ensembles of BASELINE.json.

`single_zone_test_building` restates what `simple_test_models::get_single_zone_test_building` (un-vendored
dev-dependency) produces, as far as the thermal path reads it: one space, one south-facing vertical wall whose
front is Outdoor and whose back is the space, optionally a window (a fenestration of the same construction) —
geometry (inferred) from tests/validate_wall_heat_transfer.rs:95-191,713-738 and the IDFs under tests/*/in.idf.
"""
from __future__ import annotations

import math
from typing import Sequence

import numpy as np

from . import desc as D

# TestMat::{Concrete, Polyurethane} — tests/massive_full/in.idf:47-56, tests/mixed_full/in.idf:59-68,
# src/surface.rs:1048-1075
CONCRETE = dict(conductivity=0.816, density=1700.0, specific_heat=800.0)
POLYURETHANE = dict(conductivity=0.0252, density=17.5, specific_heat=2400.0)
# TestMat::Glass is not recoverable from the reference tree; soda-lime glass values, builder-defined.
GLASS = dict(conductivity=1.0, density=2500.0, specific_heat=840.0)


def concrete(thickness, emissivity=D.NAN, solar_abs=D.NAN):
    return D.solid_layer(thickness, **CONCRETE, front_emissivity=emissivity, back_emissivity=emissivity,
                         front_solar_abs=solar_abs, back_solar_abs=solar_abs)


def polyurethane(thickness, emissivity=D.NAN, solar_abs=D.NAN):
    return D.solid_layer(thickness, **POLYURETHANE, front_emissivity=emissivity, back_emissivity=emissivity,
                         front_solar_abs=solar_abs, back_solar_abs=solar_abs)


def glass(thickness, tau, emissivity=D.NAN, solar_abs=D.NAN):
    return D.solid_layer(thickness, **GLASS, front_emissivity=emissivity, back_emissivity=emissivity,
                         front_solar_abs=solar_abs, back_solar_abs=solar_abs, solar_transmittance=tau)


def polygon_props(vertices: Sequence[Sequence[float]]):
    """normal (unit), area, perimeter, centroid of a planar polygon (geometry3d, un-vendored; Newell's method)."""
    v = np.asarray(vertices, float)
    nxt = np.roll(v, -1, axis=0)
    nvec = np.array([
        np.sum((v[:, 1] - nxt[:, 1]) * (v[:, 2] + nxt[:, 2])),
        np.sum((v[:, 2] - nxt[:, 2]) * (v[:, 0] + nxt[:, 0])),
        np.sum((v[:, 0] - nxt[:, 0]) * (v[:, 1] + nxt[:, 1])),
    ])
    area = 0.5 * float(np.linalg.norm(nvec))
    normal = nvec / np.linalg.norm(nvec)
    perimeter = float(np.sum(np.linalg.norm(nxt - v, axis=1)))
    centroid = v.mean(axis=0)
    return normal, area, perimeter, centroid


def single_zone_test_building(construction, *, zone_volume=40.0, surface_height=2.0, surface_width=2.0,
                              window_height=0.0, window_width=0.0, emissivity=0.84, solar_absorbtance=0.7,
                              steps_per_hour=20, front_hs=D.NAN, back_hs=D.NAN, back_emissivity_override=D.NAN):
    """One space + one exterior wall (+ optional window). `construction`: list of ('concrete'|'polyurethane', th)."""
    makers = {"concrete": concrete, "polyurethane": polyurethane}
    layers = [makers[name](th, emissivity, solar_absorbtance) for name, th in construction]
    win_area = window_height * window_width
    normal = (0.0, -1.0, 0.0)
    surfaces = [D.surface(0, surface_height * surface_width - win_area, 2.0 * (surface_height + surface_width), normal,
                          surface_height / 2.0, front=(D.BOUNDARY_OUTDOOR, 0), back=(D.BOUNDARY_SPACE, 0),
                          front_hs=front_hs, back_hs=back_hs, back_emissivity=back_emissivity_override)]
    if win_area > 0.0:
        surfaces.append(D.surface(0, win_area, 2.0 * (window_height + window_width), normal, surface_height / 2.0,
                                  front=(D.BOUNDARY_OUTDOOR, 0), back=(D.BOUNDARY_SPACE, 0), is_fenestration=True))
    return D.single_model([layers], [zone_volume], surfaces, steps_per_hour)


def validation_case(name: str, steps_per_hour: int = 20) -> D.BatchDesc:
    """The EnergyPlus-driven cases of tests/validate_wall_heat_transfer.rs:792-994 (+ the commented-out Trombe wall)."""
    kind, _, variant = name.partition("_")
    if kind in ("massive", "mixed", "nomass"):
        eps = 0.9 if variant in ("full", "yes_ir_no_solar") else 0.0
        alpha = 0.7 if variant in ("full", "no_ir_yes_solar") else 0.0
        cons = {"massive": [("concrete", 0.2)], "nomass": [("polyurethane", 0.02)],
                "mixed": [("polyurethane", 0.02), ("concrete", 0.2), ("polyurethane", 0.02)]}[kind]
        # march_test_model (:713-738) + `thermal_model.surfaces[0].back_emissivity = 0.0` (:630)
        return single_zone_test_building(cons, zone_volume=600.0, surface_height=3.0, surface_width=20.0,
                                         emissivity=eps, solar_absorbtance=alpha, steps_per_hour=steps_per_hour,
                                         back_emissivity_override=0.0)
    if name in ("tilted", "horizontal"):
        verts = {"tilted": [(0, 0, 0), (20, 0, 0), (20, 2.12132034357, 2.12132034357), (0, 2.12132034357, 2.12132034357)],
                 "horizontal": [(10, 3, 14.9), (10, 0, 14.9), (30, 0, 14.9), (30, 3, 14.9)]}[name]
        normal, area, perimeter, centroid = polygon_props(verts)  # tests/{tilted,horizontal}/back.spl
        layers = [concrete(0.2, 0.9, 0.7)]
        s = D.surface(0, area, perimeter, tuple(normal), float(centroid[2]), front=(D.BOUNDARY_OUTDOOR, 0),
                      back=(D.BOUNDARY_SPACE, 0), back_emissivity=0.0)
        return D.single_model([layers], [600.0], [s], steps_per_hour)
    if name == "trombe_wall_full":
        # validate_wall_heat_transfer.rs:1091-1099 (commented out in the reference; E+ columns are a missing blob)
        layers = [concrete(0.2, 0.9, 0.08), D.gas_layer(0.05, D.GAS_AIR), glass(0.03, 0.82, 0.9, 0.08)]
        s = D.surface(0, 60.0, 46.0, (0.0, -1.0, 0.0), 1.5, front=(D.BOUNDARY_OUTDOOR, 0), back=(D.BOUNDARY_SPACE, 0),
                      back_emissivity=0.0)
        return D.single_model([layers], [600.0], [s], steps_per_hour)
    raise KeyError(name)


# ------------------------------------------------------------------------------------------------------------------
# Synthetic ensembles (BASELINE.json configs 3-5; distributions in SURVEY.md §8(d))
# ------------------------------------------------------------------------------------------------------------------
class SplitMix64:
    """splitmix64 stream → uniform doubles; vectorised."""

    def __init__(self, seed: int):
        self.state = np.uint64(seed)

    def next_u64(self, n: int) -> np.ndarray:
        with np.errstate(over="ignore"):
            idx = np.arange(1, n + 1, dtype=np.uint64)
            z = self.state + idx * np.uint64(0x9E3779B97F4A7C15)
            self.state = self.state + np.uint64(n) * np.uint64(0x9E3779B97F4A7C15)
            z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
            z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
            return z ^ (z >> np.uint64(31))

    def uniform(self, n: int, lo=0.0, hi=1.0) -> np.ndarray:
        u = (self.next_u64(n) >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)
        return lo + (hi - lo) * u


def _random_layers(rng: SplitMix64, n: int):
    k = np.exp(rng.uniform(n, math.log(0.05), math.log(2.5)))
    rho = rng.uniform(n, 300.0, 2500.0)
    cp = rng.uniform(n, 700.0, 2000.0)
    th = rng.uniform(n, 0.03, 0.30)
    return k, rho, cp, th


RK4_STABILITY_LIMIT = 2.785293563405282   # |R(z)| <= 1 of the classical RK4 on the negative real axis: -2.7853 <= z <= 0
FILM_MAX = 30.0                           # W/m²K: TARP forced convection at the synthetic weather's 6 m/s + linearised IR


def rk4_number_np(mass, u, massive_chunks, dt: float, film: float = FILM_MAX) -> float:
    """dt·λmax(C⁻¹K) over the massive chunks of a wall — numpy statement of hb_surface_rk4_number (include/heat_b200.h)
    for planners that must not load the library (bench.py's reference arm) and for the cross-check in tests/."""
    mass, u = np.asarray(mass, np.float64), np.asarray(u, np.float64)
    link = np.where(u < 0.0, 8.0, u)
    n_nodes, worst = len(mass), 0.0
    for ini, fin in np.asarray(massive_chunks).reshape(-1, 2):
        n = int(fin - ini)
        if n <= 0:
            continue
        g = np.arange(ini, fin)
        left = np.where(g == 0, film, link[np.maximum(g - 1, 0)])
        right = np.where(g == n_nodes - 1, film, link[g])
        a = np.diag((left + right) / mass[g])
        if n > 1:
            e = link[g[:-1]] / np.sqrt(mass[g[:-1]] * mass[g[1:]])
            a -= np.diag(e, 1) + np.diag(e, -1)
        worst = max(worst, float(np.linalg.eigvalsh(a)[-1]))
    return worst * dt


def unstable_wall(steps_per_hour: int = 1) -> D.BatchDesc:
    """Wall 347550 of the raw 800 000-wall draw of seed 42 (the 8-GPU configs[2] job): four solid layers, the third —
    4 cm at k = 2.34 — left without inner nodes by the reference's rule between two light layers.  At the rule's
    sub-step (30.5 s) its interface nodes give dt·λmax = 2.807 > RK4_STABILITY_LIMIT: the reference's march diverges
    within a simulated day and panics."""
    stack = [(0.26476804, 0.79516467, 473.21812667, 1954.91784947), (0.10924581, 0.05198719, 385.13555222, 934.12034676),
             (0.04085655, 2.33743229, 462.85227834, 1455.90056047), (0.05593663, 0.0571488, 1276.53881926, 754.52769089)]
    layers = [D.solid_layer(th, conductivity=k, density=rho, specific_heat=cp, front_emissivity=0.87629394,
                            back_emissivity=0.87629394, front_solar_abs=0.8935592, back_solar_abs=0.8935592)
              for th, k, rho, cp in stack]
    s = D.surface(0, 60.0, 46.0, (-0.26801150678176583, -0.9634157110160533, 0.0), 1.5, front=(D.BOUNDARY_OUTDOOR, 0),
                  back=(D.BOUNDARY_SPACE, 0))
    return D.single_model([layers], [351.44779206], [s], steps_per_hour)


def wall_ensemble(n_walls: int, seed: int = 42, *, mixed_fraction: float = 0.0, steps_per_hour: int = 1,
                  max_layers: int = 4, max_nodes: int = 128, planner=None) -> D.BatchDesc:
    """`n_walls` independent models, each one exterior wall + its own zone (massive_full topology).

    Massive walls: 1..max_layers random solid layers. Mixed walls (fraction `mixed_fraction`): polyurethane /
    random solid / polyurethane, whose outer layers the reference's rule classifies as no-mass.
    Walls that the reference's discretization rule cuts into more than `max_nodes` nodes (thick layers of very low
    diffusivity) or leaves without any massive node are re-drawn, up to 8 times, then replaced by a 20 cm concrete wall.
    So are the few stacks (about 1 in 10^5: a thin, very conductive layer the rule leaves without inner nodes, between
    lighter ones) on which the reference's own explicit RK4 is unstable at the sub-step its rule picks — dt·λmax above
    RK4_STABILITY_LIMIT, see hb_surface_rk4_number: the reference marches such a wall to NaN and panics, and so, to the
    same message, does this library (tests/parity_cases.py check_unstable_wall_fails_like_reference); they are not valid
    input of the path.

    `planner(desc) -> (n_nodes[n_walls], n_massive_chunks[n_walls], rk4_number[n_walls])`: who evaluates the
    discretization rule for the re-draw test; default: this library's host planning (`hb_create` with HB_DEVICE_NONE).
    """
    d = _wall_ensemble_draw(n_walls, seed, mixed_fraction, steps_per_hour, max_layers)
    if not max_nodes:
        return d
    if planner is None:
        def planner(desc):
            from .model import ThermalBatch
            plan = ThermalBatch(desc, device=-1)  # host-only planning handle (discretization runs on the host anyway)
            try:
                return _bulk_n_nodes(plan, n_walls)
            finally:
                plan.close()
    for attempt in range(1, 10):
        n_nodes, n_massive, rk4 = planner(d)
        # re-draw too, walls whose every layer the rule classifies no-mass (nothing left to march)
        bad = np.nonzero((n_nodes > max_nodes) | (n_massive == 0) | (rk4 > RK4_STABILITY_LIMIT))[0]
        if bad.size == 0:
            return d
        # the layers to re-draw: every layer of a bad wall except the fixed polyurethane skins of mixed walls
        idx = np.concatenate([np.arange(d.construction_offset[w], d.construction_offset[w + 1]) for w in bad])
        skin = (d.layers["thickness"][idx] == 0.02) & (d.layers["conductivity"][idx] == POLYURETHANE["conductivity"])
        idx = idx[~skin]
        if attempt < 9:
            k, rho, cp, th = _random_layers(SplitMix64(seed + 7919 * attempt), idx.size)
        else:
            k, rho, cp = (np.full(idx.size, CONCRETE[f]) for f in ("conductivity", "density", "specific_heat"))
            th = np.full(idx.size, 0.05)
        d.layers["conductivity"][idx], d.layers["density"][idx] = k, rho
        d.layers["specific_heat"][idx], d.layers["thickness"][idx] = cp, th
    return d


def _bulk_n_nodes(plan, n: int) -> np.ndarray:
    import ctypes as C
    from . import _lib
    out = np.empty(n, np.int64)
    massive = np.empty(n, np.int64)
    rk4 = np.empty(n, np.float64)
    info, number = _lib.SurfaceInfoC(), C.c_double(0.0)
    fn, fr, h = plan._lib.hb_get_surface_info, plan._lib.hb_surface_rk4_number, plan._h
    pi, pn, film = C.byref(info), C.byref(number), C.c_double(FILM_MAX)
    for s in range(n):
        fn(h, s, pi)
        fr(h, s, film, pn)
        out[s] = info.n_nodes
        massive[s] = info.n_massive_chunks
        rk4[s] = number.value
    return out, massive, rk4


def _wall_ensemble_draw(n_walls, seed, mixed_fraction, steps_per_hour, max_layers) -> D.BatchDesc:
    rng = SplitMix64(seed)
    n_lay = 1 + (rng.next_u64(n_walls) % np.uint64(max_layers)).astype(np.int64)
    is_mixed = rng.uniform(n_walls) < mixed_fraction
    n_lay = np.where(is_mixed, 3, n_lay)
    off = np.zeros(n_walls + 1, np.int64)
    np.cumsum(n_lay, out=off[1:])
    total = int(off[-1])
    k, rho, cp, th = _random_layers(rng, total)
    eps = rng.uniform(n_walls, 0.7, 0.95)
    alpha = rng.uniform(n_walls, 0.2, 0.9)
    azimuth = rng.uniform(n_walls, 0.0, 2.0 * math.pi)
    volume = rng.uniform(n_walls, 50.0, 600.0)

    layers = np.zeros(total, D.LAYER_DTYPE)
    layers["kind"] = D.LAYER_SOLID
    layers["thickness"], layers["conductivity"], layers["density"], layers["specific_heat"] = th, k, rho, cp
    wall_of_layer = np.repeat(np.arange(n_walls), n_lay)
    for f, v in (("front_emissivity", eps), ("back_emissivity", eps), ("front_solar_abs", alpha), ("back_solar_abs", alpha)):
        layers[f] = v[wall_of_layer]
    layers["solar_transmittance"] = D.NAN
    # mixed walls: first and third layers become 2 cm polyurethane
    first = off[:-1][is_mixed]
    for idx in (first, first + 2):
        layers["thickness"][idx] = 0.02
        layers["conductivity"][idx] = POLYURETHANE["conductivity"]
        layers["density"][idx] = POLYURETHANE["density"]
        layers["specific_heat"][idx] = POLYURETHANE["specific_heat"]

    surfaces = np.zeros(n_walls, D.SURFACE_DTYPE)
    surfaces["construction"] = np.arange(n_walls)
    surfaces["front_kind"], surfaces["back_kind"] = D.BOUNDARY_OUTDOOR, D.BOUNDARY_SPACE
    surfaces["front_zone"], surfaces["back_zone"] = -1, 0
    surfaces["area"], surfaces["perimeter"], surfaces["centroid_z"] = 60.0, 46.0, 1.5
    surfaces["normal"][:, 0] = np.sin(azimuth)
    surfaces["normal"][:, 1] = np.cos(azimuth)
    for f in ("front_hs_override", "back_hs_override", "front_emissivity_override", "back_emissivity_override"):
        surfaces[f] = D.NAN
    # Mixed walls end in a no-mass node facing the zone.  The reference's damped fixed point for such a node
    # (surface.rs:790-898) is not contractive once 4εσT³ > 3(u + h): its result is then chaotic at the 0.1 °C level
    # (tests/test_reference_sensitivity.py).  The reference's own validation driver sets the
    # zone-facing emissivity to 0 (validate_wall_heat_transfer.rs:630); the ensembles do the same for mixed walls.
    surfaces["back_emissivity_override"][is_mixed] = 0.0
    ar = np.arange(n_walls + 1, dtype=np.int32)
    return D.BatchDesc(layers, off.astype(np.int32), ar, ar, volume, surfaces, steps_per_hour)


def shoebox_ensemble(n_buildings: int, seed: int = 43, *, steps_per_hour: int = 1, glazing=None, params=None):
    """`n_buildings` single-zone shoeboxes: 4 walls + roof (Outdoor ↔ Space), floor (AmbientTemperature front) and one
    single-layer window, parametric sweep. Returns (BatchDesc, q_internal[n_buildings]).

    `glazing`: list of layers of the window construction shared by all buildings (default: one 6 mm pane); e.g.
    `double_glazing()` — panes + gas cavity, ISO 9050 optics (glazing.rs:180-286) and Cavity::u_value (cavity.rs:59-69).
    `params`: (n_lay[n], k, rho, cp, th [Σ n_lay], eps[n], alpha[n], volume[n], gains[n]) instead of the seeded draw."""
    rng = SplitMix64(seed)
    n = n_buildings
    # one opaque construction per building (1-3 layers) + one glazing construction shared by all
    n_lay = 1 + (rng.next_u64(n) % np.uint64(3)).astype(np.int64)
    off = np.zeros(n + 2, np.int64)
    np.cumsum(n_lay, out=off[1:n + 1])
    total = int(off[n])
    glazing = [glass(0.006, 0.6, 0.84, 0.2)] if glazing is None else list(glazing)
    off[n + 1] = total + len(glazing)
    k, rho, cp, th = _random_layers(rng, total)
    eps = rng.uniform(n, 0.7, 0.95)
    alpha = rng.uniform(n, 0.2, 0.9)
    volume = rng.uniform(n, 50.0, 600.0)
    gains = rng.uniform(n, 0.0, 500.0)
    if params is not None:
        n_lay, k, rho, cp, th, eps, alpha, volume, gains = (np.asarray(a) for a in params)
        np.cumsum(n_lay, out=off[1:n + 1])
        total = int(off[n])
        off[n + 1] = total + len(glazing)
    layers = np.zeros(total + len(glazing), D.LAYER_DTYPE)
    layers["kind"] = D.LAYER_SOLID
    layers["thickness"][:total], layers["conductivity"][:total] = th, k
    layers["density"][:total], layers["specific_heat"][:total] = rho, cp
    b_of_layer = np.repeat(np.arange(n), n_lay)
    for f, v in (("front_emissivity", eps), ("back_emissivity", eps), ("front_solar_abs", alpha), ("back_solar_abs", alpha)):
        layers[f][:total] = v[b_of_layer]
    layers["solar_transmittance"] = D.NAN
    for i, g in enumerate(glazing):
        layers[total + i] = g

    height = 3.0
    side = np.sqrt(volume / height)
    surfaces = np.zeros((n, 7), D.SURFACE_DTYPE)
    surfaces["construction"] = np.arange(n)[:, None]
    surfaces["construction"][:, 6] = n
    surfaces["is_fenestration"][:, 6] = 1
    surfaces["front_kind"], surfaces["back_kind"] = D.BOUNDARY_OUTDOOR, D.BOUNDARY_SPACE
    surfaces["front_zone"], surfaces["back_zone"] = -1, 0
    normals = np.array([(0, -1, 0), (1, 0, 0), (0, 1, 0), (-1, 0, 0), (0, 0, 1), (0, 0, -1), (0, -1, 0)], float)
    surfaces["normal"] = normals[None]
    wall_area = side * height
    win_area = 0.3 * wall_area
    for j in range(4):
        surfaces["area"][:, j] = wall_area - (win_area if j == 0 else 0.0)
        surfaces["perimeter"][:, j] = 2.0 * (side + height)
        surfaces["centroid_z"][:, j] = height / 2.0
    surfaces["area"][:, 4] = side * side
    surfaces["perimeter"][:, 4] = 4.0 * side
    surfaces["centroid_z"][:, 4] = height
    surfaces["area"][:, 5] = side * side
    surfaces["perimeter"][:, 5] = 4.0 * side
    surfaces["centroid_z"][:, 5] = 0.0
    surfaces["front_kind"][:, 5] = D.BOUNDARY_AMBIENT       # floor: ground modelled as a fixed temperature
    surfaces["front_ambient_t"][:, 5] = 15.0
    surfaces["area"][:, 6] = win_area
    surfaces["perimeter"][:, 6] = 2.0 * (np.sqrt(win_area) * 2.0)
    surfaces["centroid_z"][:, 6] = height / 2.0
    for f in ("front_hs_override", "back_hs_override", "front_emissivity_override", "back_emissivity_override"):
        surfaces[f] = D.NAN
    zo = np.arange(n + 1, dtype=np.int32)
    so = (np.arange(n + 1, dtype=np.int64) * 7).astype(np.int32)
    return D.BatchDesc(layers, off.astype(np.int32), zo, so, volume, surfaces.reshape(-1), steps_per_hour), gains


def slow_nomass_shoebox():
    """Building 872809 of the 10^6-shoebox draw of seed 43 (the 8-GPU configs[3] job): one 3 cm light, conductive layer,
    so all seven surfaces are pairs of no-mass nodes and the zone is the only mass.  In hour 65 of the synthetic weather
    the reference's damped fixed point (surface.rs:808-896) creeps: about 1.4·10^5 passes in one main step, more than
    10^5 of them in a single solve — far below the reference's own limit of 99 199 000, which is the only limit there is.
    Returns (BatchDesc, q_internal[1])."""
    one = lambda v: np.array([v])
    return shoebox_ensemble(1, params=(one(1), one(1.8548797806167274), one(361.55561998828756), one(748.820901506258),
                                       one(0.031159117865444532), one(0.8685396147411738), one(0.4314982273132669),
                                       one(303.80703082049365), one(108.11485792021885)))


def double_glazing(gas=D.GAS_ARGON, gap=0.012):
    """6 mm pane | gas cavity | 4 mm low-absorption pane (the pattern of tests/test_host_planning.py)."""
    return [glass(0.006, 0.7, 0.84, 0.1), D.gas_layer(gap, gas), glass(0.004, 0.8, 0.84, 0.05)]


def two_zone_ensemble(n_buildings: int, seed: int = 44, *, steps_per_hour: int = 2):
    """`n_buildings` two-zone buildings: zones 0 and 1 with four outdoor walls each and one partition whose front face
    looks at zone 0 and whose back face looks at zone 1 (a surface between two spaces, model.rs:79-96), plus a
    single-pane window on zone 1.  Returns (BatchDesc, q_internal[2·n_buildings])."""
    rng = SplitMix64(seed)
    n = n_buildings
    n_lay = 1 + (rng.next_u64(n) % np.uint64(2)).astype(np.int64)
    off = np.zeros(n + 2, np.int64)
    np.cumsum(n_lay, out=off[1:n + 1])
    total = int(off[n])
    off[n + 1] = total + 1
    k, rho, cp, th = _random_layers(rng, total)
    layers = np.zeros(total + 1, D.LAYER_DTYPE)
    layers["kind"] = D.LAYER_SOLID
    layers["thickness"][:total], layers["conductivity"][:total] = th, k
    layers["density"][:total], layers["specific_heat"][:total] = rho, cp
    for f, v in (("front_emissivity", 0.9), ("back_emissivity", 0.85), ("front_solar_abs", 0.6), ("back_solar_abs", 0.5)):
        layers[f][:total] = v
    layers["solar_transmittance"] = D.NAN
    layers[total] = glass(0.004, 0.7, 0.84, 0.15)
    volume = rng.uniform(2 * n, 40.0, 300.0)
    gains = rng.uniform(2 * n, 0.0, 400.0)
    NSURF = 10
    surfaces = np.zeros((n, NSURF), D.SURFACE_DTYPE)
    surfaces["construction"] = np.arange(n)[:, None]
    surfaces["front_kind"], surfaces["back_kind"] = D.BOUNDARY_OUTDOOR, D.BOUNDARY_SPACE
    surfaces["front_zone"] = -1
    surfaces["back_zone"][:, 0:4] = 0
    surfaces["back_zone"][:, 4:8] = 1
    surfaces["front_kind"][:, 8] = D.BOUNDARY_SPACE       # the partition
    surfaces["front_zone"][:, 8] = 0
    surfaces["back_zone"][:, 8] = 1
    surfaces["construction"][:, 9] = n                     # window on zone 1 (fenestrations come last)
    surfaces["is_fenestration"][:, 9] = 1
    surfaces["back_zone"][:, 9] = 1
    normals = np.array([(0, -1, 0), (1, 0, 0), (0, 1, 0), (0, 0, 1)] * 2 + [(1, 0, 0), (0, -1, 0)], float)
    surfaces["normal"] = normals[None]
    area = rng.uniform(n, 6.0, 20.0)
    surfaces["area"] = area[:, None]
    surfaces["area"][:, 9] = 0.25 * area
    surfaces["perimeter"] = 4.0 * np.sqrt(surfaces["area"])
    surfaces["centroid_z"] = 1.5
    for f in ("front_hs_override", "back_hs_override", "front_emissivity_override", "back_emissivity_override"):
        surfaces[f] = D.NAN
    zo = (np.arange(n + 1, dtype=np.int64) * 2).astype(np.int32)
    so = (np.arange(n + 1, dtype=np.int64) * NSURF).astype(np.int32)
    return D.BatchDesc(layers, off.astype(np.int32), zo, so, volume, surfaces.reshape(-1), steps_per_hour), gains


def synthetic_weather(n_rows: int, seed: int = 7):
    """Smooth annual + diurnal series: returns weather[n_rows][3] and inputs[n_rows][4]
    (front solar, back solar, front IR, back IR), hourly rows."""
    t = np.arange(n_rows, dtype=np.float64)
    day = 2.0 * math.pi * t / 24.0
    year = 2.0 * math.pi * t / 8760.0
    t_out = 15.0 - 10.0 * np.cos(year) - 5.0 * np.cos(day - 0.6)
    wind_speed = 3.0 + 3.0 * np.sin(year * 3.0 + 1.0) * np.cos(day * 0.5)
    wind_dir = np.mod(t * 7.3 + 40.0 * np.sin(day), 360.0)
    solar = np.maximum(0.0, 800.0 * np.sin(day - math.pi / 2.0)) * (0.6 + 0.4 * np.cos(year - math.pi))
    ir = 5.670374419e-8 * (t_out + 263.15) ** 4
    weather = np.stack([t_out, wind_dir, np.abs(wind_speed)], axis=1)
    inputs = np.stack([solar, np.zeros_like(solar), ir, ir], axis=1)
    return np.ascontiguousarray(weather), np.ascontiguousarray(inputs)

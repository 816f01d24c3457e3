"""Flat model descriptions shared by the C ABI (include/heat_b200.h) and the host mirror.

The reference reads these values out of `simple_model::SimpleModel` (un-vendored crate): substances and
materials are merged into *layers*, constructions are ranges of layers, spaces are *zones*.  Numpy structured
dtypes below are bit-compatible with `hb_layer` / `hb_surface`.
"""
from __future__ import annotations

import dataclasses
from typing import Iterable, List, Optional, Sequence

import numpy as np

NAN = float("nan")

# enums of include/heat_b200.h
BOUNDARY_OUTDOOR, BOUNDARY_SPACE, BOUNDARY_AMBIENT, BOUNDARY_GROUND = 0, 1, 2, 3
LAYER_SOLID, LAYER_GAS = 0, 1
GAS_AIR, GAS_ARGON, GAS_KRYPTON, GAS_XENON = 0, 1, 2, 3
TERRAIN_NONE, TERRAIN_COUNTRY, TERRAIN_SUBURBS, TERRAIN_CITY, TERRAIN_OCEAN, TERRAIN_URBAN, TERRAIN_UNSET = (
    -1, 0, 1, 2, 3, 4, 5)

LAYER_DTYPE = np.dtype(
    [
        ("kind", "<i4"), ("gas", "<i4"),
        ("thickness", "<f8"), ("conductivity", "<f8"), ("density", "<f8"), ("specific_heat", "<f8"),
        ("front_emissivity", "<f8"), ("back_emissivity", "<f8"),
        ("front_solar_abs", "<f8"), ("back_solar_abs", "<f8"), ("solar_transmittance", "<f8"),
    ],
    align=True,
)
SURFACE_DTYPE = np.dtype(
    [
        ("construction", "<i4"), ("is_fenestration", "<i4"),
        ("front_kind", "<i4"), ("back_kind", "<i4"), ("front_zone", "<i4"), ("back_zone", "<i4"),
        ("front_ambient_t", "<f8"), ("back_ambient_t", "<f8"),
        ("area", "<f8"), ("perimeter", "<f8"), ("centroid_z", "<f8"), ("normal", "<f8", (3,)),
        ("front_hs_override", "<f8"), ("back_hs_override", "<f8"),
        ("front_emissivity_override", "<f8"), ("back_emissivity_override", "<f8"),
    ],
    align=True,
)
assert LAYER_DTYPE.itemsize == 80 and SURFACE_DTYPE.itemsize == 120


def solid_layer(thickness, conductivity, density, specific_heat, *, front_emissivity=NAN, back_emissivity=NAN,
                front_solar_abs=NAN, back_solar_abs=NAN, solar_transmittance=NAN):
    """A Material on a Normal substance. NaN = property not set (reference defaults apply)."""
    r = np.zeros((), LAYER_DTYPE)
    r["kind"] = LAYER_SOLID
    r["thickness"], r["conductivity"], r["density"], r["specific_heat"] = thickness, conductivity, density, specific_heat
    r["front_emissivity"], r["back_emissivity"] = front_emissivity, back_emissivity
    r["front_solar_abs"], r["back_solar_abs"] = front_solar_abs, back_solar_abs
    r["solar_transmittance"] = solar_transmittance
    return r


def gas_layer(thickness, gas=GAS_AIR):
    r = np.zeros((), LAYER_DTYPE)
    r["kind"], r["gas"], r["thickness"] = LAYER_GAS, gas, thickness
    for f in ("front_emissivity", "back_emissivity", "front_solar_abs", "back_solar_abs", "solar_transmittance"):
        r[f] = NAN
    return r


def surface(construction, area, perimeter, normal, centroid_z, *, front=(BOUNDARY_OUTDOOR, 0),
            back=(BOUNDARY_OUTDOOR, 0), is_fenestration=False, front_hs=NAN, back_hs=NAN,
            front_emissivity=NAN, back_emissivity=NAN):
    """`front`/`back` = (BOUNDARY_OUTDOOR, _) | (BOUNDARY_SPACE, zone index) | (BOUNDARY_AMBIENT, temperature)."""
    r = np.zeros((), SURFACE_DTYPE)
    r["construction"], r["is_fenestration"] = construction, int(is_fenestration)
    r["area"], r["perimeter"], r["centroid_z"], r["normal"] = area, perimeter, centroid_z, normal
    for side, (kind, val) in (("front", front), ("back", back)):
        r[side + "_kind"] = kind
        r[side + "_zone"] = int(val) if kind == BOUNDARY_SPACE else -1
        r[side + "_ambient_t"] = float(val) if kind == BOUNDARY_AMBIENT else 0.0
    r["front_hs_override"], r["back_hs_override"] = front_hs, back_hs
    r["front_emissivity_override"], r["back_emissivity_override"] = front_emissivity, back_emissivity
    return r


@dataclasses.dataclass
class BatchDesc:
    """A batch of independent models over one pool of constructions (mirrors `hb_batch_desc`)."""

    layers: np.ndarray                # LAYER_DTYPE [n_layers]
    construction_offset: np.ndarray   # int32 [n_constructions + 1]
    model_zone_offset: np.ndarray     # int32 [n_models + 1]
    model_surface_offset: np.ndarray  # int32 [n_models + 1]
    zone_volume: np.ndarray           # f8 [total zones]
    surfaces: np.ndarray              # SURFACE_DTYPE [total surfaces]
    steps_per_hour: int
    model_terrain: Optional[np.ndarray] = None  # int32 [n_models]

    def __post_init__(self):
        self.layers = np.ascontiguousarray(self.layers, LAYER_DTYPE)
        self.surfaces = np.ascontiguousarray(self.surfaces, SURFACE_DTYPE)
        self.construction_offset = np.ascontiguousarray(self.construction_offset, np.int32)
        self.model_zone_offset = np.ascontiguousarray(self.model_zone_offset, np.int32)
        self.model_surface_offset = np.ascontiguousarray(self.model_surface_offset, np.int32)
        self.zone_volume = np.ascontiguousarray(self.zone_volume, np.float64)
        if self.model_terrain is not None:
            self.model_terrain = np.ascontiguousarray(self.model_terrain, np.int32)

    @property
    def n_models(self) -> int:
        return len(self.model_zone_offset) - 1

    @property
    def n_constructions(self) -> int:
        return len(self.construction_offset) - 1

    def slice_models(self, m0: int, m1: int) -> "BatchDesc":
        """Models [m0, m1) as their own batch (a rank's shard); the constructions pool is shared, not copied."""
        z0, z1 = int(self.model_zone_offset[m0]), int(self.model_zone_offset[m1])
        s0, s1 = int(self.model_surface_offset[m0]), int(self.model_surface_offset[m1])
        return BatchDesc(self.layers, self.construction_offset, self.model_zone_offset[m0:m1 + 1] - z0,
                         self.model_surface_offset[m0:m1 + 1] - s0, self.zone_volume[z0:z1], self.surfaces[s0:s1],
                         self.steps_per_hour, None if self.model_terrain is None else self.model_terrain[m0:m1])

    def select_models(self, idx) -> "BatchDesc":
        """The models `idx` (any order, any subset) as their own batch; the constructions pool is shared, not copied."""
        idx = np.asarray(idx, np.int64)
        zo, so = self.model_zone_offset.astype(np.int64), self.model_surface_offset.astype(np.int64)
        nz, ns = zo[idx + 1] - zo[idx], so[idx + 1] - so[idx]

        def gather(off, cnt):
            total = int(cnt.sum())
            if total == 0:
                return np.zeros(0, np.int64)
            start = np.repeat(off[idx], cnt)
            within = np.arange(total) - np.repeat(np.cumsum(cnt) - cnt, cnt)
            return start + within

        zsel, ssel = gather(zo, nz), gather(so, ns)
        return BatchDesc(self.layers, self.construction_offset, np.concatenate([[0], np.cumsum(nz)]),
                         np.concatenate([[0], np.cumsum(ns)]), self.zone_volume[zsel], self.surfaces[ssel],
                         self.steps_per_hour, None if self.model_terrain is None else self.model_terrain[idx])

    def model(self, m: int) -> "BatchDesc":
        """The single-model batch holding model `m` (constructions pool shared)."""
        z0, z1 = self.model_zone_offset[m], self.model_zone_offset[m + 1]
        s0, s1 = self.model_surface_offset[m], self.model_surface_offset[m + 1]
        return BatchDesc(self.layers, self.construction_offset, [0, z1 - z0], [0, s1 - s0], self.zone_volume[z0:z1],
                         self.surfaces[s0:s1], self.steps_per_hour,
                         None if self.model_terrain is None else self.model_terrain[m:m + 1])


def single_model(constructions: Sequence[Sequence[np.ndarray]], zone_volumes: Iterable[float],
                 surfaces: Sequence[np.ndarray], steps_per_hour: int, terrain: int = TERRAIN_NONE) -> BatchDesc:
    """One model: `constructions` = list of layer lists; `surfaces` = surfaces first, then fenestrations."""
    layers: List[np.ndarray] = []
    off = [0]
    for c in constructions:
        layers.extend(c)
        off.append(len(layers))
    zv = list(zone_volumes)
    return BatchDesc(np.array(layers, LAYER_DTYPE), off, [0, len(zv)], [0, len(surfaces)], zv,
                     np.array(surfaces, SURFACE_DTYPE), steps_per_hour, np.array([terrain], np.int32))

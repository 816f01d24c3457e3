"""hb_reduce_noalloc — the K3 reduction with its scratch in a static device array (no cudaMalloc in the call) — against
hb_reduce, through the C ABI.  Collected last on purpose: the entry point has no counterpart in the reference (it was
added) and a failure here must not hide a parity test under `pytest -x`."""
import ctypes as C

import numpy as np
import pytest

from heat_b200 import HeatB200Error, _lib, cases

pytestmark = pytest.mark.gpu

SUM_FIELDS = ("sum_zone_t", "sum_node_t", "front_conv_energy_j", "back_conv_energy_j")
EXACT_FIELDS = ("min_zone_t", "max_zone_t", "min_node_t", "max_node_t", "n_zones", "n_nodes", "nomass_iterations",
                "status_flags")
SUM_RTOL = 1e-12   # hb_reduce adds with atomics in launch order (tests/test_parity_gpu.py uses the same bound)


def same_stats(got, want):
    for f in EXACT_FIELDS:
        assert getattr(got, f) == getattr(want, f), (f, getattr(got, f), getattr(want, f))
    for f in SUM_FIELDS:
        a, b = getattr(got, f), getattr(want, f)
        assert abs(a - b) <= SUM_RTOL * max(abs(a), abs(b), 1.0), (f, a, b)


def words(st):
    return bytes(C.string_at(C.addressof(st), C.sizeof(st)))


def marched(desc, gains=None, hours=6):
    from heat_b200.model import ThermalBatch

    weather, inputs = cases.synthetic_weather(2 * hours + 1)
    b = ThermalBatch(desc)
    if gains is not None:
        b.set_zone_inputs(q_internal=gains)
    b.march_series(hours, weather, weather_mode=_lib.WEATHER_LINEAR, inputs=inputs[:hours], input_mode=_lib.INPUT_SITE)
    return b, weather, inputs


def test_reduce_noalloc_matches_reduce_on_two_calls_in_a_row():
    hours = 6
    # warp tiles of every geometry, with and without no-mass ends; then buildings (zones shared by seven surfaces)
    b, weather, inputs = marched(cases.wall_ensemble(3000, seed=17, mixed_fraction=0.4), hours=hours)
    first = b.reduce_noalloc()
    same_stats(first, b.reduce())
    assert first.n_zones == 3000 and first.n_nodes == b.nodes and first.nomass_iterations > 0
    # second call in a row, on a state that moved on: the scratch still holds the first call's partials
    b.march_series(hours, weather[hours:], weather_mode=_lib.WEATHER_LINEAR, inputs=inputs[hours:2 * hours],
                   input_mode=_lib.INPUT_SITE)
    second = b.reduce_noalloc()
    same_stats(second, b.reduce())
    assert second.sum_node_t != first.sum_node_t
    # folded in tile order: the same state gives the same bits
    assert words(b.reduce_noalloc()) == words(second)

    # a smaller batch of another kind right after: partials of the larger one beyond its tiles are never read
    d2, gains = cases.shoebox_ensemble(40, seed=5, steps_per_hour=2)
    b2, _, _ = marched(d2, gains)
    same_stats(b2.reduce_noalloc(), b2.reduce())
    same_stats(b.reduce_noalloc(), second)   # and back


class Cudart:
    """The CALLER's device memory for `out_dev`, through libcudart by ctypes (no torch here: the test must not depend on
    which NCCL an earlier test left in the process)."""

    def __init__(self):
        self.rt = C.CDLL("libcudart.so.12")
        for f in (self.rt.cudaMalloc, self.rt.cudaMemcpy, self.rt.cudaFree, self.rt.cudaSetDevice, self.rt.cudaMemset):
            f.restype = C.c_int

    def ok(self, rc):
        assert rc == 0, f"cudart error {rc}"

    def malloc(self, nbytes):
        self.ok(self.rt.cudaSetDevice(0))
        p = C.c_void_p()
        self.ok(self.rt.cudaMalloc(C.byref(p), C.c_size_t(nbytes)))
        self.ok(self.rt.cudaMemset(p, 0, C.c_size_t(nbytes)))
        return p

    def to_host(self, p, nbytes):
        buf = C.create_string_buffer(nbytes)
        self.ok(self.rt.cudaMemcpy(buf, p, C.c_size_t(nbytes), 2))   # cudaMemcpyDeviceToHost
        return buf.raw

    def free(self, p):
        self.ok(self.rt.cudaFree(p))


def test_reduce_noalloc_device_output_and_errors():
    b, _, _ = marched(cases.wall_ensemble(500, seed=23))
    rt = Cudart()
    out_dev = rt.malloc(96)   # allocated by the caller, before the calls: the entry point itself allocates nothing
    assert b.reduce_noalloc(out_dev=out_dev.value, host=False) is None   # no synchronize in this form …
    want = b.reduce_noalloc()                                            # … this one waits for the handle's stream
    assert rt.to_host(out_dev, 96) == words(want)
    both = b.reduce_noalloc(out_dev=out_dev.value)
    assert words(both) == words(want) and rt.to_host(out_dev, 96) == words(want)
    same_stats(want, b.reduce())
    rt.free(out_dev)
    with pytest.raises(HeatB200Error, match="both outputs are NULL"):
        b.reduce_noalloc(host=False)


def test_reduce_noalloc_rejects_batches_beyond_the_static_scratch():
    from heat_b200.model import ThermalBatch

    b = ThermalBatch(cases.wall_ensemble(400_000, seed=3))   # 19 045 tiles > HB_REDUCE_NOALLOC_MAX_TILES
    with pytest.raises(HeatB200Error, match="HB_REDUCE_NOALLOC_MAX_TILES = 16384") as e:
        b.reduce_noalloc()
    assert e.value.status == 1   # HB_ERR_INVALID
    assert b.reduce().n_zones == 400_000   # the allocating entry point still serves it
    b.close()

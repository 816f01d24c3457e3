"""N > 1 host logic on CPU: two `gloo` ranks shard an ensemble by contiguous model ranges, each marches its shard
(kernel data flow re-enacted by tests/emu — no GPU here), and the ensemble statistics are combined with the same
`heat_b200.dist.combine_stats` all-reduce bench.py uses over NCCL.  The union must equal the single-process result.
"""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
for p in (ROOT, HERE, os.path.join(HERE, "emu"), os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

from heat_b200 import cases, dist as D  # noqa: E402

N_WALLS, N_STEPS = 14, 6


def _layout(desc):
    """(zone slice, node slices) per model from the state layout of include/heat_b200.h."""
    from heat_b200.model import ThermalBatch
    plan = ThermalBatch(desc, device=-1)
    out = []
    for m in range(desc.n_models):
        mi = plan.model_info(m)
        nodes = []
        for s in range(mi.surface_begin, mi.surface_begin + mi.n_surfaces):
            si = plan.surface_info(s)
            nodes.append(slice(si.state_offset + 8, si.state_offset + 8 + si.n_nodes))
        out.append((slice(mi.state_offset, mi.state_offset + mi.n_zones), nodes))
    plan.close()
    return out


def _march(desc):
    import emu_py
    weather, inputs = cases.synthetic_weather(N_STEPS + 1)
    b = emu_py.EmuBatch(desc)
    b.march_series(N_STEPS, weather, weather_mode=1, inputs=inputs[:N_STEPS], input_mode=1)
    return b.get_state()


def _worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        full = cases.wall_ensemble(N_WALLS, seed=5, mixed_fraction=0.4)
        m0, m1 = D.shard_bounds(full.n_models, rank, world)
        shard = full.slice_models(m0, m1)
        state = _march(shard)
        local = D.stats_from_state(state, _layout(shard))
        local.status_flags = 1 if rank == 0 else 8   # different status bits per rank: every rank must end with the OR
        total = D.combine_stats(local)
        t = D.max_over_ranks([float(rank + 1)])
        u = D.sum_over_ranks([float(m1 - m0)])
        q.put((rank, m0, m1, state, total, t[0], u[0]))
    finally:
        dist.destroy_process_group()


def test_shard_bounds_cover_everything():
    for n in (0, 1, 7, 8, 100_003):
        for world in (1, 2, 3, 8):
            b = [D.shard_bounds(n, r, world) for r in range(world)]
            assert b[0][0] == 0 and b[-1][1] == n
            assert all(b[i][1] == b[i + 1][0] for i in range(world - 1))
            sizes = [x[1] - x[0] for x in b]
            assert max(sizes) - min(sizes) <= 1


def test_deal_models_balances_tails_and_work():
    rng = np.random.default_rng(3)
    S = rng.choice([2, 4, 10, 20, 58, 118], size=10_000, p=[.3, .3, .2, .15, .04, .01])
    N = rng.integers(2, 128, size=10_000)
    for world in (2, 4, 8):
        parts = [D.deal_models(S, N, r, world) for r in range(world)]
        allidx = np.sort(np.concatenate(parts))
        assert np.array_equal(allidx, np.arange(S.size))            # a partition
        work = np.array([(S[p] * N[p]).sum() for p in parts], float)
        assert work.max() / work.min() < 1.01
        tails = [int((S[p] == 118).sum()) for p in parts]
        assert max(tails) - min(tails) <= 1


def test_deal_models_by_kind():
    """`kind` (e.g. walls with no-mass ends: other kernels) is the first sort key: every rank gets the same number (± 1)
    of models of every kind in every sub-step class, and the partition / work balance hold as without it."""
    rng = np.random.default_rng(4)
    S = rng.choice([18, 54, 118], size=8_000, p=[.8, .15, .05])
    N = rng.integers(2, 128, size=8_000)
    K = (rng.random(8_000) < 0.02).astype(np.int64)
    for world in (2, 8):
        parts = [D.deal_models(S, N, r, world, kind=K) for r in range(world)]
        assert np.array_equal(np.sort(np.concatenate(parts)), np.arange(S.size))
        for s in (18, 54, 118):
            counts = [int(((S[p] == s) & (K[p] == 1)).sum()) for p in parts]
            assert max(counts) - min(counts) <= 1, (world, s, counts)
        work = np.array([(S[p] * N[p]).sum() for p in parts], float)
        assert work.max() / work.min() < 1.02


def test_two_rank_gloo_equals_single_process():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = sorted((q.get(timeout=300) for _ in procs), key=lambda x: x[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0

    full = cases.wall_ensemble(N_WALLS, seed=5, mixed_fraction=0.4)
    state_full = _march(full)
    ref = D.stats_from_state(state_full, _layout(full))
    # shards are contiguous and their states concatenate to the single-process state (models are independent)
    assert got[0][1] == 0 and got[0][2] == got[1][1] and got[1][2] == N_WALLS
    assert np.array_equal(np.concatenate([got[0][3], got[1][3]]), state_full)
    for _, _, _, _, total, tmax, usum in got:       # every rank holds the combined statistics
        assert total.n_zones == ref.n_zones and total.n_nodes == ref.n_nodes
        assert total.min_node_t == ref.min_node_t and total.max_node_t == ref.max_node_t
        assert total.min_zone_t == ref.min_zone_t and total.max_zone_t == ref.max_zone_t
        assert abs(total.sum_node_t - ref.sum_node_t) <= 1e-9 * abs(ref.sum_node_t)
        assert abs(total.sum_zone_t - ref.sum_zone_t) <= 1e-9 * abs(ref.sum_zone_t)
        assert tmax == 2.0 and usum == float(N_WALLS)
        assert total.status_flags == 9

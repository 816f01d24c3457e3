"""Multi-GPU plumbing of the batched entry point: one process per GPU, the ensemble's models are sharded
by contiguous index ranges, no exchange during the march; the only collective is the final all-reduce of the ensemble
statistics (`hb_reduce` per rank → sum / min / max across ranks; NCCL on GPUs, gloo in the CPU tests).
"""
from __future__ import annotations

import dataclasses
from typing import Optional, Tuple

import numpy as np


def shard_bounds(n_models: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous, balanced range [m0, m1) of models owned by `rank` (the first n % world ranks get one more)."""
    if not (0 <= rank < world):
        raise ValueError("rank out of range")
    base, extra = divmod(n_models, world)
    m0 = rank * base + min(rank, extra)
    return m0, m0 + base + (1 if rank < extra else 0)


def deal_models(substeps, nodes, rank: int, world: int, kind=None) -> np.ndarray:
    """Indices of the models `rank` marches when ONE global ensemble is split over `world` GPUs.

    A rank's time is set as much by its longest dependent chains (sub-steps per main step x horizon, in sequence) as by
    its total work, so contiguous ranges of the sorted ensemble would hand every long chain to rank 0.  The models are
    sorted by (sub-steps, nodes), both descending, and dealt `world` at a time, the heaviest of each hand to the rank
    with the least work so far: every rank gets the same share of every part of the (S, N) distribution — equal tails
    and equal sums of N·S (longest-processing-time rule, one hand at a time).

    `kind` (optional, one small integer per model): what else decides which kernel a model goes to — e.g. 1 for walls
    with a no-mass end.  It becomes the first sort key, so that every rank also gets the same number (± 1) of models of
    every kind in every sub-step class, hence the same launch groups with the same numbers of tiles."""
    if not (0 <= rank < world):
        raise ValueError("rank out of range")
    substeps, nodes = np.asarray(substeps), np.asarray(nodes)
    kind = np.zeros(substeps.shape, np.int64) if kind is None else np.asarray(kind)
    order = np.lexsort((-nodes, -substeps, -kind))
    work = (substeps * nodes).astype(np.float64)[order]
    load = np.zeros(world)
    mine = []
    for i in range(0, order.size, world):
        hand = work[i:i + world]                       # already heaviest first (sorted by S, then N)
        ranks = np.argsort(load, kind="stable")[:hand.size]
        load[ranks] += hand
        hit = np.nonzero(ranks == rank)[0]
        if hit.size:
            mine.append(order[i + hit[0]])
    return np.sort(np.asarray(mine, np.int64))


@dataclasses.dataclass
class EnsembleStats:
    """Mirror of `hb_ensemble_stats` (include/heat_b200.h) as plain Python numbers."""

    sum_zone_t: float
    min_zone_t: float
    max_zone_t: float
    sum_node_t: float
    min_node_t: float
    max_node_t: float
    front_conv_energy_j: float
    back_conv_energy_j: float
    n_zones: int
    n_nodes: int
    nomass_iterations: int = 0
    status_flags: int = 0

    @classmethod
    def from_c(cls, s) -> "EnsembleStats":
        return cls(*(getattr(s, f.name) for f in dataclasses.fields(cls)))

    @property
    def mean_zone_t(self) -> float:
        return self.sum_zone_t / self.n_zones

    @property
    def mean_node_t(self) -> float:
        return self.sum_node_t / self.n_nodes


STATUS_BITS = 8   # ST_NAN_H, ST_NAN_ERR, ST_NAN_ZONE, ST_ITER_CAP, ST_BAD_GAMMA (hb_dev.hpp) + spare


def combine_stats(local: EnsembleStats, device: Optional[str] = None, group=None) -> EnsembleStats:
    """All-reduce the per-rank statistics over `torch.distributed` (sum / min / max / bitwise-or of the status words).

    Works with any initialised backend: tensors are placed on `device` ("cuda" for NCCL, None/"cpu" for gloo).
    Without an initialised process group the local statistics are returned unchanged.
    """
    import torch
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()):
        return local
    dev = torch.device(device) if device else torch.device("cpu")
    sums = torch.tensor([local.sum_zone_t, local.sum_node_t, local.front_conv_energy_j, local.back_conv_energy_j,
                         float(local.n_zones), float(local.n_nodes), float(local.nomass_iterations)],
                        dtype=torch.float64, device=dev)
    mins = torch.tensor([local.min_zone_t, local.min_node_t], dtype=torch.float64, device=dev)
    # the status words are OR-ed bit by bit: NCCL has no bitwise-or reduction, a MAX per bit is the same thing
    bits = [float((int(local.status_flags) >> k) & 1) for k in range(STATUS_BITS)]
    maxs = torch.tensor([local.max_zone_t, local.max_node_t] + bits, dtype=torch.float64, device=dev)
    dist.all_reduce(sums, op=dist.ReduceOp.SUM, group=group)
    dist.all_reduce(mins, op=dist.ReduceOp.MIN, group=group)
    dist.all_reduce(maxs, op=dist.ReduceOp.MAX, group=group)
    s, mn, mx = sums.tolist(), mins.tolist(), maxs.tolist()
    flags = sum(1 << k for k in range(STATUS_BITS) if mx[2 + k] > 0.5)
    return EnsembleStats(s[0], mn[0], mx[0], s[1], mn[1], mx[1], s[2], s[3], int(round(s[4])), int(round(s[5])),
                         int(round(s[6])), flags)


def max_over_ranks(values, device: Optional[str] = None, group=None):
    """Element-wise max of a short list of floats over all ranks (timings are reported as the slowest rank's)."""
    import torch
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()):
        return list(values)
    t = torch.tensor(list(values), dtype=torch.float64, device=torch.device(device) if device else torch.device("cpu"))
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return t.tolist()


def sum_over_ranks(values, device: Optional[str] = None, group=None):
    import torch
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()):
        return list(values)
    t = torch.tensor(list(values), dtype=torch.float64, device=torch.device(device) if device else torch.device("cpu"))
    dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    return t.tolist()


def stats_from_state(state: np.ndarray, layout) -> EnsembleStats:
    """Statistics of a reference-layout state vector computed on the host (tests: what hb_reduce must agree with).

    `layout`: iterable of (zone_slice, [node_slice, ...]) per model.
    """
    zs, ns = [], []
    for zsl, nodes in layout:
        zs.append(state[zsl])
        ns.extend(state[sl] for sl in nodes)
    z = np.concatenate(zs) if zs else np.zeros(0)
    n = np.concatenate(ns) if ns else np.zeros(0)
    return EnsembleStats(float(z.sum()), float(z.min()), float(z.max()), float(n.sum()), float(n.min()), float(n.max()),
                         0.0, 0.0, int(z.size), int(n.size))

#!/usr/bin/env python
"""bench.py — wall node-substep updates/s of the wall-conduction time-stepping path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload walls|buildings|mixed|buildings2g]

A "step" is one pass of the hot path over one batch of synthetic input: the whole ensemble marched over the whole
horizon (default: BASELINE.json configs[2] — 10^5 independent massive walls with randomised layer stacks, each with
its own zone, one year of hourly synthetic weather, n = 1 step/hour) in ONE persistent kernel launch per register
tier.  Multi-GPU (torchrun, one rank per GPU) is weak scaling: every rank marches its own shard of the same size;
the only collective is the final NCCL all-reduce of the ensemble energy/temperature statistics.

Printed JSON keys follow the driver contract; `roofline` is against the FP64 pipe (the binding bound with node state
resident on chip; peak measured live with a DFMA probe because MEASURED_PEAKS.json has no FP64 entry) and
`roofline_hbm` gives the HBM side against MEASURED_PEAKS.json.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

# the launch groups of a series run on streams of their own: give each a hardware work queue (read at context creation)
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "oracle")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

FLOP_PER_UPDATE = 37.0          # RK4 tridiagonal arithmetic per node-substep (surface.rs:268-305)
BYTES_PER_NODE_LAUNCH = 40.0    # (3 diagonals + T in + T out)·8 B per node per launch, state resident in between
BYTES_PER_SURFACE_LAUNCH = 8 * 8.0 * 2


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="walls", choices=["walls", "buildings", "mixed", "buildings2g"])
    ap.add_argument("--units", type=int, default=0, help="walls / buildings per GPU (0 = the config's size)")
    ap.add_argument("--hours", type=int, default=8760, help="main steps per bench step (n = 1/h)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


def make_workload(kind: str, units: int, rank: int):
    from heat_b200 import cases
    if kind == "walls":       # configs[2]
        n = units or 100_000
        return cases.wall_ensemble(n, seed=42 + 1000 * rank), None, f"{n} massive walls + own zone (configs[2]), seed 42"
    if kind == "mixed":       # configs[4] per-GPU shard
        n = units or 1_250_000
        return (cases.wall_ensemble(n, seed=44 + 1000 * rank, mixed_fraction=0.5), None,
                f"{n} walls 50% massive / 50% poly-X-poly (configs[4] shard), seed 44")
    if kind == "buildings2g":  # probe: configs[3] topology with pane | argon | pane windows (general kernel)
        n = units or 20_000
        d, gains = cases.shoebox_ensemble(n, seed=43 + 1000 * rank, glazing=cases.double_glazing())
        return d, gains, f"{n} shoebox buildings, 6 surfaces + double-glazed window (probe), seed 43"
    n = units or 125_000      # configs[3] per-GPU shard of 10^6 buildings / 8
    d, gains = cases.shoebox_ensemble(n, seed=43 + 1000 * rank)
    return d, gains, f"{n} shoebox buildings, 6 surfaces + window (configs[3] shard), seed 43"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe)."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, n_gpus: int):
        """One sampler for the GPUs 0..n_gpus-1 of the box (run by rank 0 only: one NVML client, not one per rank)."""
        self.rows, self.proc, self.n_gpus = [], None, n_gpus

    def start(self):
        try:
            ids = ",".join(str(i) for i in range(self.n_gpus))
            self.proc = subprocess.Popen(["nvidia-smi", "-i", ids, f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except subprocess.TimeoutExpired:
                self.proc.kill()
        sm, mx, reasons = [], 0.0, set()
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx = max(mx, float(r[1]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons),
                "samples": len(sm)}


def cpu_reference_rate(desc, weather, inputs, n_models: int, hours: int, threads: int):
    """The oracle (kind "port": faithful C++ restatement of the reference) on `threads` host threads, one model per
    thread at a time — the reference's own CPU implementation cannot be built here (Rust, un-vendored deps)."""
    import oracle_py as O
    models = [O.OracleModel(desc, m) for m in range(n_models)]
    t0 = time.perf_counter()
    updates, _ = O.run_ensemble(models, hours, weather, inputs, weather_linear=True, threads=threads)
    dt = time.perf_counter() - t0
    return updates / dt, updates, dt


CPU_SAMPLE_SECONDS = float(os.environ.get("HEAT_B200_CPU_SAMPLE_S", "20"))   # CPU work per bounded sample of the workload


def cpu_sample_size(desc, weather, inputs, hours: int, threads: int):
    """Number of leading models of the workload (x `h` hourly steps) that keep all host threads busy for about
    CPU_SAMPLE_SECONDS: a pilot of 2 models per thread sets the scale."""
    h = min(hours, 2190)
    n0 = min(desc.n_models, 2 * threads)
    _, _, dt0 = cpu_reference_rate(desc, weather, inputs, n0, h, threads)
    n = int(n0 * CPU_SAMPLE_SECONDS / max(dt0, 1e-3))
    n = max(n0, min(desc.n_models, (n // threads) * threads or threads))
    return n, h


def main():
    a = parse()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    from heat_b200 import cases

    hours = a.hours
    weather, inputs = cases.synthetic_weather(hours + 1)
    inputs = np.ascontiguousarray(inputs[:hours])
    cores = os.cpu_count() or 1

    # ---------------------------------------------------------------------------------------------------------
    if a.impl == "reference":
        if rank != 0:
            return
        desc, gains, wname = make_workload(a.workload, a.units, 0)
        n_sample, h_sample = cpu_sample_size(desc, weather, inputs, hours, cores)
        vals = []
        for it in range(a.warmup + a.steps):
            rate, updates, dt = cpu_reference_rate(desc, weather, inputs, n_sample, h_sample, cores)
            if it >= a.warmup:
                vals.append((rate, dt))
        value = float(np.mean([v[0] for v in vals]))
        sample = f"first {n_sample} models of the workload x {h_sample} hourly steps per bench step, {cores} threads"
        print(json.dumps({
            "impl": "reference", "metric": "wall node-substep updates/sec", "value": value, "unit": "updates/s",
            "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": float(np.mean([v[1] for v in vals]) * 1e3), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": wname, "hours": hours, "steps_per_hour": 1},
            "cpu_baseline": {"value": value, "unit": "updates/s", "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": "updates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        }))
        return

    # ---------------------------------------------------------------------------------------------------------
    from heat_b200 import _lib
    from heat_b200.model import ThermalBatch, measure_fp64_peak

    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        # NCCL prints its version banner on fd 1 when the communicator is created: keep stdout for the JSON line
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
            dist.barrier()
            torch.cuda.synchronize()
        finally:
            os.dup2(saved, 1)
            os.close(saved)

    desc, gains, wname = make_workload(a.workload, a.units, rank)
    batch = ThermalBatch(desc, device=local_rank)
    if gains is not None:
        batch.set_zone_inputs(q_internal=gains)
    updates_per_step = batch.updates_per_step * hours
    initial = batch.get_state()

    w_dev = batch.upload(weather)
    i_dev = batch.upload(inputs)

    def barrier():
        if dist is not None:
            dist.barrier()

    fp64_peak = measure_fp64_peak(local_rank)

    # ---- value: inputs resident in HBM, CUDA-event timing on the launching stream ----
    launches0 = None
    ms = []
    sampler = ClockSampler(world) if rank == 0 else None
    for it in range(a.warmup + a.steps):
        batch.set_state(initial)
        batch.flush_l2()
        barrier()
        if it == a.warmup:
            if sampler:
                sampler.start()
            launches0 = batch.launch_count
        t = batch.march_series_dev(hours, w_dev, hours + 1, 1, weather_mode=_lib.WEATHER_LINEAR, inputs_dev=i_dev,
                                   input_mode=_lib.INPUT_SITE)
        barrier()
        if it >= a.warmup:
            ms.append(t)
    gpu_launches = batch.launch_count - launches0
    clocks = sampler.stop() if sampler else None
    stats = batch.reduce()

    # ---- e2e: the public host-buffer API; H2D of the step's inputs and D2H of its result inside the timed region ----
    e2e_s = []
    for it in range(1 + a.steps):
        batch.set_state(initial)
        barrier()
        t0 = time.perf_counter()
        batch.march_series(hours, weather, weather_mode=_lib.WEATHER_LINEAR, inputs=inputs, input_mode=_lib.INPUT_SITE)
        st = batch.reduce()           # device→host read of the step's result (ensemble statistics)
        final = batch.get_state()     # … and of the full final state
        barrier()
        if it >= 1:
            e2e_s.append(time.perf_counter() - t0)
    h2d = weather.nbytes + inputs.nbytes
    d2h = final.nbytes + 12 * 8

    from heat_b200.dist import EnsembleStats, combine_stats, max_over_ranks, sum_over_ranks
    dev = "cuda" if dist is not None else None
    # slowest rank's time; units of all ranks; the path's only collective: the final ensemble statistics (NCCL)
    t_local = float(np.mean(ms)) * 1e-3
    by_rank = [0.0] * world
    by_rank[rank] = t_local * 1e3
    by_rank = sum_over_ranks(by_rank, dev)   # every rank's own device time, for the record
    t_step, t_e2e = max_over_ranks([t_local, float(np.mean(e2e_s))], dev)
    total_updates = sum_over_ranks([float(updates_per_step)], dev)[0]
    es = combine_stats(EnsembleStats.from_c(stats), dev)
    reduced = {"mean_zone_t": es.mean_zone_t, "mean_node_t": es.mean_node_t,
               "front_conv_energy_j": es.front_conv_energy_j, "back_conv_energy_j": es.back_conv_energy_j,
               "min_node_t": es.min_node_t, "max_node_t": es.max_node_t, "status_flags": es.status_flags}

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except OSError:
            pass
        traffic, traffic_src = None, None
        try:   # DRAM bytes per step measured once with ncu for the default workload (profiles/r1p_traffic.json)
            tj = json.load(open(os.path.join(ROOT, "profiles", "r1p_traffic.json")))
            if tj.get("workload") == wname and world == 1:
                traffic = int(tj["dram_bytes_read_per_step"]) + int(tj["dram_bytes_written_per_step"])
                traffic_src = "profiles/r1p_traffic.json (ncu dram__bytes_read+write.sum, all launches of one step)"
        except (OSError, ValueError, KeyError):
            pass
        hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
        hbm_src = "MEASURED_PEAKS.json" if "hbm_gbs" in peaks else "fallback 6.65 TB/s (B200_PROFILING.md)"
        value = total_updates / t_step
        per_gpu = updates_per_step / t_step
        fp64_achieved = FLOP_PER_UPDATE * per_gpu / 1e12
        alg_bytes = BYTES_PER_NODE_LAUNCH * batch.nodes + BYTES_PER_SURFACE_LAUNCH * batch.n_surfaces + h2d
        out = {
            "metric": "wall node-substep updates/sec", "value": value, "unit": "updates/s", "n_gpus": world,
            "steps": a.steps, "warmup": a.warmup, "ms_per_step": t_step * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": wname, "hours": hours, "steps_per_hour": 1, "weather": "linear between hourly rows",
                       "updates_per_step_per_gpu": int(updates_per_step), "nodes_per_gpu": int(batch.nodes),
                       "l2": "flushed between timed iterations (256 MiB write)",
                       "ms_per_step_by_rank": [round(x, 2) for x in by_rank], "ensemble_stats": reduced},
            "roofline": {"bound": "fp64", "achieved": fp64_achieved, "peak": fp64_peak, "unit": "TFLOP/s",
                         "frac": fp64_achieved / fp64_peak if fp64_peak else None, "traffic": traffic,
                         "traffic_source": traffic_src,
                         "flop_per_update": FLOP_PER_UPDATE,
                         "peak_source": "DFMA probe kernel run in this process (hb_measure_fp64_peak)"},
            "roofline_hbm": {"bound": "hbm", "achieved": alg_bytes / t_step / 1e9, "peak": hbm_peak, "unit": "GB/s",
                             "frac": alg_bytes / t_step / 1e9 / hbm_peak, "peak_source": hbm_src},
            "e2e": {"value": total_updates / t_e2e, "unit": "updates/s", "h2d_bytes_per_step": int(h2d),
                    "d2h_bytes_per_step": int(d2h)},
            "gpu_launches": int(gpu_launches), "clocks": clocks,
        }
        if world == 1 and not a.no_cpu_baseline:
            n_sample, h_sample = cpu_sample_size(desc, weather, inputs, hours, cores)
            rate, upd, dt = cpu_reference_rate(desc, weather, inputs, n_sample, h_sample, cores)
            out["cpu_baseline"] = {"value": rate, "unit": "updates/s", "cores": cores, "kind": "port",
                                   "sample": f"first {n_sample} models x {h_sample} hourly steps, {dt:.1f} s"}
        print(json.dumps(out))
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()

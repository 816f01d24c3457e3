#!/usr/bin/env python
"""bench.py — wall node-substep updates/s of the wall-conduction time-stepping path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--workload walls|buildings|mixed|buildings2g] [--units U] [--hours H] [--no-secondary]

A "step" is one pass of the hot path over one batch of synthetic input: the whole ensemble marched over the whole
horizon (default: BASELINE.json configs[2] — 10^5 independent massive walls with randomised layer stacks, each with
its own zone, one year of hourly synthetic weather, n = 1 step/hour) in ONE persistent kernel launch per geometry.
Multi-GPU (torchrun, one rank per GPU) is weak scaling: ONE global ensemble of N x the per-GPU size is dealt to the
ranks so that every rank gets the same share of long chains and of work (heat_b200.dist.deal_models); nothing is
exchanged while marching; the only collective is the final NCCL all-reduce of the ensemble statistics, done inside the
C ABI (hb_comm_init / hb_reduce_all).

Printed JSON keys follow the driver contract.  `roofline` is against the FP64 pipe (the binding bound with node state
resident on chip); `roofline_hbm` gives the HBM side against MEASURED_PEAKS.json.  `secondary` carries configs[3] and
configs[4] at their per-GPU sizes over a 240 h horizon; `drop_in` the single-model `ThermalModel.march` rate
(configs[0]) beside the CPU oracle's.
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
FP64_FMA_PER_CLK_PER_SM = 64    # B200: 64 DFMA lanes per SM and clock -> 128 FLOP/clk/SM

# per-GPU sizes of the BASELINE.json configs
WORKLOADS = {
    "walls": dict(units=100_000, seed=42, config=2, what="massive walls + own zone (configs[2])"),
    "buildings": dict(units=125_000, seed=43, config=3, what="shoebox buildings, 6 surfaces + window (configs[3] shard)"),
    "mixed": dict(units=1_250_000, seed=44, config=4, what="walls 50% massive / 50% poly-X-poly (configs[4] shard)"),
    "buildings2g": dict(units=20_000, seed=43, config=None, what="shoebox buildings, 6 surfaces + double-glazed window (probe)"),
}
CPU_N_SAMPLE, CPU_H_SAMPLE = 1920, 2190   # bounded CPU sample of a workload: models x hourly steps (≈ 20 s on 16 cores)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="walls", choices=sorted(WORKLOADS))
    ap.add_argument("--units", type=int, default=0, help="walls / buildings per GPU (0 = the config's size)")
    ap.add_argument("--hours", type=int, default=8760, help="main steps per bench step (n = 1/h)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-secondary", action="store_true", help="skip configs[3] / configs[4] / drop-in lines")
    return ap.parse_args()


def workload_name(kind: str, units: int) -> str:
    w = WORKLOADS[kind]
    return f"{units} {w['what']}, seed {w['seed']}"


def config_of(kind: str, units: int, hours: int, n_models: int) -> dict:
    """The `config` both arms print: the workload and the bounded sample the CPU legs march."""
    n_sample = min(n_models, CPU_N_SAMPLE)
    h_sample = min(hours, CPU_H_SAMPLE)
    return {"workload": workload_name(kind, units), "hours": hours, "steps_per_hour": 1,
            "weather": "linear between hourly rows", "l2": "flushed between timed iterations (256 MiB write)",
            "cpu_sample": {"n_sample": n_sample, "h_sample": h_sample,
                           "models": "every (n_models // n_sample)-th model of the one-GPU ensemble (base seed)",
                           "hours": "every (hours // h_sample)-th hourly row of the year"}}


def oracle_planner(desc):
    """Discretization of every wall by the CPU oracle (the reference arm must not load the product library)."""
    import ctypes as C
    import oracle_py as O
    n = desc.n_models
    from heat_b200 import cases
    n_nodes, n_massive, rk4 = np.zeros(n, np.int64), np.zeros(n, np.int64), np.zeros(n)
    mass, u, buf = np.zeros(1100), np.zeros(1100), np.zeros(2 * 1100, np.int32)
    for m in range(n):
        om = O.OracleModel(desc, m)
        k = n_nodes[m] = om.n_nodes(0)
        n_massive[m] = om._lib.ho_surface_chunks(om._m, 0, 1, buf.ctypes.data, buf.size)
        om._lib.ho_surface_segments(om._m, 0, mass.ctypes.data, u.ctypes.data)
        # dt·λmax <= 2·dt·max (left + right link) / mass (Gershgorin): only walls near the limit need the eigenvalue
        link = np.where(u[:k] < 0.0, 8.0, u[:k])
        row = (np.concatenate(([cases.FILM_MAX], link[:k - 1])) + np.concatenate((link[:k - 1], [cases.FILM_MAX]))) / mass[:k]
        rk4[m] = 2.0 * om.dt * float(row.max())
        if rk4[m] > 0.9 * cases.RK4_STABILITY_LIMIT:
            rk4[m] = cases.rk4_number_np(mass[:k], u[:k], buf[:2 * n_massive[m]], om.dt)
    return n_nodes, n_massive, rk4


def make_ensemble(kind: str, units: int, seed: int, planner=None):
    """(BatchDesc, gains or None) of `units` models of workload `kind`."""
    from heat_b200 import cases
    if kind == "walls":
        return cases.wall_ensemble(units, seed=seed, planner=planner), None
    if kind == "mixed":
        return cases.wall_ensemble(units, seed=seed, mixed_fraction=0.5, planner=planner), None
    if kind == "buildings2g":
        return cases.shoebox_ensemble(units, seed=seed, glazing=cases.double_glazing())
    return cases.shoebox_ensemble(units, seed=seed)


def shard_of(kind: str, units: int, rank: int, world: int, planner=None, plan_info=None):
    """This rank's share of the job's ensemble.

    walls / buildings: ONE global ensemble of world x units models (the config's generator and seed), dealt by
    (sub-steps, nodes) — heat_b200.dist.deal_models.  mixed (10^7 walls at 8 GPUs): generating the global ensemble on
    every rank would take minutes; each rank draws its own 1.25 M walls (seed + 1000·rank) — at that size the shards'
    (S, N) distributions agree to a fraction of a percent (profiles/r1u_mixed_10M_8gpu_240h.json: 0.3 % spread).
    """
    seed = WORKLOADS[kind]["seed"]
    if world == 1 or kind == "mixed":
        return make_ensemble(kind, units, seed + (1000 * rank if kind == "mixed" else 0), planner)
    from heat_b200 import dist
    d, gains = make_ensemble(kind, units * world, seed, planner)
    S, N, K = plan_info(d)
    idx = dist.deal_models(S, N, rank, world, kind=K)
    return d.select_models(idx), (None if gains is None else gains[idx])


def gpu_plan_info(desc):
    """(sub-steps, nodes, kind) per model from the library's host planning (hb_create with HB_DEVICE_NONE); kind: 1 for a
    one-wall model whose wall has no-mass chunks (it marches on other kernels than an all-massive wall), else 0."""
    import ctypes as C
    from heat_b200 import _lib
    from heat_b200.model import ThermalBatch
    plan = ThermalBatch(desc, device=-1)
    try:
        mi, si = _lib.ModelInfoC(), _lib.SurfaceInfoC()
        S, N = np.zeros(desc.n_models, np.int64), np.zeros(desc.n_models, np.int64)
        K = np.zeros(desc.n_models, np.int64)
        one_wall = int(desc.model_surface_offset[-1]) == desc.n_models
        for m in range(desc.n_models):
            plan._lib.hb_get_model_info(plan._h, m, C.byref(mi))
            S[m], N[m] = mi.dt_subdivisions, mi.nodes
            if one_wall:
                plan._lib.hb_get_surface_info(plan._h, m, C.byref(si))
                K[m] = 1 if si.n_nomass_chunks > 0 else 0
        return S, N, K
    finally:
        plan.close()


def cpu_sample(desc, weather, inputs, hours: int):
    """The bounded sample of a workload the CPU legs march: models and hourly rows spread over the ensemble / the year."""
    n_sample, h_sample = min(desc.n_models, CPU_N_SAMPLE), min(hours, CPU_H_SAMPLE)
    idx = np.arange(n_sample) * (desc.n_models // n_sample)
    stride = max(1, hours // h_sample)
    rows = np.arange(h_sample + 1) * stride
    rows[-1] = min(rows[-1], len(weather) - 1)
    return desc.select_models(idx), np.ascontiguousarray(weather[rows]), np.ascontiguousarray(inputs[rows[:-1]]), h_sample


def cpu_reference_rate(desc, weather, inputs, hours: int, threads: int, gains=None):
    """The oracle (kind "port": faithful C++ restatement of the reference) on `threads` host threads, one model per
    thread at a time — the reference's own CPU implementation cannot be built here (Rust, un-vendored deps)."""
    import oracle_py as O
    models = [O.OracleModel(desc, m) for m in range(desc.n_models)]
    t0 = time.perf_counter()
    updates, _ = O.run_ensemble(models, hours, weather, inputs, weather_linear=True, threads=threads, q_internal=gains)
    dt = time.perf_counter() - t0
    return updates / dt, updates, dt


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


class Job:
    """Rank / world of this process and the little torch.distributed plumbing the bench needs (barrier, max, sum)."""

    def __init__(self):
        self.rank = int(os.environ.get("RANK", "0"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.dist = None
        if self.world > 1:
            import torch
            import torch.distributed as dist
            torch.cuda.set_device(self.local_rank)
            # NCCL prints its version banner on fd 1 when the communicator is created: keep stdout for the JSON line
            sys.stdout.flush()
            saved = os.dup(1)
            os.dup2(2, 1)
            try:
                dist.init_process_group("nccl", device_id=torch.device("cuda", self.local_rank))
                dist.barrier()
                torch.cuda.synchronize()
            finally:
                os.dup2(saved, 1)
                os.close(saved)
            self.dist = dist
        self.dev = "cuda" if self.dist is not None else None

    def barrier(self):
        if self.dist is not None:
            self.dist.barrier()

    def unique_id(self, make) -> bytes:
        """The 128-byte NCCL id of the library's own communicator: made on rank 0, shipped with the job's plumbing."""
        box = [make() if self.rank == 0 else None]
        if self.dist is not None:
            sys.stdout.flush()
            saved = os.dup(1)
            os.dup2(2, 1)
            try:
                self.dist.broadcast_object_list(box, src=0)
            finally:
                os.dup2(saved, 1)
                os.close(saved)
        return box[0]

    def close(self):
        if self.dist is not None:
            self.dist.destroy_process_group()


def run_workload(job: Job, kind: str, units: int, hours: int, steps: int, warmup: int, sample_clocks: bool):
    """March this rank's shard of `kind` `warmup + steps` times; returns the rank-0 result dict (None on other ranks)."""
    from heat_b200 import _lib, cases
    from heat_b200.dist import max_over_ranks, sum_over_ranks
    from heat_b200.model import ThermalBatch

    weather, inputs = cases.synthetic_weather(hours + 1)
    inputs = np.ascontiguousarray(inputs[:hours])
    desc, gains = shard_of(kind, units, job.rank, job.world, plan_info=gpu_plan_info)
    batch = ThermalBatch(desc, device=job.local_rank)
    if gains is not None:
        batch.set_zone_inputs(q_internal=gains)
    if job.world > 1:   # the path's only collective lives inside the library
        batch.comm_init(job.unique_id(ThermalBatch.comm_unique_id), job.rank, job.world)
    updates_per_step = batch.updates_per_step * hours
    initial = batch.get_state()
    w_dev, i_dev = batch.upload(weather), batch.upload(inputs)

    # ---- value: inputs resident in HBM, CUDA-event timing on the launching stream ----
    launches0, ms = None, []
    sampler = ClockSampler(job.world) if (job.rank == 0 and sample_clocks) else None
    for it in range(warmup + steps):
        batch.set_state(initial)
        batch.flush_l2()
        job.barrier()
        if it == warmup:
            if sampler:
                sampler.start()
            launches0 = batch.launch_count
        t = batch.march_series_dev(hours, w_dev, hours + 1, 1, weather_mode=_lib.WEATHER_LINEAR, inputs_dev=i_dev,
                                   input_mode=_lib.INPUT_SITE)
        job.barrier()
        if it >= warmup:
            ms.append(t)
    gpu_launches = batch.launch_count - launches0
    clocks = sampler.stop() if sampler else None
    stats = batch.reduce_all()        # local K3 reduction + NCCL all-reduce inside the C ABI
    if stats.status_flags:
        raise RuntimeError(f"physics status flags {stats.status_flags:#x} after the timed marches: the run is poisoned")

    # ---- e2e: the public host-buffer API; H2D of the step's inputs and D2H of its result inside the timed region ----
    e2e_s = []
    for it in range(1 + max(1, steps - 1)):
        batch.set_state(initial)
        job.barrier()
        t0 = time.perf_counter()
        batch.march_series(hours, weather, weather_mode=_lib.WEATHER_LINEAR, inputs=inputs, input_mode=_lib.INPUT_SITE)
        batch.reduce_all()            # device→host read of the step's result (ensemble statistics)
        final = batch.get_state()     # … and of the full final state
        job.barrier()
        if it >= 1:
            e2e_s.append(time.perf_counter() - t0)
    h2d = weather.nbytes + inputs.nbytes
    d2h = final.nbytes + 12 * 8

    t_local = float(np.mean(ms)) * 1e-3
    by_rank = [0.0] * job.world
    by_rank[job.rank] = t_local * 1e3
    by_rank = sum_over_ranks(by_rank, job.dev)   # every rank's own device time, for the record
    t_step, t_e2e = max_over_ranks([t_local, float(np.mean(e2e_s))], job.dev)
    total_updates = sum_over_ranks([float(updates_per_step)], job.dev)[0]
    nodes, n_surfaces, n_models = batch.nodes, batch.n_surfaces, batch.n_models
    out = None
    if job.rank == 0:
        out = {
            "kind": kind, "desc": desc, "gains": gains, "weather": weather, "inputs": inputs, "n_models": n_models,
            "value": total_updates / t_step, "per_gpu": updates_per_step / t_step, "t_step": t_step,
            "e2e": {"value": total_updates / t_e2e, "unit": "updates/s", "h2d_bytes_per_step": int(h2d),
                    "d2h_bytes_per_step": int(d2h)},
            "alg_bytes": BYTES_PER_NODE_LAUNCH * nodes + BYTES_PER_SURFACE_LAUNCH * n_surfaces + h2d,
            "gpu_launches": int(gpu_launches), "clocks": clocks,
            "details": {"updates_per_step_per_gpu": int(updates_per_step), "nodes_per_gpu": int(nodes),
                        "models_per_gpu": int(n_models), "ms_per_step_by_rank": [round(x, 2) for x in by_rank],
                        "ensemble_stats": {
                            "mean_zone_t": stats.sum_zone_t / max(1, stats.n_zones),
                            "mean_node_t": stats.sum_node_t / max(1, stats.n_nodes),
                            "front_conv_energy_j": stats.front_conv_energy_j,
                            "back_conv_energy_j": stats.back_conv_energy_j, "min_node_t": stats.min_node_t,
                            "max_node_t": stats.max_node_t, "n_zones": int(stats.n_zones), "n_nodes": int(stats.n_nodes),
                            "status_flags": int(stats.status_flags),
                            "reduced_by": "hb_reduce_all (NCCL inside the C ABI)" if job.world > 1 else "hb_reduce"}},
        }
    batch.free_dev(w_dev)
    batch.free_dev(i_dev)
    batch.close()
    return out


def drop_in_rate(n_steps: int = 2000):
    """configs[0]: the drop-in `ThermalModel.march` (one call per main step: hb_set_state + one-step series +
    hb_get_state, caller-owned state) on massive_full over hourly steps, beside the CPU oracle's `march`."""
    import oracle_py as O
    from heat_b200 import cases, model as M
    d = cases.validation_case("massive_full", 1)
    weather, inputs = cases.synthetic_weather(n_steps + 1)
    header = M.SimulationStateHeader()
    tm = M.ThermalModel.new(None, (), d, header, 1)
    mem = tm.allocate_memory()
    state = header.take_values()
    date = M.Date(1, 1, 0.0)

    class Hourly:   # weather::Weather with linear interpolation between hourly rows
        def get_weather_data(self, dd):
            h = dd.hour_of_year()
            i = min(int(h), n_steps - 1)
            f = h - i
            r = weather[i] + (weather[i + 1] - weather[i]) * f
            return M.CurrentWeather(r[0], r[2], r[1])

    w = Hourly()
    t0 = time.perf_counter()
    for i in range(n_steps):
        state[tm.surface_slot(0, 4)] = inputs[i, 0]
        state[tm.surface_slot(0, 6)] = inputs[i, 2]
        tm.march(date, w, d, state, mem)
        date.add_hours(1.0)
    t_gpu = time.perf_counter() - t0
    om = O.OracleModel(d)
    so = om.initial_state()
    rows = np.empty((om.dt_subdivisions, 3))
    t0 = time.perf_counter()
    for i in range(n_steps):
        so[tm.surface_slot(0, 4)] = inputs[i, 0]
        so[tm.surface_slot(0, 6)] = inputs[i, 2]
        for s in range(om.dt_subdivisions):
            rows[s] = weather[i] + (weather[i + 1] - weather[i]) * ((s + 1) / om.dt_subdivisions)
        om.march(so, rows)
    t_cpu = time.perf_counter() - t0
    upd = n_steps * om.dt_subdivisions * om.n_nodes(0)
    return {"workload": "massive_full, one model, hourly steps, ThermalModel.march per step (configs[0])",
            "steps": n_steps, "gpu_steps_per_s": n_steps / t_gpu, "cpu_steps_per_s": n_steps / t_cpu,
            "gpu_updates_per_s": upd / t_gpu, "cpu_updates_per_s": upd / t_cpu,
            "max_abs_diff_vs_oracle": float(np.abs(state - so).max()),
            "note": "three host synchronisations per main step for ONE model: the drop-in call is latency-bound by "
                    "design; ensembles go through the batched entry point"}


def main():
    a = parse()
    job_rank = int(os.environ.get("RANK", "0"))
    kind = a.workload
    units = a.units or WORKLOADS[kind]["units"]
    hours = a.hours
    cores = os.cpu_count() or 1

    # ---------------------------------------------------------------------------------------------------------
    if a.impl == "reference":
        # The reference's own CPU implementation of the path (the oracle port: the Rust crate cannot be built here), all
        # host threads, on a bounded sample of the GPU arm's workload.  Rank 0 only; loads nothing but oracle/.
        if job_rank != 0:
            return
        from heat_b200 import cases
        weather, inputs = cases.synthetic_weather(hours + 1)
        planner = oracle_planner if kind in ("walls", "mixed") else None
        desc, gains = make_ensemble(kind, units, WORKLOADS[kind]["seed"], planner)
        sdesc, sw, si, h_sample = cpu_sample(desc, weather, inputs[:hours], hours)
        sgains = None if gains is None else gains[np.arange(sdesc.n_models) * (desc.n_models // sdesc.n_models)]
        vals = []
        for it in range(a.warmup + a.steps):
            rate, updates, dt = cpu_reference_rate(sdesc, sw, si, h_sample, cores, sgains)
            if it >= a.warmup:
                vals.append((rate, dt))
        value = float(np.mean([v[0] for v in vals]))
        sample = f"{sdesc.n_models} models x {h_sample} hourly steps per bench step (config.cpu_sample), {cores} threads"
        print(json.dumps({
            "impl": "reference", "metric": "wall node-substep updates/sec", "value": value, "unit": "updates/s",
            "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": float(np.mean([v[1] for v in vals]) * 1e3), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": config_of(kind, units, hours, desc.n_models),
            "cpu_baseline": {"value": value, "unit": "updates/s", "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": "updates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        }))
        return

    # ---------------------------------------------------------------------------------------------------------
    from heat_b200.model import measure_fp64_peak
    job = Job()
    probe_peak = measure_fp64_peak(job.local_rank)
    main_run = run_workload(job, kind, units, hours, a.steps, a.warmup, sample_clocks=True)

    secondary = []
    if not a.no_secondary and kind == "walls" and not a.units:
        for k2 in ("buildings", "mixed"):
            r = run_workload(job, k2, WORKLOADS[k2]["units"], 240, 2, 3, sample_clocks=False)
            if r is not None:
                secondary.append(r)

    if job.rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except OSError:
            pass
        n_sm = 148
        try:
            import torch
            n_sm = torch.cuda.get_device_properties(job.local_rank).multi_processor_count
        except Exception:
            pass
        clocks = main_run["clocks"] or {}
        f_max = (clocks.get("sm_max_mhz") or peaks.get("sm_max_mhz") or 1965.0) * 1e6
        arch_peak = n_sm * FP64_FMA_PER_CLK_PER_SM * 2 * f_max / 1e12
        fp64_peak = max(probe_peak, arch_peak)
        hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
        hbm_src = "MEASURED_PEAKS.json" if "hbm_gbs" in peaks else "fallback 6.65 TB/s (B200_PROFILING.md)"

        def roofline_of(r, with_traffic):
            achieved = FLOP_PER_UPDATE * r["per_gpu"] / 1e12
            traffic, traffic_src = None, None
            if with_traffic:
                try:   # DRAM bytes per step measured once with ncu for the default workload
                    tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
                    if tj.get("workload") == workload_name(kind, units) and job.world == 1 and tj.get("hours") == hours:
                        traffic = int(tj["dram_bytes_read_per_step"]) + int(tj["dram_bytes_written_per_step"])
                        traffic_src = tj.get("source")
                except (OSError, ValueError, KeyError):
                    pass
            return {"bound": "fp64", "achieved": achieved, "peak": fp64_peak, "unit": "TFLOP/s",
                    "frac": achieved / fp64_peak, "traffic": traffic, "traffic_source": traffic_src,
                    "flop_per_update": FLOP_PER_UPDATE,
                    "peak_source": "max(DFMA probe kernel run in this process, SMs x 128 FLOP/clk x max SM clock)",
                    "peak_dfma_probe": probe_peak, "peak_architectural": arch_peak}

        out = {
            "metric": "wall node-substep updates/sec", "value": main_run["value"], "unit": "updates/s",
            "n_gpus": job.world, "steps": a.steps, "warmup": a.warmup, "ms_per_step": main_run["t_step"] * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": config_of(kind, units, hours, main_run["n_models"]),
            "details": main_run["details"],
            "roofline": roofline_of(main_run, True),
            "roofline_hbm": {"bound": "hbm", "achieved": main_run["alg_bytes"] / main_run["t_step"] / 1e9, "peak": hbm_peak,
                             "unit": "GB/s", "frac": main_run["alg_bytes"] / main_run["t_step"] / 1e9 / hbm_peak,
                             "peak_source": hbm_src,
                             "note": "algorithmic bytes = one pass over the state per series; the launch groups of the "
                                     "shortest chains run as time segments and re-read their state per segment: "
                                     "roofline.traffic is the measured total, still ~1e-4 of the HBM roofline"},
            "e2e": main_run["e2e"], "gpu_launches": main_run["gpu_launches"], "clocks": main_run["clocks"],
        }
        if secondary:
            out["secondary"] = [{
                "config": {"workload": workload_name(r["kind"], WORKLOADS[r["kind"]]["units"]), "hours": 240,
                           "steps_per_hour": 1, "baseline_config": WORKLOADS[r["kind"]]["config"]},
                "value": r["value"], "unit": "updates/s", "ms_per_step": r["t_step"] * 1e3, "steps": 2, "warmup": 3,
                "roofline": {k: v for k, v in roofline_of(r, False).items() if k in ("bound", "achieved", "peak", "unit", "frac")},
                "e2e": r["e2e"], "gpu_launches": r["gpu_launches"], "details": r["details"]} for r in secondary]
        if job.world == 1 and not a.no_secondary and kind == "walls" and not a.units:
            out["drop_in"] = drop_in_rate()
        if job.world == 1 and not a.no_cpu_baseline:
            sdesc, sw, si, h_sample = cpu_sample(main_run["desc"], main_run["weather"], main_run["inputs"], hours)
            g = main_run["gains"]
            sg = None if g is None else g[np.arange(sdesc.n_models) * (main_run["desc"].n_models // sdesc.n_models)]
            rate, upd, dt = cpu_reference_rate(sdesc, sw, si, h_sample, cores, sg)
            out["cpu_baseline"] = {"value": rate, "unit": "updates/s", "cores": cores, "kind": "port",
                                   "sample": f"{sdesc.n_models} models x {h_sample} hourly steps (config.cpu_sample), {dt:.1f} s"}
            out["vs_cpu_baseline"] = out["e2e"]["value"] / rate
        print(json.dumps(out))
    job.close()


if __name__ == "__main__":
    main()

"""Scratch: march every rank's shard of an N-GPU bench job on ONE GPU, one after the other, and report physics failures
(the N-GPU run itself costs N x the box time).   python tools/sweep_shards.py WORLD [walls:8760 buildings:240 mixed:240]"""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT]
import bench
from heat_b200 import cases, _lib, dist
from heat_b200.model import ThermalBatch

world = int(sys.argv[1])
jobs = [a.split(":") for a in (sys.argv[2:] or ["walls:8760", "buildings:240", "mixed:240"])]
for kind, hours in jobs:
    hours = int(hours)
    units = bench.WORKLOADS[kind]["units"]
    weather, inputs = cases.synthetic_weather(hours + 1)
    inputs = np.ascontiguousarray(inputs[:hours])
    t0 = time.time()
    if kind != "mixed":
        d, gains = bench.make_ensemble(kind, units * world, bench.WORKLOADS[kind]["seed"])
        S, N, K = bench.gpu_plan_info(d)
        print(f"{kind}: global ensemble of {d.n_models} in {time.time() - t0:.1f} s", flush=True)
    for rank in (int(r) for r in os.environ.get("SWEEP_RANKS", ",".join(map(str, range(world)))).split(",")):
        if kind == "mixed":
            sd, sg = bench.make_ensemble(kind, units, bench.WORKLOADS[kind]["seed"] + 1000 * rank)
        else:
            idx = dist.deal_models(S, N, rank, world, kind=K)
            sd, sg = d.select_models(idx), (None if gains is None else gains[idx])
        b = ThermalBatch(sd)
        if sg is not None:
            b.set_zone_inputs(q_internal=sg)
        w, i = b.upload(weather), b.upload(inputs)
        try:
            s0 = b.get_state()
            allms = []
            for _ in range(int(os.environ.get("SWEEP_REPS", "1")) - 1):
                allms.append(b.march_series_dev(hours, w, hours + 1, 1, weather_mode=_lib.WEATHER_LINEAR, inputs_dev=i, input_mode=_lib.INPUT_SITE))
                b.set_state(s0)
            if allms:
                print("   earlier reps:", " ".join(f"{x:.0f}" for x in allms), flush=True)
            ms = b.march_series_dev(hours, w, hours + 1, 1, weather_mode=_lib.WEATHER_LINEAR, inputs_dev=i, input_mode=_lib.INPUT_SITE)
            st = b.reduce()
            print(f"{kind} shard {rank}/{world}: ok, {ms:.1f} ms, {b.updates_per_step * hours / ms / 1e6:.1f} G updates/s, "
                  f"flags {st.status_flags}, mean zone {st.sum_zone_t / st.n_zones:.4f}", flush=True)
        except Exception as e:
            print(f"{kind} shard {rank}/{world}: FAILED {e}", flush=True)
        b.close()

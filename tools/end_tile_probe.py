"""Scratch: the S = 118 no-mass-end walls of <= 16 nodes of one shard of the 8-GPU walls job (one G = 2 warp tile, the
slowest launch group of the shard), marched alone: all together, by kind of end, and one by one."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT]
import bench
from heat_b200 import cases, _lib, dist
from heat_b200.model import ThermalBatch

rank, hours = int(sys.argv[1]), int(sys.argv[2])
d, _ = bench.make_ensemble("walls", 800000, 42)
S, N, K = bench.gpu_plan_info(d)
sd = d.select_models(dist.deal_models(S, N, rank, 8, kind=K))
p = ThermalBatch(sd, device=-1)
pick = []
for m in range(sd.n_models):
    if p.model_info(m).dt_subdivisions >= 96:
        si = p.surface_info(m)
        if si.n_nomass_chunks > 0 and si.n_nodes <= 16:
            a = p.surface_arrays(m)
            pick.append((m, si.n_nodes, int(a["nomass_chunks"][0][0]) == 0))
p.close()
weather, inputs = cases.synthetic_weather(hours + 1)
inputs = np.ascontiguousarray(inputs[:hours])


def run(models, label):
    b = ThermalBatch(sd.select_models(np.array(models)))
    w, i = b.upload(weather), b.upload(inputs)
    for _ in range(2):
        ms = b.march_series_dev(hours, w, hours + 1, 1, weather_mode=_lib.WEATHER_LINEAR, inputs_dev=i, input_mode=_lib.INPUT_SITE)
    print(f"{label}: {ms:.2f} ms -> {ms * 1e-3 * 1.965e9 / (hours * 118):.0f} cycles per sub-step, iterations {b.reduce().nomass_iterations / (2 * hours * 118 * len(models)):.2f} per solve", flush=True)
    b.close()


run([m for m, _, _ in pick], f"all {len(pick)} walls")
run([m for m, _, f in pick if f], "front no-mass end")
run([m for m, _, f in pick if not f], "back no-mass end")
run([m for m, n, _ in pick if n <= 8], "n <= 8")
run([m for m, n, _ in pick if n > 8], "n > 8")
for m, n, f in pick:
    run([m], f"wall {m} n={n} {'front' if f else 'back'}")

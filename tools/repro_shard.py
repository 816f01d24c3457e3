"""Scratch: march ONE rank's shard of the N-GPU configs[2] job on one GPU in chunks of hours and report the first
physics failure (model, wall data), then march the failing model alone."""
import os, re, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT]
import bench
from heat_b200 import cases, _lib, dist
from heat_b200.model import ThermalBatch

rank, world, hours, chunk = (int(a) for a in sys.argv[1:5])
d, _ = bench.make_ensemble("walls", 100_000 * world, 42)
S, N, K = bench.gpu_plan_info(d)
idx = dist.deal_models(S, N, rank, world, kind=K)
sd = d.select_models(idx)
weather, inputs = cases.synthetic_weather(hours + 1)
inputs = np.ascontiguousarray(inputs[:hours])


def march(desc, label, whole=False):
    b = ThermalBatch(desc)
    done = 0
    try:
        while done < hours:
            n = hours if whole else min(chunk, hours - done)
            b.march_series(n, weather[done:done + n + 1], weather_mode=_lib.WEATHER_LINEAR, inputs=inputs[done:done + n],
                           input_mode=_lib.INPUT_SITE)
            done += n
        st = b.get_state()
        print(label, "ok", hours, "h; state finite:", bool(np.isfinite(st).all()), "zone mean", b.reduce().mean_zone_t, flush=True)
        return None
    except Exception as e:
        print(label, "FAILED in hours", done, "…", done + n, ":", e, flush=True)
        m = re.search(r"model (\d+)", str(e))
        return int(m.group(1)) if m else -1
    finally:
        b.close()


bad = march(sd, f"shard {rank}/{world} whole series", whole=True)
bad2 = march(sd, f"shard {rank}/{world} chunks of {chunk}")
for m in {bad, bad2} - {None, -1}:
    g = idx[m]
    print("model", m, "global", g, "S", S[g], "N", N[g], "surface", sd.surfaces[m], flush=True)
    c0, c1 = d.construction_offset[d.surfaces[g]["construction"]], d.construction_offset[d.surfaces[g]["construction"] + 1]
    print("layers", d.layers[c0:c1], flush=True)
    march(sd.select_models(np.array([m])), "failing model alone")
    lo = max(0, m - 64)
    march(sd.select_models(np.arange(lo, min(sd.n_models, m + 64))), "with shard neighbours")
os.environ["HB_NO_WTILE"] = "1"
march(sd, "block-tile kernels (HB_NO_WTILE=1) whole series", whole=True)

"""Scratch diagnostic: where does the GPU leave the oracle on epw/massive_full?"""
import sys, os
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle")]
import parity_cases as pc

name = sys.argv[1] if len(sys.argv) > 1 else "epw/massive_full"
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 8760
zo, fo, so, m = pc.run_oracle(name, steps, record_state=True)
zg, fg, sg, b = pc.run_gpu(name, steps, record_state=True)
err = np.abs(sg - so)
err[:, m.state_offset(0) + 4: m.state_offset(0) + 8] = 0
per_step = err.max(axis=1)
bad = np.nonzero(per_step > 1e-9)[0]
print("steps", steps, "max err", per_step.max(), "first step >1e-9:", bad[:5])
c = pc.make_case(name, steps)
if bad.size:
    k = bad[0]
    print("weather rows around", k, "\n", c.weather[max(0, k - 1):k + 3])
    print("inputs", c.inputs[max(0, k - 1):k + 2])
    print("state gpu", sg[k][:12], "\nstate ora", so[k][:12])
    print("err by slot", err[k][:16])
import torch
S = m.dt_subdivisions
w = c.weather
f = (np.arange(S) + 1.0) / S
wd = w[:-1, 1:2] + (w[1:, 1:2] - w[:-1, 1:2]) * f[None, :]
ang = wd * (np.pi / 180.0)
cg = torch.cos(torch.tensor(ang, device="cuda", dtype=torch.float64)).cpu().numpy()
cn = np.cos(ang)
flip = (-(cg) > 0) != (-(cn) > 0)
print("windward sign flips torch-cuda vs numpy:", int(flip.sum()), "at", np.argwhere(flip)[:5])
print("min |cos|", np.abs(cn).min())

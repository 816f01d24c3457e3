"""Scratch: per-(G) tile statistics of the walls workload (mirrors hb_host.cc packing)."""
import sys, os, ctypes as C, numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path[:0] = [ROOT]
from heat_b200 import cases, _lib
from heat_b200.model import ThermalBatch
n = int(sys.argv[1]) if len(sys.argv) > 1 else 100000
d = cases.wall_ensemble(n, seed=42)
p = ThermalBatch(d, device=-1)
info = _lib.SurfaceInfoC(); mi = _lib.ModelInfoC()
N = np.zeros(n, int); S = np.zeros(n, int); NM = np.zeros(n, int)
for s in range(n):
    p._lib.hb_get_surface_info(p._h, s, C.byref(info)); N[s] = info.n_nodes; NM[s] = info.n_nomass_chunks
    p._lib.hb_get_model_info(p._h, s, C.byref(mi)); S[s] = mi.dt_subdivisions
G = np.ones(n, int)
for k in range(n):
    g = 1
    while g * 8 < N[k] + 2: g *= 2
    G[k] = g
for g in (1, 2, 4, 8, 16, 32):
    m = (G == g) & (NM == 0)
    W = 32 if g <= 8 else 16 if g == 16 else 8
    # tiles: sorted by S, closed on S change
    tiles = 0; bs = 0
    for sv in np.unique(S[m]):
        c = int((m & (S == sv)).sum()); t = -(-c // W); tiles += t; bs += t * sv
    print(f"G={g:2d} walls {m.sum():6d} tiles {tiles:5d} block-substeps/step {bs:7d} updates/step {int((N[m]*S[m]).sum()):9d} fill {m.sum()/max(tiles*W,1):.2f}")

#!/usr/bin/env python
"""Developer timing of the band fill alone (Mercator band of -r R): python tools/bandbench.py [r] [reps]"""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
from ocean_model_grid_generator_b200 import _cabi, device as D
r = float(sys.argv[1]) if len(sys.argv) > 1 else 4
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 20
D.ready()
lib = _cabi.load()
Ni, nyp = int(720 * r), int(350 * r) + 1
lam = D.to_dev(-300.0 + np.arange(Ni + 1) * 360.0 / Ni)
phi = D.to_dev(np.linspace(-66.0, 64.0, nyp))
nxp = Ni + 1
x, y, a = D.empty(nyp, nxp), D.empty(nyp, nxp), D.empty(nyp, nxp)
dx, dy, ar = D.empty(nyp, nxp - 1), D.empty(nyp - 1, nxp), D.empty(nyp - 1, nxp - 1)
st = D.stream()
def run():
    _cabi.call("ogg_band_fill_metrics", D.ptr(lam), D.ptr(phi), nyp, nxp, 6371.0e3, 0, D.ptr(x), D.ptr(y), D.ptr(dx), D.ptr(dy), D.ptr(ar), D.ptr(a), st)
for _ in range(3):
    run()
torch.cuda.synchronize()
lib.ogg_timing_reset(); lib.ogg_timing_enable(1)
for _ in range(reps):
    run()
torch.cuda.synchronize()
lib.ogg_timing_enable(0)
for k in range(lib.ogg_timing_kernels()):
    tot, n = ctypes.c_double(), ctypes.c_longlong()
    lib.ogg_timing_get(k, ctypes.byref(tot), ctypes.byref(n))
    if n.value and lib.ogg_timing_name(k).decode() == "band_fill":
        ms = tot.value / n.value
        print(os.environ.get("OGG_B200_LIB", "default"), "r=%g %dx%d band_fill %.4f ms  %.0f GB/s" % (r, nyp, nxp, ms, 48.0 * nyp * nxp / ms / 1e6))

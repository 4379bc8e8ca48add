#!/usr/bin/env python
"""Developer A/B timing of the two cap quadrature entry points at the 1/4-degree bench sizes.

    OGG_B200_LIB=/path/to/variant.so python tools/kbench.py [r] [reps]

Prints per-kernel mean ms (CUDA events around every launch, ogg_timing_*) and a sha256 of the outputs so that
variants can be checked bit-identical to each other."""
import ctypes
import hashlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from ocean_model_grid_generator_b200 import _cabi, device as D  # noqa: E402
from ocean_model_grid_generator_b200 import ocean_grid_generator as G  # noqa: E402

r = float(sys.argv[1]) if len(sys.argv) > 1 else 4
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
world = int(sys.argv[3]) if len(sys.argv) > 3 else 1   # time one rank's row-slab of a `world`-GPU run
D.ready()
lib = _cabi.load()
Ni = int(r * 720)
Nj_bp, Nj_sc = int(120 * r), int(r * 40) * 7 // 4
lat0_bp, lon0 = 64.05895973, -300.0
rp = G.bipolar_rp(lat0_bp)
jmin = int(-(-0.49 * Nj_sc // 1))
jmin += jmin % 2
st = D.stream()
out = {}


def run():
    dx, dy, da = D.empty(Nj_bp + 1, Ni), D.empty(Nj_bp, Ni + 1), D.empty(Nj_bp, Ni)
    _cabi.call("ogg_bipolar_cap_metrics_quad", 5, Ni, Nj_bp, Nj_bp - Nj_bp // world, Nj_bp, lat0_bp, lon0, rp, 6371.0e3, 0, D.ptr(dx), D.ptr(dy), D.ptr(da), st)
    n = Nj_sc - jmin
    sx, sy, sa = D.empty(n + 1, Ni), D.empty(n, Ni + 1), D.empty(n, Ni)
    _cabi.call("ogg_dp_metrics_quad", 4, Ni, Nj_sc, Nj_sc - n // world, Nj_sc, lon0, -78.0, 80.0, 0.2, 6371.0e3, 0, D.ptr(sx), D.ptr(sy), D.ptr(sa), st)
    out.update(bp=(dx[:-1], dy, da), dp=(sx[:-1], sy, sa))


for _ in range(3):
    run()
torch.cuda.synchronize()
lib.ogg_timing_reset(); lib.ogg_timing_enable(1)
for _ in range(reps):
    run()
torch.cuda.synchronize()
lib.ogg_timing_enable(0)
res = {}
for k in range(lib.ogg_timing_kernels()):
    tot, n = ctypes.c_double(), ctypes.c_longlong()
    lib.ogg_timing_get(k, ctypes.byref(tot), ctypes.byref(n))
    if n.value:
        res[lib.ogg_timing_name(k).decode()] = tot.value / reps
h = {k: hashlib.sha256(b"".join(t.cpu().numpy().tobytes() for t in v)).hexdigest()[:16] for k, v in out.items()}
print(os.environ.get("OGG_B200_LIB", "default"), "r=%g world=%d" % (r, world), " ".join("%s=%.4f" % kv for kv in sorted(res.items(), key=lambda kv: -kv[1])[:4]), h)

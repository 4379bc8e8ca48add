"""O1 oracle: import the UNMODIFIED reference script with the repo's portable-trig twin.

TEST INFRASTRUCTURE; works only where /root/reference exists (the build container).
`load_reference()` returns the module object of /root/reference/ocean_grid_generator.py imported
with  oracle/numpypi  as `numpypi` and oracle/shim/netCDF4.py as `netCDF4`.

Canonical semantics need numpy's strict (non-FMA, non-SVML) loops, which numpy only selects at
import time: run under  NPY_DISABLE_CPU_FEATURES=<STRICT_FEATURES>  (make_golden.py re-execs
itself that way).  `strict_numpy()` tells whether the current process satisfies that.
"""
import importlib.util
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REFERENCE = os.environ.get("OGG_REFERENCE", "/root/reference/ocean_grid_generator.py")
STRICT_FEATURES = ("SSSE3 SSE41 POPCNT SSE42 AVX F16C FMA3 AVX2 AVX512F AVX512CD AVX512_SKX "
                   "AVX512_CLX AVX512_CNL AVX512_ICL AVX512_SPR")


def strict_numpy():
    import numpy as np
    a = np.array([0.1 + 0.7j, 1e-3 - 3.3j, 0.3 + 0.9j])
    b = np.array([0.9 - 0.3j, 7.7 + 0.13j, 1.1 - 1e-3j])
    p = a * b
    re = a.real * b.real - a.imag * b.imag
    im = a.real * b.imag + a.imag * b.real
    return bool((p.real == re).all() and (p.imag == im).all())


def load_reference():
    if not os.path.exists(REFERENCE):
        raise FileNotFoundError(REFERENCE)
    for p in (os.path.join(HERE, "shim"), HERE):
        if p not in sys.path:
            sys.path.insert(0, p)
    spec = importlib.util.spec_from_file_location("ogg_reference", REFERENCE)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod

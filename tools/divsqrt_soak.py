#!/usr/bin/env python
"""Longer run of ogg_selftest_divsqrt than the test suite's: python tools/divsqrt_soak.py [seeds] [log2 n]"""
import ctypes, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from ocean_model_grid_generator_b200 import _cabi, device as D
seeds = int(sys.argv[1]) if len(sys.argv) > 1 else 8
n = 1 << (int(sys.argv[2]) if len(sys.argv) > 2 else 28)
D.ready()
lib = _cabi.load()
tot = {m: [0, 0, 0] for m in range(5)}
t0 = time.time()
for s in range(seeds):
    for m in range(5):
        c = (ctypes.c_uint64 * 3)()
        assert lib.ogg_selftest_divsqrt(n if m != 4 else n // 4, 0x1234567 + 7919 * s + 104729 * m, m, c) == 0
        for k in range(3):
            tot[m][k] += c[k]
for m in range(5):
    print("mode %d: accepted %d  differing %d  in-range rejected %d" % (m, tot[m][0], tot[m][1], tot[m][2]))
print("total accepted %.3e, differing %d, %.1f s" % (sum(v[0] for v in tot.values()), sum(v[1] for v in tot.values()), time.time() - t0))

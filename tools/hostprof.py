#!/usr/bin/env python
"""Developer profile of the host-side cost of Supergrid.build() (Python + C-ABI enqueue time).  python tools/hostprof.py [r]"""
import contextlib, cProfile, io, os, pstats, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from ocean_model_grid_generator_b200 import supergrid as SG, ocean_grid_generator as G, device as D
r = sys.argv[1] if len(sys.argv) > 1 else "4"
D.ready()
args = vars(G._build_parser().parse_args(("-r %s --r_dp 0.2" % r).split()))
pa = {k: v for k, v in args.items() if k not in ("gridfilename", "debug", "plotem", "no_changing_meta", "write_subgrid_files")}
with contextlib.redirect_stdout(io.StringIO()):
    plan = SG.plan_supergrid(**pa)
sg = SG.Supergrid(plan)
pinned = sg.pinned_like()
for _ in range(5):
    sg.build()
torch.cuda.synchronize()
for host in (None, pinned):
    for conc in (True, False):
        ts = []
        for _ in range(20):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            sg.build(concurrent=conc, host=host)
            ts.append((time.perf_counter() - t0) * 1e3)
        torch.cuda.synchronize()
        print("host=%s concurrent=%s: enqueue time per build: min %.3f median %.3f ms" % (host is not None, conc, min(ts), sorted(ts)[10]))
pr = cProfile.Profile()
pr.enable()
for _ in range(50):
    sg.build()
pr.disable()
torch.cuda.synchronize()
s = io.StringIO()
pstats.Stats(pr, stream=s).sort_stats("cumulative").print_stats(22)
print(s.getvalue()[:5000])

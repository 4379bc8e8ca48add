#!/usr/bin/env python
"""Developer timing of one whole-supergrid build: concurrent piece streams vs one stream.  python tools/stepbench.py [r] [steps]"""
import contextlib, io, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from ocean_model_grid_generator_b200 import supergrid as SG, ocean_grid_generator as G, device as D
r = sys.argv[1] if len(sys.argv) > 1 else "4"
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 50
D.ready()
args = vars(G._build_parser().parse_args(("-r %s --r_dp 0.2" % r).split()))
pa = {k: v for k, v in args.items() if k not in ("gridfilename", "debug", "plotem", "no_changing_meta", "write_subgrid_files")}
with contextlib.redirect_stdout(io.StringIO()):
    plan = SG.plan_supergrid(**pa)
sg = SG.Supergrid(plan)
for conc in (False, True, False, True):
    for _ in range(5):
        sg.build(concurrent=conc)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        sg.build(concurrent=conc)
    e1.record()
    torch.cuda.synchronize()
    print("r=%s concurrent=%s  %.4f ms/step  area=%.17g" % (r, conc, e0.elapsed_time(e1) / steps, sg.area_sum()))

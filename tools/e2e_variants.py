#!/usr/bin/env python
"""Developer A/B of device->host strategies for the end-to-end call.  python tools/e2e_variants.py [r]"""
import contextlib, io, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from ocean_model_grid_generator_b200 import supergrid as SG, ocean_grid_generator as G, device as D, _cabi
r = sys.argv[1] if len(sys.argv) > 1 else "4"
D.ready()
args = vars(G._build_parser().parse_args(("-r %s --r_dp 0.2" % r).split()))
pa = {k: v for k, v in args.items() if k not in ("gridfilename", "debug", "plotem", "no_changing_meta", "write_subgrid_files")}
with contextlib.redirect_stdout(io.StringIO()):
    plan0 = SG.plan_supergrid(**pa)
sg0 = SG.Supergrid(plan0)
pinned = sg0.pinned_like()
area = D.empty(1)
copy_stream = torch.cuda.Stream()

def run(mode, replan=True):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    if replan:
        with contextlib.redirect_stdout(io.StringIO()):
            plan = SG.plan_supergrid(**pa)
        sg = SG.Supergrid(plan)
    else:
        sg = sg0
    t1 = time.perf_counter()
    if mode == "piece_streams":
        sg.build(host=pinned)
    elif mode == "after":
        sg.build()
        for k, v in sg.arrays().items():
            pinned[k].copy_(v, non_blocking=True)
    elif mode == "one_stream":
        sg.build(host=pinned, copy_stream=copy_stream)
    _cabi.call("ogg_sum_strided", D.ptr(sg.area), sg.area.numel(), 1, D.ptr(area), D.stream())
    t2 = time.perf_counter()
    s = float(area[0])
    torch.cuda.synchronize()
    t3 = time.perf_counter()
    return (t1 - t0) * 1e3, (t2 - t1) * 1e3, (t3 - t2) * 1e3, (t3 - t0) * 1e3

for mode in ("piece_streams", "after") * 2:
    try:
        res = [run(mode) for _ in range(6)]
    except TypeError as e:
        print(mode, "unsupported", e); continue
    best = min(res[1:], key=lambda x: x[3])
    print("%-14s plan+alloc %.2f  queue %.2f  wait %.2f  total %.2f ms   (all totals: %s)" % ((mode,) + best + (" ".join("%.2f" % x[3] for x in res),)))

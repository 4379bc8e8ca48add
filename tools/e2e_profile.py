#!/usr/bin/env python
"""Developer breakdown of the end-to-end call (plan + build + D2H).  python tools/e2e_profile.py [r]"""
import contextlib, io, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from ocean_model_grid_generator_b200 import supergrid as SG, ocean_grid_generator as G, device as D
r = sys.argv[1] if len(sys.argv) > 1 else "4"
D.ready()
args = vars(G._build_parser().parse_args(("-r %s --r_dp 0.2" % r).split()))
pa = {k: v for k, v in args.items() if k not in ("gridfilename", "debug", "plotem", "no_changing_meta", "write_subgrid_files")}
pinned = None
for it in range(4):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    with contextlib.redirect_stdout(io.StringIO()):
        plan = SG.plan_supergrid(**pa)
    torch.cuda.synchronize(); t1 = time.perf_counter()
    sg = SG.Supergrid(plan)
    torch.cuda.synchronize(); t2 = time.perf_counter()
    sg.build()
    torch.cuda.synchronize(); t3 = time.perf_counter()
    arrs = sg.arrays()
    if pinned is None:
        pinned = {k: torch.empty(v.shape, dtype=torch.float64, pin_memory=True) for k, v in arrs.items()}
    t3 = time.perf_counter()
    for k, v in arrs.items():
        pinned[k].copy_(v, non_blocking=True)
    torch.cuda.synchronize(); t4 = time.perf_counter()
    s = float(pinned["area"].sum()); t5 = time.perf_counter()
    print("iter %d: plan %.2f ms, alloc %.2f, build %.2f, d2h %.2f (%.1f GB/s), host sum %.2f, total %.2f"
          % (it, (t1 - t0) * 1e3, (t2 - t1) * 1e3, (t3 - t2) * 1e3, (t4 - t3) * 1e3,
             sum(v.numel() * 8 for v in pinned.values()) / (t4 - t3) / 1e9, (t5 - t4) * 1e3, (t5 - t0) * 1e3))

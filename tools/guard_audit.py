#!/usr/bin/env python
"""Audit of the operand-range guards the cap kernels leave out as proven (csrc/ogg_trig.cuh "ELIDE", dp_vertex_fast).

    tools/build_variant.sh audit -DOGG_GUARD_AUDIT
    OGG_B200_LIB=$PWD/build/libogg_audit.so python tools/guard_audit.py

Builds a set of supergrids (the reference's recipes incl. pole rows, odd resolutions, displaced poles close to the cap
edge) with a library that evaluates every elided claim, and prints how many units passed their guards and for how many
of those a claim was false (must be 0)."""
import contextlib, ctypes, io, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from ocean_model_grid_generator_b200 import _cabi, device as D, supergrid as SG, ocean_grid_generator as G  # noqa: E402

D.ready()
lib = _cabi.load()
counts = (ctypes.c_uint64 * 2)()
assert lib.ogg_guard_audit(counts, 1) == 0
RECIPES = ["-r 4 --r_dp 0.2", "-r 1", "-r 2 --r_dp 0.2 --south_cutoff_row 2", "-r 1.5 --r_dp 0.05 --lon_dp 123.0",
           "-r 3 --r_dp 0.6 --lon_dp 10.0", "-r 2 --lat_dp -85.85", "-r 0.5 --r_dp 0.9",
           "-r 8 --r_dp 0.2 --south_cutoff_row 5 --match_dy bp so p125sc --ensure_nj_even"]
if "--big" in sys.argv:   # soak: the 1/12- and 1/30-degree grids (1.1e9 and 6.7e9 samples)
    RECIPES = ["-r 12 --r_dp 0.2", "-r 30 --r_dp 0.2 --ensure_nj_even"]
out = []
for rec in RECIPES:
    try:
        args = vars(G._build_parser().parse_args(rec.split()))
    except SystemExit:
        print("skipped (flags not accepted):", rec); continue
    pa = {k: v for k, v in args.items() if k not in ("gridfilename", "debug", "plotem", "no_changing_meta", "write_subgrid_files")}
    with contextlib.redirect_stdout(io.StringIO()):
        plan = SG.plan_supergrid(**pa)
        sg = SG.Supergrid(plan)
        sg.build()
    torch.cuda.synchronize()
    assert lib.ogg_guard_audit(counts, 1) == 0
    out.append({"recipe": rec, "units_passed": int(counts[0]), "claims_false": int(counts[1])})
    print(json.dumps(out[-1]), flush=True)
    del sg
    torch.cuda.empty_cache()
# whole caps incl. the pole rows (the supergrid recipes above cut the displaced-pole cap's inner rows away)
for order, Ni, Nj in ((4, 1440, 140), (2, 360, 35)):
    sx, sy, sa = D.empty(Nj + 1, Ni), D.empty(Nj, Ni + 1), D.empty(Nj, Ni)
    _cabi.call("ogg_dp_metrics_quad", order, Ni, Nj, 0, Nj, -300.0, -78.0, 80.0, 0.2, 6371.0e3, 0, D.ptr(sx), D.ptr(sy), D.ptr(sa), D.stream())
    torch.cuda.synchronize()
    assert lib.ogg_guard_audit(counts, 1) == 0
    out.append({"recipe": "whole displaced-pole cap order %d %dx%d" % (order, Ni, Nj), "units_passed": int(counts[0]), "claims_false": int(counts[1])})
    print(json.dumps(out[-1]))
tot = sum(o["units_passed"] for o in out)
bad = sum(o["claims_false"] for o in out)
print(json.dumps({"audit_build": tot > 0, "units_passed": tot, "claims_false": bad}))
sys.exit(0 if (tot > 0 and bad == 0) else 1)

#!/usr/bin/env python
"""Derives the per-unit FP64 operation counts behind bench.py's roofline from the CPU oracle (oracle/opcount.cpp compiles
oracle/ogg_oracle.c with an operator-counting scalar in place of double) and writes profiles/opcount.json.

    python tools/opcount.py [--resolution R] [--out profiles/opcount.json] [--print]

Three weightings of the same tallies are added under "totals" (what bench.py reports side by side):
  ops_8d     SURVEY 8(d): one op per add/sub/mul/div/sqrt/fmod/compare-select (rint counted too), on the "minimal" tally
  ops_issue  FP64-pipe instructions of the IEEE-correct expansions on sm_100a: add/sub/mul/DSETP 1, division 8 (5 DFMA to
             refine the MUFU seed, DMUL, 2 DFMA), division by a cached divisor 3, square root 8; rint/selects/sign ops 0
  ops_reference  SURVEY 8(d) weights on the un-hoisted, both-polynomials restatement (what the reference itself evaluates
             per unique sample) -- reported for orientation only
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
W_8D = dict(add_sub=1, mul=1, div=1, div_by_constant=1, div_by_one=1, sqrt=1, fmod=1, compare=1, round=1)
W_ISSUE = dict(add_sub=1, mul=1, div=8, div_by_constant=3, div_by_one=0, sqrt=8, fmod=0, compare=1, round=0)


def weighted(tally, w):
    return sum(tally[k] * w.get(k, 0) for k in tally)


def derive(resolution=1.0):
    exe = os.path.join(tempfile.mkdtemp(prefix="opcount"), "opcount")
    subprocess.check_call(["g++", "-O1", "-std=c++17", "-ffp-contract=off", "-fno-fast-math", "-fpermissive", "-w", "-o", exe,
                           os.path.join(ROOT, "oracle", "opcount.cpp")])
    d = json.loads(subprocess.check_output([exe, str(resolution)]).decode())
    d["weights"] = {"ops_8d": W_8D, "ops_issue": W_ISSUE}
    d["totals"] = {}
    for unit in ("bipolar_sample", "dp_vertex", "haversine", "fd_combine_x2", "dp_hsample"):
        d["totals"][unit] = {"ops_8d": weighted(d[unit]["minimal"], W_8D), "ops_issue": weighted(d[unit]["minimal"], W_ISSUE),
                             "ops_8d_hoisted_both_polynomials": weighted(d[unit]["hoisted"], W_8D),
                             "ops_reference": weighted(d[unit]["reference"], W_8D)}
    d["C_f_totals"] = {f: {"reference": weighted(v["reference"], W_8D), "minimal": weighted(v["minimal"], W_8D)}
                       for f, v in d["C_f"].items()}
    d["derived_by"] = "tools/opcount.py (oracle/opcount.cpp: operator-counting scalar through oracle/ogg_oracle.c)"
    return d


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--resolution", type=float, default=4.0)
    ap.add_argument("--out", default=os.path.join(ROOT, "profiles", "opcount.json"))
    ap.add_argument("--print", action="store_true")
    a = ap.parse_args()
    d = derive(a.resolution)
    if a.print:
        json.dump(d["totals"], sys.stdout, indent=1)
        print()
    else:
        with open(a.out, "w") as f:
            json.dump(d, f, indent=1, sort_keys=True)
        print("wrote", a.out, {k: round(v["ops_8d"], 2) for k, v in d["totals"].items()})

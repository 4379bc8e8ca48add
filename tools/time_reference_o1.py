#!/usr/bin/env python
"""Wall time of the UNMODIFIED reference (O1: /root/reference/ocean_grid_generator.py run as __main__, with
oracle/numpypi standing in for the un-vendored numpypi and oracle/shim/netCDF4.py for the absent netCDF4) in BOTH
numpy dispatch modes:
  strict   NPY_DISABLE_CPU_FEATURES=<oracle/ref_harness.STRICT_FEATURES> -- the mode the goldens are generated in
  default  numpy's own runtime dispatch (AVX2/AVX-512 loops): what a user of the reference gets
Works only where /root/reference exists (this container); the result is committed as profiles/cpu_o1_<host>.json
and quoted in BASELINE.md.  Usage: python tools/time_reference_o1.py [--recipes "-r 2" "-r 4 --r_dp 0.2"]
"""
import argparse
import json
import os
import platform
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import ref_harness as H  # noqa: E402

CHILD = r"""
import contextlib, io, sys, time
sys.path.insert(0, %(oracle)r)
import ref_harness as H
H.load_reference.__globals__  # touch
for p in (%(shim)r, %(oracle)r):
    if p not in sys.path: sys.path.insert(0, p)
src = open(H.REFERENCE).read()
sys.argv = ["ocean_grid_generator.py", "-f", "/tmp/o1_timing.nc"] + %(flags)r.split()
buf = io.StringIO()
t0 = time.time()
with contextlib.redirect_stdout(buf):
    exec(compile(src, H.REFERENCE, "exec"), {"__name__": "__main__", "__file__": H.REFERENCE})
wall = time.time() - t0
rt = [float(l.split()[-1]) for l in buf.getvalue().splitlines() if l.startswith("runtime(secs)")]
shape = [l for l in buf.getvalue().splitlines() if l.startswith("shapes:")]
print("O1RESULT", wall, rt[0] if rt else -1.0, H.strict_numpy(), shape[-1] if shape else "")
"""


def run(flags, strict):
    env = dict(os.environ)
    env.pop("NPY_DISABLE_CPU_FEATURES", None)
    if strict:
        env["NPY_DISABLE_CPU_FEATURES"] = H.STRICT_FEATURES
    code = CHILD % {"oracle": os.path.join(ROOT, "oracle"), "shim": os.path.join(ROOT, "oracle", "shim"), "flags": flags}
    t0 = time.time()
    out = subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True)
    line = [l for l in out.stdout.splitlines() if l.startswith("O1RESULT")]
    if not line:
        raise RuntimeError(out.stderr[-2000:])
    f = line[0].split(None, 4)
    return {"flags": flags, "mode": "strict" if strict else "default", "wall_secs": float(f[1]),
            "reference_runtime_secs": float(f[2]), "strict_loops_active": f[3] == "True",
            "shapes_line": f[4] if len(f) > 4 else "", "process_secs": time.time() - t0}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--recipes", nargs="*", default=["-r 2", "-r 4 --r_dp 0.2"])
    ap.add_argument("--out", default=os.path.join(ROOT, "profiles", "cpu_o1_container.json"))
    a = ap.parse_args()
    import numpy as np
    res = {"host": platform.node(), "cpu": platform.processor() or open("/proc/cpuinfo").read().split("model name")[1].split("\n")[0].strip(": \t"),
           "cores_visible": os.cpu_count(), "numpy": np.__version__, "python": platform.python_version(),
           "what": "unmodified reference CLI (O1), one process, numpy's own threading (none for these loops)", "runs": []}
    for flags in a.recipes:
        for strict in (True, False):
            r = run(flags, strict)
            print(r, flush=True)
            res["runs"].append(r)
            json.dump(res, open(a.out, "w"), indent=1)


if __name__ == "__main__":
    main()

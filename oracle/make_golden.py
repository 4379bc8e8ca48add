"""Generates tests/golden/* by running the UNMODIFIED reference (O1) -- TEST INFRASTRUCTURE.

Runs only in the build container (needs /root/reference).  The reference is imported with the portable-trig
twin (oracle/numpypi) standing in for the un-vendored numpypi and oracle/shim/netCDF4.py for the absent
netCDF4, under numpy's strict loops (NPY_DISABLE_CPU_FEATURES, see ref_harness.py) so that complex multiply
and every non-overridden ufunc are plain IEEE / glibc.  Outputs:
  tests/golden/functions.npz   small per-function input/output arrays (bit patterns)
  tests/golden/configs.json    per-array sha256 + shape + sums of the six supergrid arrays for the
                               reference's own CLI recipes (extras/Makefile:28-39)
Usage: python oracle/make_golden.py [--configs quick|all] [--functions]
"""
import contextlib
import hashlib
import io
import json
import os
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, HERE)
import ref_harness as H  # noqa: E402

if os.environ.get("NPY_DISABLE_CPU_FEATURES") != H.STRICT_FEATURES:
    env = dict(os.environ, NPY_DISABLE_CPU_FEATURES=H.STRICT_FEATURES)
    os.execve(sys.executable, [sys.executable] + sys.argv, env)

import numpy as np  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")

CONFIGS = {  # extras/Makefile:28-39 (file name / subgrid flags dropped; they do not change the arrays)
    "res4.0": "-r 0.25 --ensure_nj_even --no_changing_meta",
    "res1.0": "-r 1.0 --south_cutoff_row 2 --no_changing_meta",
    "res0.5": "-r 2 --no_changing_meta",
    "res0.5_equenh": "-r 2 --enhanced_equatorial=4 --no_changing_meta",
    "res0.25": "-r 4 --r_dp 0.2 --south_cutoff_row 83 --no_changing_meta",
    "res0.25_om5proto": "-r 4 --south_ocean_lower_lat -88.57 --match_dy so --no_south_cap --no_changing_meta",
    # BASELINE.json configs[0] geometry through the CLI that exists in this snapshot
    "readme_dp": "-r 2 --lon_dp 60.0 --lat_dp -85.85 --no_changing_meta",
    "res1.0_dp": "-r 1.0 --r_dp 0.2 --south_cutoff_row 2 --no_changing_meta",
    # BASELINE.json configs[1]: the benchmark's headline workload, no south cut (all 143 kept displaced-pole rows compared)
    "res0.25_full": "-r 4 --r_dp 0.2 --no_changing_meta",
    # the reference's own 1/8-degree recipe (extras/Makefile:40-41): match_dy bp + so + p125sc (OGG:1050-1057, 1094-1097, 1123-1125)
    "res0.125": "-r 8 --r_dp 0.2 --south_cutoff_row 5 --match_dy bp so p125sc --ensure_nj_even --no_changing_meta",
    # plan branches no extras/Makefile recipe reaches: match_dy bp alone, p125sc alone, the angle cut (OGG:1276-1279),
    # a row cut that removes the whole southern cap and part of the Southern Ocean band (OGG:1296-1313)
    "res1.0_bp": "-r 1.0 --r_dp 0.2 --match_dy bp --south_cutoff_row 2 --no_changing_meta",
    "res1.0_p125sc": "-r 1.0 --r_dp 0.2 --match_dy p125sc --ensure_nj_even --south_cutoff_row 2 --no_changing_meta",
    "res1.0_cutang": "-r 1.0 --r_dp 0.2 --south_cutoff_ang -80.4 --no_changing_meta",
    "res1.0_cutso": "-r 1.0 --r_dp 0.2 --south_cutoff_row 45 --no_changing_meta",
    "res0.5_cutang_regular": "-r 2 --south_cutoff_ang -81.1 --no_changing_meta",
    # --ensure_nj_even on a default-flag grid, where it shortens the Mercator band (OGG:329-333) AND cuts the southern row
    # of the doughnut-cut displaced-pole cap (OGG:1188-1194): pins oracle/supergrid_cpu.build_cpu(ensure_nj_even=True),
    # the checker of the 1/30-degree slabs (BASELINE configs[3] is `-r 30 --r_dp 0.2 --ensure_nj_even`)
    "res0.667_dp_even": "-r 1.5 --r_dp 0.2 --ensure_nj_even --no_changing_meta",
}
QUICK = ["res4.0", "res1.0", "res0.5", "res0.5_equenh", "readme_dp", "res1.0_dp"]


def quiet(f, *a, **k):
    with contextlib.redirect_stdout(io.StringIO()):
        return f(*a, **k)


LAST_STDOUT = [""]


def run_config(R, flags):
    import argparse  # noqa: F401
    import netCDF4
    src = open(H.REFERENCE).read()
    # reuse the reference's own argparse block: exec the module text as __main__ with argv set
    argv = ["ocean_grid_generator.py", "-f", "/tmp/golden_tmp.nc"] + flags.split()
    old = sys.argv
    sys.argv = argv
    try:
        g = {"__name__": "__main__", "__file__": H.REFERENCE}
        buf = io.StringIO()
        with contextlib.redirect_stdout(buf):
            exec(compile(src, H.REFERENCE, "exec"), g)
        LAST_STDOUT[0] = buf.getvalue()
    finally:
        sys.argv = old
    return netCDF4.LAST["/tmp/golden_tmp.nc"]


def describe(a):
    a = np.ascontiguousarray(a, dtype=np.float64)
    return {"shape": list(a.shape), "sha256": hashlib.sha256(a.tobytes()).hexdigest(),
            "sum": float(np.sum(a)), "min": float(a.min()), "max": float(a.max())}


def make_configs(which):
    R = quiet(H.load_reference)
    path = os.path.join(GOLD, "configs.json")
    out = json.load(open(path)) if os.path.exists(path) else {}
    for name in which:
        t0 = time.time()
        v = run_config(R, CONFIGS[name])
        out[name] = {"flags": CONFIGS[name], "arrays": {k: describe(v[k]) for k in ("x", "y", "dx", "dy", "area", "angle_dx")}}
        # the reference's own stdout for this recipe: the CHECK_metrics self-check lines (OGG:1016, 1069, 1111, 1156, 1172),
        # the plan messages and its wall time (`runtime(secs)`, OGG:1449; strict numpy loops, one core)
        lines = LAST_STDOUT[0].splitlines()
        out[name]["stdout_checks"] = [ln for ln in lines if "CHECK_metrics" in ln]
        out[name]["stdout_plan"] = [ln for ln in lines if ln.startswith("   Cutting") or "need to be cut" in ln
                                    or "No SC grid remained" in ln or ln.startswith("   Equator is at") or ln.startswith("shapes:")]
        out[name]["o1_runtime_secs"] = [float(ln.split()[-1]) for ln in lines if ln.startswith("runtime(secs)")]
        print(name, out[name]["arrays"]["x"]["shape"], "%.1fs" % (time.time() - t0), flush=True)
        json.dump(out, open(path, "w"), indent=1, sort_keys=True)


def make_functions():
    R = quiet(H.load_reference)
    rng = np.random.default_rng(20261104)
    G = {}
    Ni, Nj, lat0, lon_bp = 48, 10, 64.05895973, -300.0
    rp = float(R.np.tan(0.5 * (90 - lat0) * R.PI_180))
    G["bp_params"] = np.array([Ni, Nj, lat0, lon_bp, rp])
    for k, a in zip(("bp_lams", "bp_phis", "bp_hi", "bp_hj"), quiet(R.generate_bipolar_cap_mesh, Ni, Nj, lat0, lon_bp)):
        G[k] = a
    for order in (2, 3, 4, 5):
        for k, a in zip(("dxq", "dyq", "daq"), quiet(R.bipolar_cap_metrics_quad_fast, order, 12, 5, lat0, lon_bp, rp)):
            G["bpq%d_%s" % (order, k)] = a
    lamg = lon_bp + rng.uniform(0, 360, (6, 21)); lamg[0, :5] = lon_bp + np.array([0, 90, 180, 270, 360.0])
    phig = rng.uniform(lat0, 90, (6, 21)); phig[1, :3] = 90.0
    G["bpp_lamg"], G["bpp_phig"] = lamg, phig
    for k, a in zip(("lams", "phis", "hi", "hj"), R.bipolar_projection(lamg, phig, lon_bp, rp)):
        G["bpp_" + k] = a
    i = np.array([0.0, 0.25, 1.0, 7.5, 11.999, 12.0]); j = np.array([0.0, 0.5, 3.3, 4.999, 5.0])
    G["bpij_i"], G["bpij_j"] = i, j
    G["bpij_hi"], G["bpij_hj"] = R.bipolar_cap_ij_array(i, j, 12, 5, lat0, lon_bp, rp)
    # displaced pole
    dpar = (64, 12, -300.0, -78.0, 80.0, 0.2)
    G["dp_params"] = np.array(dpar)
    x, y, londp, latdp = quiet(R.generate_displaced_pole_grid, *dpar)
    G["dp_x"], G["dp_y"], G["dp_pole"] = x, y, np.array([londp, latdp])
    for order in (2, 4):
        for k, a in zip(("dxq", "dyq", "daq"), quiet(R.displacedPoleCap_metrics_quad, order, 16, 6, *dpar[2:])):
            G["dpq%d_%s" % (order, k)] = a
    jj = np.arange(0, 13) * 1.0; ii = np.arange(0, 65) * 1.0
    for order in (2, 4, 6):
        G["dp_hi%d" % order] = R.numerical_hi(jj, ii, *dpar, 1e-3, order)
        G["dp_hj%d" % order] = R.numerical_hj(jj, ii, *dpar, 1e-3, order)
    G["dp_gad"] = R.great_arc_distance(jj, ii + 0.5, jj, ii - 0.25, *dpar)
    r_joint = R.np.tan((90 + -78.0) * R.PI_180)
    z_0 = 0.2 * (R.np.cos(80.0 * R.PI_180) + 1j * R.np.sin(80.0 * R.PI_180))
    G["dp_consts"] = np.array([r_joint, z_0.real, z_0.imag])
    lon = np.tile(np.sort(rng.uniform(-300, 60, 33)), (7, 1)); lat = np.tile(np.sort(rng.uniform(-90, -78, 7)).reshape(7, 1), (1, 33))
    G["dpp_lon"], G["dpp_lat"] = lon, lat
    G["dpp_lam"], G["dpp_phi"] = R.displacedPoleCap_projection(lon, lat, z_0, r_joint)
    mb = rng.uniform(-180, 180, (5, 40)); G["mb_in"] = mb.copy(); G["mb_out"] = R.monotonic_bounding(mb.copy(), -300.0)
    # bands and general metrics
    lam, phi = quiet(R.generate_latlon_grid, 48, 9, -300., 360., -78., 11.3, ensure_nj_even=True)
    G["ll_x"], G["ll_y"] = lam, phi
    meshes = {"ll": (lam, phi), "bp": (G["bp_lams"], G["bp_phis"]), "dp": (x, y)}
    for nm, (mx, my) in meshes.items():
        for fix in (True, False):
            for k, a in zip(("dx", "dy", "area"), R.generate_grid_metrics_MIDAS(mx, my, latlon_areafix=fix)):
                G["midas_%s_%d_%s" % (nm, fix, k)] = a
        G["angle_" + nm] = R.angle_x(mx, my)
    for r_, ps, pn, ee in [(0.25, -66.85954725, 64.05895973, 0), (1, -66.85954725, 64.05895973, 0), (2, -68.0, 65.0, 0),
                           (2, -68.0, 65.0, 4), (1, -77.8, 60.0, 1), (4, -66.85954725, 64.05895973, 8)]:
        Ni_ = int(r_ * 2 * 360)
        _, yy = quiet(R.generate_mercator_grid, Ni_, ps, pn, -300., 360, r_, ensure_nj_even=False, enhanced_equatorial=ee)
        G["merc_phi_r%s_e%d" % (r_, ee)] = yy[:, 0].copy()
    # elementary functions on fixed arguments (pins the numpy twin itself)
    t = np.concatenate([rng.uniform(-7, 7, 400), [0.0, -0.0, np.pi / 2, 90 * (np.pi / 180), 1e-9, -1e-300]])
    G["trig_arg"] = t
    for f in ("sin", "cos", "tan", "arctan", "sinh", "cosh", "exp"):
        G["trig_" + f] = getattr(R.np, f)(t)
    u = np.concatenate([rng.uniform(-1, 1, 400), [0.0, 1.0, -1.0, 0.5, -0.5]])
    G["trig_uarg"] = u
    G["trig_arcsin"], G["trig_arccos"] = R.np.arcsin(u), R.np.arccos(u)
    G["trig_log"] = R.np.log(np.abs(t) + 1e-3)
    G["trig_arctan2"] = R.np.arctan2(t[:405], u)
    np.savez_compressed(os.path.join(GOLD, "functions.npz"), **G)
    print("functions.npz:", len(G), "arrays", sum(a.nbytes for a in G.values()), "bytes")


if __name__ == "__main__":
    assert H.strict_numpy(), "numpy is not running its strict loops"
    os.makedirs(GOLD, exist_ok=True)
    if "--functions" in sys.argv:
        make_functions()
    if "--configs" in sys.argv:
        w = sys.argv[sys.argv.index("--configs") + 1]
        make_configs(QUICK if w == "quick" else (list(CONFIGS) if w == "all" else w.split(",")))

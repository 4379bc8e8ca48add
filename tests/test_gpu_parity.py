"""Parity tests proper: every C-ABI entry point (called through the drop-in Python mirror) against the CPU
oracle on the same inputs and against the golden vectors produced by the unmodified reference.  Bar: bit-exact."""
import numpy as np
import pytest

from util import assert_bits, sha

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def G():
    from ocean_model_grid_generator_b200 import device, ocean_grid_generator
    device.ready()
    return ocean_grid_generator


@pytest.fixture(scope="module")
def dnp():
    from ocean_model_grid_generator_b200 import devnp
    return devnp


UNARY = ["sin", "cos", "tan", "arcsin", "arccos", "arctan", "exp", "log", "sinh", "cosh"]


@pytest.mark.parametrize("name", UNARY)
def test_trig_device_equals_oracle(dnp, oracle, name):
    rng = np.random.default_rng(17)
    lo, hi = {"arcsin": (-1, 1), "arccos": (-1, 1), "log": (1e-6, 1e6), "exp": (-30, 30), "sinh": (-6, 6),
              "cosh": (-6, 6), "tan": (-7, 7)}.get(name, (-400, 400))
    x = np.concatenate([rng.uniform(lo, hi, 1000000), rng.uniform(lo, hi, 100000) * 1e-4,
                        [0.0, -0.0, np.pi / 2, -np.pi / 2, 90 * (np.pi / 180), 0.4375, 0.6875, 1.1875, 2.4375, 1e-300]])
    if name == "log":
        x = np.abs(x) + 1e-300
    if name in ("arcsin", "arccos"):
        x = np.clip(x, -1, 1)
        x[:6] = [0.0, 1.0, -1.0, 0.5, -0.5, -0.0]
    assert_bits(getattr(dnp, name)(x), oracle.unary(name, x), name)


@pytest.mark.parametrize("mode", [0, 1, 2, 3, 4])
def test_fast_division_and_sqrt_are_ieee(G, mode):
    """The cap kernels replace the compiler's guarded-and-branching IEEE division / square root by the same
    instruction sequence with operand-range guards and a per-sample redo (MathFast, csrc/ogg_trig.cuh).  Every result
    the guards accept must be the IEEE result, bit for bit; in-range operands must not be rejected."""
    import ctypes
    from ocean_model_grid_generator_b200 import _cabi
    lib = _cabi.load()
    total = 0
    for seed in (1, 2, 3):
        counts = (ctypes.c_uint64 * 3)()
        n = 1 << 27 if mode != 4 else 1 << 25
        assert lib.ogg_selftest_divsqrt(n, seed * 0x9e3779b9, mode, counts) == 0
        accepted, differing, rejected = counts[0], counts[1], counts[2]
        assert differing == 0, "mode %d seed %d: %d of %d accepted results differ from IEEE" % (mode, seed, differing, accepted)
        assert rejected == 0, "mode %d: %d in-range operand sets rejected" % (mode, rejected)
        total += accepted
        if mode in (0, 2, 3):
            assert accepted == 5 * n          # everything generated in range is accepted
        elif mode == 1:
            assert 0 < accepted < 5 * n       # raw bit patterns: the guards reject a share and accept a share
        else:
            assert accepted > 4 * n           # trig: only exact zeros / out-of-range quotients are redone
    assert total > 0


def test_elided_guards_audit(G):
    """The cap kernels skip operand-range guards that follow from an earlier guard of the same unit (ELIDE, ZEROOPS in
    csrc/ogg_trig.cuh / dp_vertex_fast).  A library built with -DOGG_GUARD_AUDIT evaluates every skipped claim while it
    builds a set of supergrids and whole caps: no unit that passed its guards may have a false claim.  The shipped
    library reports {0, 0} (its counters are compiled out)."""
    import ctypes, json, os, shutil, subprocess, sys
    from ocean_model_grid_generator_b200 import _cabi
    counts = (ctypes.c_uint64 * 2)(7, 7)
    assert _cabi.load().ogg_guard_audit(counts, 0) == 0 and counts[0] == 0 and counts[1] == 0
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    lib = os.path.join(root, "build", "libogg_audit.so")
    src = [os.path.join(root, "ocean_model_grid_generator_b200", "csrc", f) for f in ("ogg_b200.cu", "ogg_trig.cuh")]
    if not os.path.exists(lib) or any(os.path.getmtime(f) > os.path.getmtime(lib) for f in src):
        if shutil.which("nvcc") is None and not os.path.exists("/usr/local/cuda/bin/nvcc"):
            pytest.skip("no nvcc to build the audit variant")
        env = dict(os.environ, PATH=os.environ.get("PATH", "") + ":/usr/local/cuda/bin")
        subprocess.check_call(["bash", os.path.join(root, "tools", "build_variant.sh"), "audit", "-DOGG_GUARD_AUDIT"], cwd=root, env=env)
    env = dict(os.environ, OGG_B200_LIB=lib)
    out = subprocess.run([sys.executable, os.path.join(root, "tools", "guard_audit.py")], cwd=root, env=env,
                         capture_output=True, text=True, timeout=900)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    last = json.loads(out.stdout.strip().splitlines()[-1])
    assert last["audit_build"] and last["units_passed"] > 10**8 and last["claims_false"] == 0, last


def test_arctan2_and_mod(dnp, oracle):
    rng = np.random.default_rng(18)
    y, x = rng.uniform(-2, 2, 500000), rng.uniform(-2, 2, 500000)
    y[:100] = 0.0; x[50:150] = 0.0; y[200:210] = -0.0
    assert_bits(dnp.arctan2(y, x), oracle.arctan2(y, x), "arctan2")
    a = np.concatenate([rng.uniform(-1000, 1000, 500000), [0.0, -0.0, 360.0, -360.0, 720.0, 359.99999999999994,
                       -1e-20, 1e-20, 360.00000000000006, -359.99999999999994, 180.0, -180.0]])
    assert_bits(dnp.mod360(a), oracle.mod360(a), "mod360")
    assert_bits(dnp.mod360(a), np.mod(a, 360.0), "mod360 vs numpy")
    assert float(dnp.tan(0.5)) == float(oracle.unary("tan", np.array([0.5]))[0])      # scalar path


def test_trig_golden(dnp, golden):
    t, u = golden["trig_arg"], golden["trig_uarg"]
    for f in ("sin", "cos", "tan", "arctan", "sinh", "cosh", "exp"):
        assert_bits(getattr(dnp, f)(t), golden["trig_" + f], f)
    assert_bits(dnp.arcsin(u), golden["trig_arcsin"], "arcsin")
    assert_bits(dnp.arccos(u), golden["trig_arccos"], "arccos")
    assert_bits(dnp.log(np.abs(t) + 1e-3), golden["trig_log"], "log")
    assert_bits(dnp.arctan2(t[:405], u), golden["trig_arctan2"], "arctan2")


# ---- bipolar cap -----------------------------------------------------------------------------------
def test_bipolar_golden(G, golden):
    Ni, Nj, lat0, lon_bp, rp = golden["bp_params"]
    Ni, Nj = int(Ni), int(Nj)
    assert G.bipolar_rp(lat0) == rp
    for g, k in zip(G.generate_bipolar_cap_mesh(Ni, Nj, lat0, lon_bp), ("bp_lams", "bp_phis", "bp_hi", "bp_hj")):
        assert_bits(g, golden[k], k)
    for order in (2, 3, 4, 5):
        for g, k in zip(G.bipolar_cap_metrics_quad_fast(order, 12, 5, lat0, lon_bp, rp), ("dxq", "dyq", "daq")):
            assert_bits(g, golden["bpq%d_%s" % (order, k)], "bpq%d %s" % (order, k))
    for g, k in zip(G.bipolar_projection(golden["bpp_lamg"], golden["bpp_phig"], lon_bp, rp), ("lams", "phis", "hi", "hj")):
        assert_bits(g, golden["bpp_" + k], "bipolar_projection " + k)
    hi, hj = G.bipolar_projection(golden["bpp_lamg"], golden["bpp_phig"], lon_bp, rp, metrics_only=True)
    assert_bits(hi, golden["bpp_hi"], "mo hi"); assert_bits(hj, golden["bpp_hj"], "mo hj")
    hi, hj = G.bipolar_cap_ij_array(golden["bpij_i"], golden["bpij_j"], 12, 5, lat0, lon_bp, rp)
    assert_bits(hi, golden["bpij_hi"], "ij hi"); assert_bits(hj, golden["bpij_hj"], "ij hj")


@pytest.mark.parametrize("Ni,Nj,lat0", [(96, 20, 64.05895973), (360, 61, 65.0), (2880, 17, 60.0), (24, 3, 80.0)])
def test_bipolar_vs_oracle(G, oracle, Ni, Nj, lat0):
    lon_bp = -300.0
    for g, r, n in zip(G.generate_bipolar_cap_mesh(Ni, Nj, lat0, lon_bp, ensure_nj_even=False),
                       oracle.generate_bipolar_cap_mesh(Ni, Nj, lat0, lon_bp, ensure_nj_even=False), "lphj"):
        assert_bits(g, r, "mesh " + n)
    rp = G.bipolar_rp(lat0)
    nyq = min(Nj, 9)
    for order in (2, 3, 4, 5):
        for g, r, n in zip(G.bipolar_cap_metrics_quad_fast(order, Ni // 4, nyq, lat0, lon_bp, rp),
                           oracle.bipolar_cap_metrics_quad_fast(order, Ni // 4, nyq, lat0, lon_bp, rp), ("dxq", "dyq", "daq")):
            assert_bits(g, r, "quad%d %s" % (order, n))
    with pytest.raises(Exception, match="Uncoded order"):
        G.bipolar_cap_metrics_quad_fast(6, 8, 4, lat0, lon_bp, rp)


def test_bipolar_row_slabs_match_whole(G):
    """Row-slab decomposition (what each GPU computes) reproduces the whole-cap result bit for bit."""
    from ocean_model_grid_generator_b200 import _cabi, device as D
    nx, ny, lat0, lon_bp = 64, 13, 64.05895973, -300.0
    rp = G.bipolar_rp(lat0)
    whole = G.bipolar_cap_metrics_quad_fast(5, nx, ny, lat0, lon_bp, rp)
    for (j0, j1) in [(0, 5), (5, 6), (6, 13)]:
        top = 1 if j1 == ny else 0
        dxq, dyq, daq = D.empty(j1 - j0 + top, nx), D.empty(j1 - j0, nx + 1), D.empty(j1 - j0, nx)
        _cabi.call("ogg_bipolar_cap_metrics_quad", 5, nx, ny, j0, j1, lat0, lon_bp, rp, 6371.0e3, 0, D.ptr(dxq), D.ptr(dyq),
                   D.ptr(daq), D.stream())
        assert_bits(D.to_host(dxq), whole[0][j0:j1 + top], "dxq slab")
        assert_bits(D.to_host(dyq), whole[1][j0:j1], "dyq slab")
        assert_bits(D.to_host(daq), whole[2][j0:j1], "daq slab")


# ---- displaced-pole cap ------------------------------------------------------------------------------
def test_displaced_pole_golden(G, golden):
    ni, nj, lon0, lat0, lon_dp, r_dp = golden["dp_params"]
    ni, nj = int(ni), int(nj)
    x, y, londp, latdp = G.generate_displaced_pole_grid(ni, nj, lon0, lat0, lon_dp, r_dp)
    assert_bits(x, golden["dp_x"], "dp x"); assert_bits(y, golden["dp_y"], "dp y")
    assert_bits(np.array([londp, latdp]), golden["dp_pole"], "pole")
    for order in (2, 4):
        for g, k in zip(G.displacedPoleCap_metrics_quad(order, 16, 6, lon0, lat0, lon_dp, r_dp), ("dxq", "dyq", "daq")):
            assert_bits(g, golden["dpq%d_%s" % (order, k)], "dpq%d %s" % (order, k))
    jj, ii = np.arange(0, 13) * 1.0, np.arange(0, 65) * 1.0
    for order in (2, 4, 6):
        assert_bits(G.numerical_hi(jj, ii, ni, nj, lon0, lat0, lon_dp, r_dp, 1e-3, order), golden["dp_hi%d" % order], "hi")
        assert_bits(G.numerical_hj(jj, ii, ni, nj, lon0, lat0, lon_dp, r_dp, 1e-3, order), golden["dp_hj%d" % order], "hj")
    assert_bits(G.great_arc_distance(jj, ii + 0.5, jj, ii - 0.25, ni, nj, lon0, lat0, lon_dp, r_dp), golden["dp_gad"], "gad")
    rj, z0 = G.dp_constants(lat0, lon_dp, r_dp)
    assert_bits(np.array([rj, z0.real, z0.imag]), golden["dp_consts"], "dp consts")
    lam, phi = G.displacedPoleCap_projection(golden["dpp_lon"], golden["dpp_lat"], z0, rj)
    assert_bits(lam, golden["dpp_lam"], "projection lam"); assert_bits(phi, golden["dpp_phi"], "projection phi")
    assert_bits(G.monotonic_bounding(golden["mb_in"].copy(), -300.0), golden["mb_out"], "monotonic_bounding")
    with pytest.raises(Exception, match="order not coded"):
        G.displacedPoleCap_metrics_quad(5, 8, 4, lon0, lat0, lon_dp, r_dp)
    with pytest.raises(Exception, match="Uncoded order"):
        G.displacedPoleCap_metrics_quad(7, 8, 4, lon0, lat0, lon_dp, r_dp)
    with pytest.raises(Exception, match="order not coded"):
        G.numerical_hi(jj, ii, ni, nj, lon0, lat0, lon_dp, r_dp, 1e-3, 3)


@pytest.mark.parametrize("ni,nj,lon_dp,r_dp", [(96, 14, 80.0, 0.2), (720, 35, 60.0, 0.35), (1440, 10, 80.0, 0.2),
                                               (300, 7, -120.0, 0.05)])
def test_displaced_pole_vs_oracle(G, oracle, ni, nj, lon_dp, r_dp):
    lon0, lat0 = -300.0, -78.0
    x, y, _, _ = G.generate_displaced_pole_grid(ni, nj, lon0, lat0, lon_dp, r_dp)
    xo, yo, _, _ = oracle.generate_displaced_pole_grid(ni, nj, lon0, lat0, lon_dp, r_dp)
    assert_bits(x, xo, "dp x"); assert_bits(y, yo, "dp y")
    nxq, nyq = min(ni, 300), min(nj, 8)
    for order in (2, 4):
        for g, r, n in zip(G.displacedPoleCap_metrics_quad(order, nxq, nyq, lon0, lat0, lon_dp, r_dp),
                           oracle.displacedPoleCap_metrics_quad(order, nxq, nyq, lon0, lat0, lon_dp, r_dp), ("dxq", "dyq", "daq")):
            assert_bits(g, r, "dp quad%d %s" % (order, n))


@pytest.mark.parametrize("lat0,ni,nj", [(-20.0, 300, 12), (30.0, 520, 9), (-44.0, 64, 40), (80.0, 96, 6)])
def test_dp_quad_haversine_quadrants(G, oracle, lat0, ni, nj):
    """k_dp_hsamples' haversine has a straight-line path for latitudes in an odd quadrant of pi/2 (every real southern
    cap: -90..-45 degrees) and the general code for the rest.  Caps reaching up to -44, -20, +30 and +80 degrees put
    latitude pairs on both sides of the quadrant boundaries at -45 and +45 (mixed pairs take the general code); 300 and
    520 columns span two and three look-back tiles, so the unwrap fast paths see tiles with and without the jump."""
    lon0, lon_dp, r_dp = -300.0, 80.0, 0.2
    for order in (2, 4):
        for g, r, n in zip(G.displacedPoleCap_metrics_quad(order, ni, nj, lon0, lat0, lon_dp, r_dp),
                           oracle.displacedPoleCap_metrics_quad(order, ni, nj, lon0, lat0, lon_dp, r_dp), ("dxq", "dyq", "daq")):
            assert_bits(g, r, "dp quad%d lat0=%g %s" % (order, lat0, n))


def test_random_caps_vs_oracle(G, oracle):
    """Seeded random cap geometries through both fused quadratures and both cap meshes against the oracle, bit for bit.
    The draws cover what the kernels specialise on: southern caps that take the first-range arctan / first Smith branch
    (lat0 < -67, r_dp <= 0.5) and caps that do not, poles displaced almost to the cap edge, bipolar caps starting at 40 to
    85 degrees (latitudes on both sides of quadrant 1 of the cosine path), column counts around the 255-column look-back
    tile and the 32-cell bipolar tile, every coded order."""
    rng = np.random.default_rng(20261105)
    for trial in range(10):
        order = int(rng.choice([2, 4]))
        ni = int(rng.choice([8, 31, 85, 86, 170, 171, 300]))
        nj = int(rng.integers(2, 14))
        lon0 = float(rng.uniform(-360, 360))
        lat0 = float(rng.choice([-80.0, -67.0, -66.9, -45.0, 10.0, 75.0])) + float(rng.uniform(-1, 1)) * (trial % 2)
        lon_dp = float(rng.uniform(-180, 180))
        r_dp = float(rng.choice([0.01, 0.2, 0.5, 0.51, 0.95]))
        what = "dp trial %d: order %d ni %d nj %d lon0 %r lat0 %r lon_dp %r r_dp %r" % (trial, order, ni, nj, lon0, lat0, lon_dp, r_dp)
        for g, r, n in zip(G.displacedPoleCap_metrics_quad(order, ni, nj, lon0, lat0, lon_dp, r_dp),
                           oracle.displacedPoleCap_metrics_quad(order, ni, nj, lon0, lat0, lon_dp, r_dp), ("dxq", "dyq", "daq")):
            assert_bits(g, r, what + " " + n)
        for g, r, n in zip(G.generate_displaced_pole_grid(ni, nj, lon0, lat0, lon_dp, r_dp)[:2],
                           oracle.generate_displaced_pole_grid(ni, nj, lon0, lat0, lon_dp, r_dp)[:2], ("x", "y")):
            assert_bits(g, r, what + " mesh " + n)
    for trial in range(10):
        order = int(rng.integers(2, 6))
        ni = int(rng.choice([8, 31, 32, 33, 96, 130]))
        nj = int(rng.integers(2, 16))
        lat0 = float(rng.choice([40.0, 46.0, 64.05895973, 85.0])) + float(rng.uniform(-1, 1)) * (trial % 2)
        lon_bp = float(rng.uniform(-360, 360))
        rp = G.bipolar_rp(lat0)
        assert rp == oracle.bipolar_rp(lat0)
        what = "bp trial %d: order %d ni %d nj %d lat0 %r lon_bp %r" % (trial, order, ni, nj, lat0, lon_bp)
        for g, r, n in zip(G.bipolar_cap_metrics_quad_fast(order, ni, nj, lat0, lon_bp, rp),
                           oracle.bipolar_cap_metrics_quad_fast(order, ni, nj, lat0, lon_bp, rp), ("dxq", "dyq", "daq")):
            assert_bits(g, r, what + " " + n)
        for g, r, n in zip(G.generate_bipolar_cap_mesh(ni, nj, lat0, lon_bp),
                           oracle.generate_bipolar_cap_mesh(ni, nj, lat0, lon_bp), ("lams", "phis", "h_i_inv", "h_j_inv")):
            assert_bits(g, r, what + " mesh " + n)


def test_monotonic_bounding_adversarial(G, oracle):
    """Random jumps exercise every branch of the 2-state scan, including chains where the shift state flips often,
    rows longer than one scan tile and values sitting on the 100-degree threshold."""
    rng = np.random.default_rng(23)
    for shape in [(3, 7), (5, 256), (4, 257), (6, 1000), (2, 5000)]:
        x = rng.uniform(-200, 200, shape)
        x[0, : min(shape[1], 6)] = [-300.0, -200.0, -100.0, 0.0, 100.0, 160.0][: min(shape[1], 6)]
        x[1] = np.round(x[1] / 50.0) * 50.0              # many differences exactly 100, 150, ...
        ref = oracle.monotonic_bounding(x.copy(), -300.0)
        got = G.monotonic_bounding(x.copy(), -300.0)
        assert_bits(got, ref, "monotonic %s" % (shape,))


def test_dp_row_slabs_match_whole(G):
    from ocean_model_grid_generator_b200 import _cabi, device as D
    nx, ny = 120, 11
    args = (-300.0, -78.0, 80.0, 0.2)
    whole = G.displacedPoleCap_metrics_quad(4, nx, ny, *args)
    for (j0, j1) in [(0, 4), (4, 5), (5, 11)]:
        top = 1 if j1 == ny else 0
        dxq, dyq, daq = D.empty(j1 - j0 + top, nx), D.empty(j1 - j0, nx + 1), D.empty(j1 - j0, nx)
        _cabi.call("ogg_dp_metrics_quad", 4, nx, ny, j0, j1, *args, 6371.0e3, 0, D.ptr(dxq), D.ptr(dyq), D.ptr(daq), D.stream())
        assert_bits(D.to_host(dxq), whole[0][j0:j1 + top], "dxq slab")
        assert_bits(D.to_host(dyq), whole[1][j0:j1], "dyq slab")
        assert_bits(D.to_host(daq), whole[2][j0:j1], "daq slab")


# ---- bands / spherical metrics ---------------------------------------------------------------------------
def test_bands_golden(G, golden):
    lam, phi = G.generate_latlon_grid(48, 9, -300., 360., -78., 11.3, ensure_nj_even=True)
    assert_bits(lam, golden["ll_x"], "latlon x"); assert_bits(phi, golden["ll_y"], "latlon y")
    meshes = {"ll": (golden["ll_x"], golden["ll_y"]), "bp": (golden["bp_lams"], golden["bp_phis"]),
              "dp": (golden["dp_x"], golden["dp_y"])}
    for nm, (mx, my) in meshes.items():
        for fix in (True, False):
            for g, k in zip(G.generate_grid_metrics_MIDAS(mx, my, latlon_areafix=fix), ("dx", "dy", "area")):
                assert_bits(g, golden["midas_%s_%d_%s" % (nm, fix, k)], "midas %s %d %s" % (nm, fix, k))
        assert_bits(G.angle_x(mx, my), golden["angle_" + nm], "angle " + nm)
    with pytest.raises(Exception, match="same shape"):
        G.angle_x(np.zeros((3, 4)), np.zeros((3, 5)))


def test_mercator_golden(G, golden):
    for r_, ps, pn, ee in [(0.25, -66.85954725, 64.05895973, 0), (1, -66.85954725, 64.05895973, 0), (2, -68.0, 65.0, 0),
                           (2, -68.0, 65.0, 4), (1, -77.8, 60.0, 1), (4, -66.85954725, 64.05895973, 8)]:
        Ni = int(r_ * 720)
        x, y = G.generate_mercator_grid(Ni, ps, pn, -300., 360, r_, ensure_nj_even=False, enhanced_equatorial=ee)
        assert_bits(y[:, 0], golden["merc_phi_r%s_e%d" % (r_, ee)], "mercator phi r=%s ee=%d" % (r_, ee))
        assert (y == y[:, :1]).all() and x.shape == (y.shape[0], Ni + 1)
        assert_bits(x[0], -300. + np.arange(Ni + 1) * 360 / float(Ni), "mercator lam")


def test_mercator_laws_fused_equal_chain(G, dnp, oracle):
    """y_mercator / phi_mercator (OGG:292-301) as one launch each: same bits as the reference's chain of element-wise
    calls (the device functions one by one, numpy's IEEE operations in between) and as the CPU oracle."""
    rng = np.random.default_rng(41)
    for Ni in (180, 2880, 21600):
        R = Ni / (2 * np.pi)
        phi = np.concatenate([rng.uniform(-1.4, 1.4, 5000), [0.0, -66.85954725 * np.pi / 180, 64.05895973 * np.pi / 180]])
        chain = R * dnp.log((1.0 + dnp.sin(phi)) / dnp.cos(phi))
        assert_bits(G.y_mercator(Ni, phi), chain, "y_mercator chain Ni=%d" % Ni)
        assert_bits(G.y_mercator(Ni, phi), oracle.y_mercator(Ni, phi), "y_mercator oracle Ni=%d" % Ni)
        y = np.arange(-int(1.3 * R) - 1, int(1.25 * R) + 1).astype(np.float64)
        chain = dnp.arctan(dnp.sinh(y / R)) * (180 / np.pi)
        assert_bits(G.phi_mercator(Ni, y), chain, "phi_mercator chain Ni=%d" % Ni)
        assert_bits(G.phi_mercator(Ni, y), oracle.phi_mercator(Ni, y), "phi_mercator oracle Ni=%d" % Ni)
        assert_bits(G._mercator_law("ogg_phi_mercator", None, R, y_first=y[0], n=y.size), chain, "phi_mercator arange")
        assert float(G.phi_mercator(Ni, 3.0)) == float(chain[np.where(y == 3.0)[0][0]])        # scalar path


@pytest.mark.parametrize("shape", [(33, 130), (64, 129), (65, 31), (2, 2), (200, 1441)])
def test_midas_general_vs_oracle(G, oracle, shape):
    rng = np.random.default_rng(31)
    nyp, nxp = shape
    x = np.cumsum(rng.uniform(0, 360.0 / nxp * 2, shape), axis=1) - 300.0 + rng.normal(0, 0.01, shape)
    y = np.cumsum(rng.uniform(0, 170.0 / nyp, shape), axis=0) - 85.0
    for fix in (True, False):
        if nyp < 2:
            continue
        for g, r, n in zip(G.generate_grid_metrics_MIDAS(x, y, latlon_areafix=fix),
                           oracle.generate_grid_metrics_MIDAS(x, y, latlon_areafix=fix), ("dx", "dy", "area")):
            assert_bits(g, r, "midas %s fix=%d %s" % (shape, fix, n))
    assert_bits(G.angle_x(x, y), oracle.angle_x(x, y), "angle")


def test_midas_halo_row_equals_whole(G, oracle):
    """Row-slab MIDAS with the next slab's first row as halo == the whole-mesh result (the multi-GPU stencil)."""
    rng = np.random.default_rng(37)
    x = np.cumsum(rng.uniform(0, 1, (41, 77)), axis=1) - 300.0
    y = np.cumsum(rng.uniform(0, 1, (41, 77)), axis=0) - 60.0
    for fix in (True, False):
        whole = oracle.generate_grid_metrics_MIDAS(x, y, latlon_areafix=fix)
        for (a, b) in [(0, 17), (17, 18), (18, 41)]:
            halo = (x[b], y[b]) if b < 41 else None
            dx, dy, area = G.generate_grid_metrics_MIDAS(x[a:b], y[a:b], latlon_areafix=fix, halo=halo)
            assert_bits(dx, whole[0][a:b], "dx"); nb = b if halo is not None else b - 1
            assert_bits(dy, whole[1][a:nb], "dy"); assert_bits(area, whole[2][a:nb], "area")


def test_band_fused_equals_general(G, oracle):
    """The fused band kernel (tile + metrics + angle) against the general kernels and the oracle on the same band."""
    from ocean_model_grid_generator_b200 import _cabi, device as D
    lam = -300.0 + np.arange(1441) * 360 / 1440.0
    phi = np.concatenate([[-90.0], np.sort(np.random.default_rng(3).uniform(-89.9, 89.9, 98)), [90.0]])
    nyp, nxp = phi.size, lam.size
    tl, tp = D.to_dev(lam), D.to_dev(phi)
    o = [D.empty(nyp, nxp), D.empty(nyp, nxp), D.empty(nyp, nxp - 1), D.empty(nyp - 1, nxp), D.empty(nyp - 1, nxp - 1), D.empty(nyp, nxp)]
    _cabi.call("ogg_band_fill_metrics", D.ptr(tl), D.ptr(tp), nyp, nxp, 6371.0e3, 0, *[D.ptr(t) for t in o], D.stream())
    x, y = np.tile(lam, (nyp, 1)), np.tile(phi.reshape(-1, 1), (1, nxp))
    assert_bits(D.to_host(o[0]), x, "x"); assert_bits(D.to_host(o[1]), y, "y")
    dx, dy, area = oracle.generate_grid_metrics_MIDAS(x, y)
    assert_bits(D.to_host(o[2]), dx, "dx"); assert_bits(D.to_host(o[3]), dy, "dy"); assert_bits(D.to_host(o[4]), area, "area")
    assert_bits(D.to_host(o[5]), oracle.angle_x(x, y), "angle")


def test_angle_x_zero_differences_take_the_ieee_branch(G, oracle):
    """angle_x on meshes whose centred differences are exact zeros -- along a parallel (dy == 0), along a meridian
    (dx == 0), both, and mixed with ordinary vertices inside one warp: k_angle sends those vertices straight to the IEEE
    expression instead of through the fast path first; the results are the oracle's either way."""
    rng = np.random.default_rng(11)
    nyp, nxp = 37, 131
    lam = -300.0 + np.arange(nxp) * 360.0 / (nxp - 1)
    phi = np.linspace(-80.0, 80.0, nyp)
    x, y = np.tile(lam, (nyp, 1)), np.tile(phi.reshape(-1, 1), (1, nxp))
    assert_bits(G.angle_x(x, y), oracle.angle_x(x, y), "parallel")                  # dy == 0 everywhere
    xm = np.tile(rng.uniform(-300, 60, (nyp, 1)), (1, nxp))                         # dx == 0 everywhere
    ym = y + rng.uniform(-1, 1, (nyp, nxp))
    assert_bits(G.angle_x(xm, ym), oracle.angle_x(xm, ym), "meridian")
    assert_bits(G.angle_x(xm, y), oracle.angle_x(xm, y), "both zero")
    mix = rng.random((nyp, nxp)) < 0.5                                               # per-vertex mixture
    x2 = np.where(mix, x, x + rng.uniform(-0.3, 0.3, (nyp, nxp)))
    y2 = np.where(mix, y, ym)
    assert_bits(G.angle_x(x2, y2), oracle.angle_x(x2, y2), "mixed")


def _band_outputs(D, nyp, nxp):
    return [D.empty(nyp, nxp), D.empty(nyp, nxp), D.empty(nyp, nxp - 1), D.empty(nyp - 1, nxp), D.empty(nyp - 1, nxp - 1), D.empty(nyp, nxp)]


def test_band_batch_equals_single_calls(G):
    """ogg_band_fill_metrics_batch (every band of a stitch in ONE launch) against one ogg_band_fill_metrics call per band:
    four bands of different shapes (ragged last row tile, a 2-column band, odd/even row lengths), with the
    SKIP_TOP_VERTEX_ROW and NO_METRICS flags, output buffers pre-filled with a sentinel so that a row the flags
    exclude must come back untouched."""
    import ctypes
    from ocean_model_grid_generator_b200 import _cabi, device as D
    rng = np.random.default_rng(11)
    shapes = [(37, 721), (9, 2), (130, 1442), (1, 300)]
    flags = [0, 1, 2, 0]          # OGG_BAND_SKIP_TOP_VERTEX_ROW = 1, OGG_BAND_NO_METRICS = 2 (include/ogg_b200.h)
    bands = (_cabi.OggBand * len(shapes))()
    keep, single = [], []
    for k, ((nyp, nxp), fl) in enumerate(zip(shapes, flags)):
        lam = D.to_dev(-300.0 + np.sort(rng.uniform(0, 360, nxp)))
        phi = D.to_dev(np.sort(rng.uniform(-80, 80, nyp)))
        a, b = _band_outputs(D, nyp, nxp), _band_outputs(D, nyp, nxp)
        for t in a + b:
            t.fill_(-777.0)
        _cabi.call("ogg_band_fill_metrics", D.ptr(lam), D.ptr(phi), nyp, nxp, 6371.0e3, fl, *[D.ptr(t) for t in a], D.stream())
        bd = bands[k]
        bd.lam, bd.phi, bd.nyp, bd.nxp, bd.flags = D.ptr(lam).value, D.ptr(phi).value, nyp, nxp, fl
        bd.x, bd.y, bd.dx, bd.dy, bd.area, bd.angle = [D.ptr(t).value for t in b]
        keep.append((lam, phi)); single.append((a, b))
    _cabi.call("ogg_band_fill_metrics_batch", bands, len(shapes), 6371.0e3, D.stream())
    for k, (a, b) in enumerate(single):
        for name, u, v in zip(("x", "y", "dx", "dy", "area", "angle"), a, b):
            assert_bits(D.to_host(v), D.to_host(u), "band %d %s" % (k, name))
    # flags did what the header says: the excluded top vertex row / the metrics stayed at the sentinel
    assert (D.to_host(single[1][1][0])[-1] == -777.0).all() and (D.to_host(single[2][1][2]) == -777.0).all()
    assert _cabi.load().ogg_band_fill_metrics_batch(bands, 5, ctypes.c_double(1.0), D.stream()) != 0   # > 4 bands: OGG_ERR_ARG


def test_band_fill_non_finite_input(G, oracle):
    """A NaN or infinite latitude / longitude leaves the band shortcuts (two constants for the angle, |t| for
    the root): the fill must then produce what the general functions produce -- NaN where they do, finite values
    elsewhere (assert_bits treats any two NaNs as equal, as everywhere in this suite)."""
    from ocean_model_grid_generator_b200 import _cabi, device as D
    lam = -300.0 + np.arange(97) * 360 / 96.0
    phi = np.linspace(-60.0, 60.0, 21)
    lam[13] = np.nan; lam[40] = np.inf; phi[5] = np.nan; phi[17] = -np.inf
    nyp, nxp = phi.size, lam.size
    o = _band_outputs(D, nyp, nxp)
    tl, tp = D.to_dev(lam), D.to_dev(phi)
    _cabi.call("ogg_band_fill_metrics", D.ptr(tl), D.ptr(tp), nyp, nxp, 6371.0e3, 0, *[D.ptr(t) for t in o], D.stream())
    x, y = np.tile(lam, (nyp, 1)), np.tile(phi.reshape(-1, 1), (1, nxp))
    with np.errstate(all="ignore"):
        dx, dy, area = oracle.generate_grid_metrics_MIDAS(x, y)
        ang = oracle.angle_x(x, y)
    assert_bits(D.to_host(o[0]), x, "x"); assert_bits(D.to_host(o[1]), y, "y")
    assert_bits(D.to_host(o[2]), dx, "dx"); assert_bits(D.to_host(o[3]), dy, "dy"); assert_bits(D.to_host(o[4]), area, "area")
    assert_bits(D.to_host(o[5]), ang, "angle")
    assert np.isnan(D.to_host(o[5])).any() and np.isfinite(D.to_host(o[5])).any()


def test_quad_helpers(G, oracle):
    rng = np.random.default_rng(41)
    for n in (2, 3, 4, 5):
        a, b = G.quad_positions(n); ao, bo = oracle.quad_positions(n)
        assert_bits(a, ao, "a"); assert_bits(b, bo, "b")
        y = rng.uniform(size=n); y2 = rng.uniform(size=(n, n))
        assert G.quad_average(y) == oracle.quad_average(y)
        assert G.quad_average_2d(y2) == oracle.quad_average_2d(y2)
    with pytest.raises(Exception, match="Uncoded order"):
        G.quad_positions(6)
    with pytest.raises(Exception, match="not squared"):
        G.quad_average_2d(np.zeros((2, 3)))
    with pytest.raises(Exception, match="Uncoded order"):
        G.quad_average(np.zeros(6))


def test_device_sum(G):
    rng = np.random.default_rng(43)
    a = rng.uniform(0, 1, (301, 77))
    assert abs(G.device_sum(a) - a.sum()) <= 1e-12 * a.sum()
    assert abs(G.device_sum(a, n=301, stride=77, offset=5) - a[:, 5].sum()) <= 1e-12 * a[:, 5].sum()


# ---- smallest shapes and the row-ownership flag ---------------------------------------------------------------------
@pytest.mark.parametrize("nx,ny", [(1, 1), (2, 1), (1, 3), (33, 1), (255, 2), (256, 2)])
def test_smallest_caps_vs_oracle(G, oracle, nx, ny):
    """One-cell caps, a single cell row, and column counts straddling the 255-column tile of the look-back kernel."""
    lat0, lon_bp = 64.05895973, -300.0
    rp = G.bipolar_rp(lat0)
    for g, r, n in zip(G.bipolar_cap_metrics_quad_fast(5, nx, ny, lat0, lon_bp, rp),
                       oracle.bipolar_cap_metrics_quad_fast(5, nx, ny, lat0, lon_bp, rp), ("dxq", "dyq", "daq")):
        assert_bits(g, r, "bipolar %dx%d %s" % (nx, ny, n))
    for g, r, n in zip(G.displacedPoleCap_metrics_quad(4, nx, ny, -300.0, -78.0, 80.0, 0.2),
                       oracle.displacedPoleCap_metrics_quad(4, nx, ny, -300.0, -78.0, 80.0, 0.2), ("dxq", "dyq", "daq")):
        assert_bits(g, r, "displaced pole %dx%d %s" % (nx, ny, n))
    for g, r, n in zip(G.generate_displaced_pole_grid(nx, ny, -300.0, -78.0, 80.0, 0.2)[:2],
                       oracle.generate_displaced_pole_grid(nx, ny, -300.0, -78.0, 80.0, 0.2)[:2], "xy"):
        assert_bits(g, r, "dp mesh %dx%d %s" % (nx, ny, n))


def test_quad_top_row_flag_leaves_row_untouched(G):
    """OGG_QUAD_NO_TOP_DX_ROW: the dx row at the cap's top edge belongs to the sub-grid above (OGG:1316-1362 `[:-1]`)."""
    from ocean_model_grid_generator_b200 import _cabi, device as D
    nx, ny = 40, 5
    lat0, lon_bp = 64.05895973, -300.0
    rp = G.bipolar_rp(lat0)
    for name, args in (("ogg_bipolar_cap_metrics_quad", (5, nx, ny, 0, ny, lat0, lon_bp, rp)),
                       ("ogg_dp_metrics_quad", (4, nx, ny, 0, ny, -300.0, -78.0, 80.0, 0.2))):
        ref = [D.full((ny + 1, nx), -7.0), D.empty(ny, nx + 1), D.empty(ny, nx)]
        _cabi.call(name, *args, 6371.0e3, 0, D.ptr(ref[0]), D.ptr(ref[1]), D.ptr(ref[2]), D.stream())
        got = [D.full((ny + 1, nx), -7.0), D.empty(ny, nx + 1), D.empty(ny, nx)]
        _cabi.call(name, *args, 6371.0e3, _cabi.OGG_QUAD_NO_TOP_DX_ROW, D.ptr(got[0]), D.ptr(got[1]), D.ptr(got[2]), D.stream())
        assert bool((got[0][ny] == -7.0).all()) and not bool((ref[0][ny] == -7.0).any())
        assert_bits(D.to_host(got[0][:ny]), D.to_host(ref[0][:ny]), name + " dx")
        assert_bits(D.to_host(got[1]), D.to_host(ref[1]), name + " dy")
        assert_bits(D.to_host(got[2]), D.to_host(ref[2]), name + " area")

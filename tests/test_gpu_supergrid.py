"""Whole-supergrid parity: the device build (plan -> in-place pieces) against the per-array sha256 of the six
arrays the UNMODIFIED reference produced for its own CLI recipes (tests/golden/configs.json), the CLI + file
round trip, row-slab (multi-GPU) decomposition, and size-independent properties at larger resolutions."""
import numpy as np
import pytest

from util import assert_bits, sha

pytestmark = pytest.mark.gpu

NAMES = ("x", "y", "dx", "dy", "area", "angle_dx")


@pytest.fixture(scope="module")
def SG():
    from ocean_model_grid_generator_b200 import device, supergrid
    device.ready()
    return supergrid


def _args(flags):
    from ocean_model_grid_generator_b200 import ocean_grid_generator as G
    return vars(G._build_parser().parse_args(flags.split()))


def _plan_args(a):
    return {k: v for k, v in a.items() if k not in ("gridfilename", "debug", "plotem", "no_changing_meta", "write_subgrid_files")}


# res0.25_full = the benchmark's headline workload `-r 4 --r_dp 0.2`; res0.125 = the reference's own 1/8-degree recipe
# (extras/Makefile:40-41: match_dy bp so p125sc); res1.0_bp / _p125sc / _cutang / _cutso, res0.5_cutang_regular = the plan
# branches no extras/Makefile recipe reaches (OGG:1050-1057, 1123-1125, 1276-1279, 1296-1313)
@pytest.mark.parametrize("name", ["res4.0", "res1.0", "res0.5", "res0.5_equenh", "readme_dp", "res1.0_dp", "res0.25",
                                  "res0.25_om5proto", "res0.25_full", "res0.125", "res1.0_bp", "res1.0_p125sc",
                                  "res1.0_cutang", "res1.0_cutso", "res0.5_cutang_regular", "res0.667_dp_even"])
def test_config_matches_reference_hashes(SG, golden_configs, name, capsys):
    cfg = golden_configs[name]
    plan = SG.plan_supergrid(**_plan_args(_args(cfg["flags"])))
    planned = capsys.readouterr().out
    for line in cfg.get("stdout_plan", []):        # the reference's own cut messages for this recipe
        if "Cutting" in line or "need to be cut" in line or "No SC grid remained" in line:
            assert line in planned.splitlines(), (line, planned)
    got = SG.build_supergrid(plan).to_host()
    for k in NAMES:
        assert list(got[k].shape) == cfg["arrays"][k]["shape"], (k, got[k].shape)
        assert sha(got[k]) == cfg["arrays"][k]["sha256"], "%s: array %s differs from the reference's" % (name, k)
    total = float(np.sum(got["area"]))
    assert abs(total - cfg["arrays"]["area"]["sum"]) <= 1e-12 * abs(total)       # global-area check, 1e-12 relative


def test_cli_writes_reference_layout(SG, golden_configs, tmp_path):
    from scipy.io import netcdf_file
    from ocean_model_grid_generator_b200 import ocean_grid_generator as G
    cfg = golden_configs["res4.0"]
    out = tmp_path / "ocean_hgrid_res4.0.nc"
    G.cli(["-f", str(out)] + cfg["flags"].split() + ["--write_subgrid_files"])
    f = netcdf_file(str(out), "r", mmap=False)
    assert f.version_byte == 2
    assert list(f.variables.keys()) == ["tile", "y", "x", "dy", "dx", "area", "angle_dx"]
    assert list(f.dimensions.keys()) == ["nyp", "nxp", "ny", "nx", "string"]
    for k in NAMES:
        assert sha(f.variables[k][:]) == cfg["arrays"][k]["sha256"], k
    assert f.variables["x"].units == b"degrees" and f.variables["dx"].units == b"meters" and f.variables["area"].units == b"m2"
    assert b"".join(f.variables["tile"][:5]) == b"tile1"
    for sub in ("Merc", "BP", "SO", "SC"):
        assert (tmp_path / ("ocean_hgrid_res4.0.nc" + sub + ".nc")).exists()


def _check_values(lines):
    """{label: tuple of floats} from the reference-format CHECK_metrics lines (labels repeat: keep them in order)."""
    import re
    out = []
    for ln in lines:
        if "CHECK_metrics" not in ln:
            continue
        label = ln.strip().split(":")[0]
        nums = re.findall(r"[-+]?(?:\d+\.\d*|\.\d+|\d+)(?:[eE][-+]?\d+)?", ln.split(")", 1)[1].replace("np.float64", ""))
        out.append((label, tuple(float(v) for v in nums)))
    return out


@pytest.mark.parametrize("name", ["res4.0", "res1.0", "res1.0_dp", "res1.0_cutso", "readme_dp", "res0.5_equenh"])
def test_cli_check_lines_match_reference(SG, golden_configs, name, tmp_path, capsys):
    """main()'s self-check lines (OGG:1016, 1069, 1111, 1156, 1172) are part of the CLI's stdout: same lines, same order
    (Mercator, bipolar cap, Southern Ocean, southern cap) as the unmodified reference printed for the recipe, the sums within
    1e-12 relative (numpy's pairwise order is not reproduced), i.e. the percentages within 2e-10.  Covers the displaced-pole
    line, which needs the rows inside the doughnut that the stitched grid never holds, the regular cap (its own top dx row)
    and a Southern Ocean band that is cut afterwards."""
    from ocean_model_grid_generator_b200 import ocean_grid_generator as G
    cfg = golden_configs[name]
    G.cli(["-f", str(tmp_path / "g.nc")] + cfg["flags"].split())
    got = _check_values(capsys.readouterr().out.splitlines())
    want = _check_values(cfg["stdout_checks"])
    assert [g[0] for g in got] == [w[0] for w in want]
    for (label, gv), (_, wv) in zip(got, want):
        assert len(gv) == len(wv), (label, gv, wv)
        for g, w in zip(gv, wv):
            assert abs(g - w) <= 2e-10 + 1e-9 * abs(w), (name, label, gv, wv)


def test_cli_check_lines_opt_out(SG, golden_configs, tmp_path, capsys, monkeypatch):
    """OGG_CHECK_LINES=0 leaves the self-check prints (and the second pass over the displaced-pole cap behind them) out;
    the file is the same file."""
    from ocean_model_grid_generator_b200 import ocean_grid_generator as G
    cfg = golden_configs["res1.0_dp"]
    G.cli(["-f", str(tmp_path / "with.nc"), "--no_changing_meta"] + cfg["flags"].split())
    assert len(_check_values(capsys.readouterr().out.splitlines())) > 0
    monkeypatch.setenv("OGG_CHECK_LINES", "0")
    G.cli(["-f", str(tmp_path / "without.nc"), "--no_changing_meta"] + cfg["flags"].split())
    out = capsys.readouterr().out
    assert _check_values(out.splitlines()) == [] and "left out (OGG_CHECK_LINES=0)" in out
    a, b = (tmp_path / "with.nc").read_bytes(), (tmp_path / "without.nc").read_bytes()
    # the history attribute holds the command line (the file name differs); everything from the data section on is equal
    assert len(a) == len(b) and a[-(len(a) // 2):] == b[-(len(b) // 2):]


def test_write_nc_device_path_equals_host_path(SG, tmp_path):
    """write_nc on device tensors (byte order reversed on the GPU, streamed through pinned staging in chunks) writes the
    same file, byte for byte, as write_nc on the copied-out numpy arrays (host conversion) -- including a variable larger
    than one staging chunk and odd element counts."""
    import torch
    from ocean_model_grid_generator_b200 import device as D, ocean_grid_generator as G
    plan = SG.plan_supergrid(**_plan_args(_args("-r 1 --r_dp 0.2")))
    a = SG.build_supergrid(plan).arrays()
    h = {k: D.to_host(v) for k, v in a.items()}
    old = G._DeviceVariable.CHUNK
    try:
        G._DeviceVariable.CHUNK = 100003          # several chunks per variable, not a multiple of anything
        G._STAGING.clear()
        G.write_nc(a["x"], a["y"], a["dx"], a["dy"], a["area"], a["angle_dx"], fnam=str(tmp_path / "dev.nc"),
                   description="d", history="h", source="s")
    finally:
        G._DeviceVariable.CHUNK = old
        G._STAGING.clear()
    G.write_nc(h["x"], h["y"], h["dx"], h["dy"], h["area"], h["angle_dx"], fnam=str(tmp_path / "host.nc"),
               description="d", history="h", source="s")
    assert (tmp_path / "dev.nc").read_bytes() == (tmp_path / "host.nc").read_bytes()
    # ogg_bswap64 itself, device -> device
    t = torch.arange(1000, dtype=torch.float64, device="cuda") * 1.25 - 7
    out = torch.empty(1000, dtype=torch.int64, device="cuda")
    from ocean_model_grid_generator_b200 import _cabi
    _cabi.call("ogg_bswap64", D.ptr(t), D.ptr(out), 1000, D.stream())
    assert out.cpu().numpy().tobytes() == t.cpu().numpy().astype(">f8").tobytes()


def test_subgrid_files_follow_reference_cut_order(SG, golden_configs, tmp_path):
    """--write_subgrid_files (OGG:1020, 1074, 1115, 1228): Mercator / bipolar / Southern Ocean files hold the sub-grids as
    generated, the southern-cap file the cap after the doughnut cut, all BEFORE the south cut of the stitched grid."""
    from scipy.io import netcdf_file
    from ocean_model_grid_generator_b200 import ocean_grid_generator as G
    out = tmp_path / "g.nc"
    G.cli(["-f", str(out), "-r", "1.0", "--r_dp", "0.2", "--south_cutoff_row", "45", "--write_subgrid_files", "--no_changing_meta"])
    shapes = {}
    for sub in ("Merc", "BP", "SO", "SC"):
        f = netcdf_file(str(out) + sub + ".nc", "r", mmap=False)
        shapes[sub] = f.variables["x"].shape
        f.close()
    f = netcdf_file(str(out), "r", mmap=False)
    whole = f.variables["x"].shape
    f.close()
    assert shapes["SO"] == (56, 721) and shapes["SC"] == (35, 721) and shapes["BP"] == (121, 721)   # uncut SO; SC: 71 rows - 36 cut
    assert whole == tuple(golden_configs["res1.0_cutso"]["arrays"]["x"]["shape"])                  # SC gone, 9 SO rows cut


def test_reference_file_hash_res4_0_stretch(SG, tmp_path):
    """The reference's own known-answer test (t/test_ocean_grid_gen.py:25-46): sha256 of the whole file written by
    `-r 0.25 --ensure_nj_even --no_changing_meta`.  Those hashes were made with upstream numpypi's polynomials and
    netCDF-C's writer, neither of which exists in this image (parity unpinned against upstream
    numpypi), so a mismatch is the expected outcome and is reported as xfail; a match would pin the trig spec."""
    import hashlib
    from ocean_model_grid_generator_b200 import ocean_grid_generator as G
    out = tmp_path / "ocean_hgrid_res4.0.nc"
    G.cli(["-f", str(out), "-r", "0.25", "--ensure_nj_even", "--no_changing_meta"])
    digest = hashlib.sha256(out.read_bytes()).hexdigest()
    accepted = {"0bc8e50b7d4110fcfb148f9f9590cf3c5e6c2d699b9c5519600aea7cc9483381",      # PAN
                "7e04457b4411260341f304a5182ec7f289ddb2724021e2825751f9994e23a6ed",      # github CI
                "8cfd0f521bdbbc2d67729d94cc98d0b079e218cc045a7ef31335262704eabcf9"}      # github CI
    if digest not in accepted:
        pytest.xfail("file sha256 %s is none of the reference's three (upstream numpypi absent: trig spec is this repo's own)" % digest)


def test_cli_argument_errors(SG):
    from ocean_model_grid_generator_b200 import ocean_grid_generator as G
    with pytest.raises(SystemExit) as e:
        G.main(4, r_dp=0.2, lat_dp=-80.0, match_dy=[], grids="all")            # OGG:904-907
    assert e.value.code == 2
    with pytest.raises(SystemExit) as e:
        G.main(4)                                                               # python defaults contain " sc" (OGG:870-888)
    assert e.value.code == 2


@pytest.mark.parametrize("flags,world", [("-r 1.0 --r_dp 0.2", 3), ("-r 0.25 --ensure_nj_even", 8), ("-r 2", 2)])
def test_row_slabs_reassemble_to_whole(SG, flags, world):
    """Each GPU's share (rank r of `world`) concatenated in rank order == the single-GPU supergrid, bit for bit."""
    plan = SG.plan_supergrid(**_plan_args(_args(flags)))
    whole = SG.build_supergrid(plan).to_host()
    # ranks hold every piece's slab; reassemble piece by piece
    parts = [SG.build_supergrid(plan, r, world) for r in range(world)]
    hosts = [p.to_host() for p in parts]
    area_total = sum(p.area_sum() for p in parts)
    assert abs(area_total - whole["area"].sum()) <= 1e-12 * whole["area"].sum()
    for k in NAMES:
        rows = []
        for pi in range(len(plan.pieces)):
            for r in range(world):
                c0, c1 = parts[r].slabs[pi]
                off = sum(b - a for a, b in parts[r].slabs[:pi])
                rows.append(hosts[r][k][off:off + (c1 - c0)])
        if k in ("x", "y", "dx", "angle_dx"):
            rows.append(hosts[world - 1][k][-1:])
        assert_bits(np.concatenate(rows, axis=0), whole[k], "%s reassembled from %d slabs" % (k, world))


def test_properties_quarter_degree(SG):
    """BASELINE configs[1] (1/4 degree full tripolar, displaced pole): properties that need no oracle run."""
    plan = SG.plan_supergrid(**_plan_args(_args("-r 4 --r_dp 0.2")))
    sg = SG.build_supergrid(plan)
    a = sg.to_host()
    assert a["x"].shape == (2243, 2881) and a["dx"].shape == (2243, 2880) and a["dy"].shape == (2242, 2881)
    assert np.isfinite(a["area"]).all() and (a["area"] > 0).all() and (a["dx"] > 0).all() and (a["dy"] >= 0).all()
    assert a["angle_dx"].min() >= -90.0 and a["angle_dx"].max() <= 90.0
    ycol = a["y"][:, plan.Ni // 4]
    assert (np.diff(ycol) > 0).all() and np.searchsorted(ycol, 0.0) % 2 == 1       # OGG:1366-1370, :1425-1434
    # bipolar fold symmetry: the last row is its own mirror image (y symmetric about the fold's centre)
    top = a["y"][-1]
    assert np.abs(top - top[::-1]).max() < 1e-9
    # sphere area north of the doughnut edge: sum(area) = 2 pi Re^2 (1 - sin(lat_south)) up to the cap's quadrature error
    Re = 6371.0e3
    lat_s = a["y"][0, 0]
    exact = 2 * np.pi * Re * Re * (1 - np.sin(np.deg2rad(-78.0)))               # bands + bipolar north of 78S (exact pieces)
    j78 = plan.pieces[0].rows - 1
    north = a["area"][j78:].sum()
    assert abs(north - exact) / exact < 1e-12 * 50, (north, exact)
    assert abs(sg.area_sum() - a["area"].sum()) <= 1e-12 * a["area"].sum()


def test_eighth_degree_slabs_checksum(SG):
    """BASELINE configs[2] (1/8 degree): 4 row-slabs == whole (sha256 per array), without any oracle run."""
    plan = SG.plan_supergrid(**_plan_args(_args("-r 8 --r_dp 0.2")))
    assert (plan.nyp, plan.nxp) == (4485, 5761)
    whole = SG.build_supergrid(plan).to_host()
    world = 4
    parts = [SG.build_supergrid(plan, r, world) for r in range(world)]
    hosts = [p.to_host() for p in parts]
    for k in NAMES:
        rows = []
        for pi in range(len(plan.pieces)):
            for r in range(world):
                c0, c1 = parts[r].slabs[pi]
                off = sum(b - a for a, b in parts[r].slabs[:pi])
                rows.append(hosts[r][k][off:off + (c1 - c0)])
        if k in ("x", "y", "dx", "angle_dx"):
            rows.append(hosts[world - 1][k][-1:])
        assert sha(np.concatenate(rows, axis=0)) == sha(whole[k]), k


# ---- device checksums and the slab sequence (grids larger than one GPU's HBM) -------------------------------------
def _mix64_host(a):
    z = np.ascontiguousarray(a, dtype=np.float64).view(np.uint64).ravel().copy()
    with np.errstate(over="ignore"):
        z += np.uint64(0x9e3779b97f4a7c15)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xbf58476d1ce4e5b9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94d049bb133111eb)
        z = z ^ (z >> np.uint64(31))
        return int(z.sum(dtype=np.uint64))


def _u64(t):
    return [int(v) & 0xffffffffffffffff for v in t.cpu().tolist()]


def test_checksum_kernel_matches_host(SG):
    """ogg_checksum_accumulate == the same splitmix64 sum computed with numpy integers, and it accumulates."""
    import ctypes
    import torch
    from ocean_model_grid_generator_b200 import _cabi, device as D
    rng = np.random.default_rng(7)
    a = rng.standard_normal(100003)
    a[:5] = [0.0, -0.0, np.inf, np.nan, 5e-324]
    t = D.to_dev(a)
    acc = torch.zeros(1, dtype=torch.int64, device="cuda")
    _cabi.call("ogg_checksum_accumulate", D.ptr(t), t.numel(), ctypes.c_void_p(acc.data_ptr()), D.stream())
    assert _u64(acc)[0] == _mix64_host(a)
    _cabi.call("ogg_checksum_accumulate", D.ptr(t), 17, ctypes.c_void_p(acc.data_ptr()), D.stream())
    assert _u64(acc)[0] == (_mix64_host(a) + _mix64_host(a[:17])) % 2 ** 64
    _cabi.call("ogg_checksum_accumulate", None, 0, ctypes.c_void_p(acc.data_ptr()), D.stream())   # empty input: no-op


@pytest.mark.parametrize("flags,V", [("-r 2 --r_dp 0.2", 3), ("-r 8 --r_dp 0.2", 8)])
def test_slab_sequence_in_shared_buffers(SG, flags, V):
    """V row-slabs built one after the other in ONE slab's buffers (how one GPU streams a grid larger than its HBM):
    the per-array checksums add up to the whole grid's, which equal the host-side checksum of the copied-out arrays."""
    import torch
    plan = SG.plan_supergrid(**_plan_args(_args(flags)))
    whole = SG.build_supergrid(plan)
    ref = _u64(whole.checksums())
    host = whole.to_host()
    assert ref == [_mix64_host(host[k]) for k in SG.FIELDS]
    sg = SG.Supergrid(plan, 0, V, capacity_world=V)
    acc = torch.zeros(6, dtype=torch.int64, device="cuda")
    area = 0.0
    for v in range(V):
        sg.set_slab(v, V).build()
        sg.checksums(acc)
        area += sg.area_sum()
    assert _u64(acc) == ref
    tot = host["area"].sum()
    assert abs(area - tot) <= 1e-12 * tot


def test_thirtieth_degree_slabs_and_area(SG):
    """BASELINE configs[3] (1/30 degree, OM5-class, 16821 x 21601 vertices = 17.4 GB): whole grid on one GPU == 8 row-slabs
    (checksum of checksums, all six arrays), sphere area north of the displaced-pole cap exact to 1e-12 relative,
    latitude strictly increasing along the symmetry meridian, equator on an odd row (OGG:1366-1370, 1425-1436)."""
    import torch
    from ocean_model_grid_generator_b200 import device as D
    plan = SG.plan_supergrid(**_plan_args(_args("-r 30 --r_dp 0.2 --ensure_nj_even")))
    assert (plan.nyp, plan.nxp) == (16821, 21601)
    whole = SG.build_supergrid(plan)
    ref = _u64(whole.checksums())
    total_area = whole.area_sum()
    a = whole.arrays()
    ycol = D.to_host(a["y"][:, plan.Ni // 4])
    assert (np.diff(ycol) > 0).all() and np.searchsorted(ycol, 0.0) % 2 == 1
    assert bool(torch.isfinite(a["area"]).all()) and float(a["area"].min()) > 0 and float(a["dx"].min()) > 0
    assert float(a["angle_dx"].min()) >= -90.0 and float(a["angle_dx"].max()) <= 90.0
    Re = 6371.0e3
    j78 = plan.pieces[0].rows - 1                    # first cell row of the Southern Ocean band (78S)
    from ocean_model_grid_generator_b200 import ocean_grid_generator as G
    north = G.device_sum(a["area"][j78:])
    exact = 2 * np.pi * Re * Re * (1 - np.sin(np.deg2rad(-78.0)))
    assert abs(north - exact) / exact < 5e-11, (north, exact)
    del a, whole
    torch.cuda.empty_cache()
    V = 8
    sg = SG.Supergrid(plan, 0, V, capacity_world=V)
    acc = torch.zeros(6, dtype=torch.int64, device="cuda")
    area = 0.0
    for v in range(V):
        sg.set_slab(v, V).build()
        sg.checksums(acc)
        area += sg.area_sum()
    assert _u64(acc) == ref
    assert abs(area - total_area) <= 1e-12 * total_area


@pytest.mark.parametrize("r", [6, 8])
def test_eighth_degree_equals_cpu_oracle(SG, oracle, r):
    """BASELINE configs[2] (1/8 degree, 4485 x 5761) and the 2-GPU weak-scaling grid (1/6 degree): every one of the six
    arrays bit-identical to the CPU oracle's composition of the same grid (oracle/supergrid_cpu.py; ~10 s of host time)."""
    import supergrid_cpu
    plan = SG.plan_supergrid(**_plan_args(_args("-r %d --r_dp 0.2" % r)))
    got = SG.build_supergrid(plan).to_host()
    ref = supergrid_cpu.build_cpu(r, r_dp=0.2)
    for k in NAMES:
        assert_bits(got[k], ref[k], "r=%d %s" % (r, k))
    tot = ref["area"].sum()
    assert abs(got["area"].sum() - tot) <= 1e-12 * tot


# Row-slabs of the large grids against the CPU oracle.  slab = (k, V): rank k's share of every sub-grid's cell rows when the
# grid is split V ways -- the same split (slabs.slab_bounds) on both sides.  The slab sizes keep the oracle's host time to
# seconds: 1/V of a grid of P points at ~5e6 points/s.  k = 0 holds every sub-grid's southern edge (the doughnut edge of the
# displaced-pole cap, the first Mercator row); the last k holds every closing row (the bipolar pole row, dx of the top
# row); a middle k has neither.
BIG_SLABS = [
    (12, "", 8, 0), (12, "", 8, 3), (12, "", 8, 7),                       # N=8 weak-scaling grid (6725 x 8641), whole in 8 slabs
    (30, "--ensure_nj_even", 32, 0), (30, "--ensure_nj_even", 32, 13), (30, "--ensure_nj_even", 32, 31),   # BASELINE configs[3]
    (120, "", 512, 0), (120, "", 512, 255), (120, "", 512, 511),          # BASELINE configs[4]: Ni = 86400, 5.8e9 points
]


@pytest.mark.parametrize("r,extra,V,k", BIG_SLABS)
def test_large_grid_slabs_equal_cpu_oracle(SG, oracle, r, extra, V, k):
    """1/12, 1/30 and 1/120 degree (BASELINE configs[3] and [4], and the 8-GPU weak grid): a row-slab of the stitched
    grid, all six arrays, bit for bit against the CPU oracle's evaluation of the same slab.  These are the sizes at which
    the 32-bit item / column / sample numbers of the cap kernels, (i*360)/Ni at Ni = 21600 / 86400 and the chunked
    quadrature are exercised (OGG:136-188, 565-601, 1316-1365)."""
    import supergrid_cpu
    import torch
    plan = SG.plan_supergrid(**_plan_args(_args(("-r %d --r_dp 0.2 " % r) + extra)))
    assert plan.Ni == 720 * r
    sg = SG.Supergrid(plan, k, V).build()
    got = sg.to_host()
    del sg
    torch.cuda.empty_cache()
    ref = supergrid_cpu.build_cpu(r, r_dp=0.2, slab=(k, V), ensure_nj_even=bool(extra))
    # the oracle's slab carries its top piece's closing vertex row, which belongs to the next slab (Supergrid.arrays()
    # holds it on the last slab only)
    n = ref["area"].shape[0] + (1 if k == V - 1 else 0)
    assert n > 1
    for name in NAMES:
        want = ref[name][:n] if name in ("x", "y", "dx", "angle_dx") else ref[name]
        assert_bits(got[name], want, "r=%d slab %d/%d %s" % (r, k, V, name))


@pytest.mark.parametrize("rank,world", [(0, 1), (1, 3), (2, 3)])
def test_build_into_pinned_host_buffers(SG, rank, world):
    """The end-to-end call of bench.py: build(host=pinned) copies each sub-grid out on its own stream while the caps
    compute.  Host buffers == a plain build read back afterwards; concurrent == single-stream build, bit for bit."""
    import torch
    plan = SG.plan_supergrid(**_plan_args(_args("-r 1 --r_dp 0.2")))
    ref = SG.Supergrid(plan, rank, world).build(concurrent=False)
    torch.cuda.synchronize()
    want = ref.to_host()
    sg = SG.Supergrid(plan, rank, world)
    pinned = sg.pinned_like()
    for v in pinned.values():
        v.fill_(-7.0)
    for _ in range(2):                                   # the second call reuses streams, events and buffers
        sg.build(host=pinned)
        torch.cuda.synchronize()
        got_dev = sg.to_host()
        for k in NAMES:
            assert_bits(pinned[k].numpy(), want[k], "pinned %s rank %d/%d" % (k, rank, world))
            assert_bits(got_dev[k], want[k], "concurrent %s rank %d/%d" % (k, rank, world))


def test_skip_metrics_leaves_minus_one(SG):
    """--skip_metrics (OGG:1009-1011, 1069-1071, ...): coordinates and angles as usual, dx / dy / area stay -1."""
    a = _args("-r 1 --r_dp 0.2")
    full = SG.build_supergrid(SG.plan_supergrid(**_plan_args(a))).to_host()
    a["skip_metrics"] = True
    got = SG.build_supergrid(SG.plan_supergrid(**_plan_args(a))).to_host()
    for k in ("x", "y", "angle_dx"):
        assert_bits(got[k], full[k], "skip_metrics " + k)
    for k in ("dx", "dy", "area"):
        assert got[k].shape == full[k].shape and (got[k] == -1.0).all(), k


def test_timing_brackets_and_pair_span(SG):
    """ogg_timing_*: per-class totals and the busy span of the two co-running cap quadratures (bench.py's roofline)."""
    import ctypes
    import torch
    from ocean_model_grid_generator_b200 import _cabi
    lib = _cabi.load()
    plan = SG.plan_supergrid(**_plan_args(_args("-r 2 --r_dp 0.2")))
    sg = SG.Supergrid(plan)
    sg.build(); torch.cuda.synchronize()
    kid = {lib.ogg_timing_name(k).decode(): k for k in range(lib.ogg_timing_kernels())}
    lib.ogg_timing_reset(); lib.ogg_timing_enable(2)
    n0 = lib.ogg_launch_count()
    for _ in range(3):
        sg.build()
    torch.cuda.synchronize()
    lib.ogg_timing_enable(0)
    assert lib.ogg_launch_count() - n0 >= 3 * 10

    def total(name):
        t, n = ctypes.c_double(), ctypes.c_longlong()
        assert lib.ogg_timing_get(kid[name], ctypes.byref(t), ctypes.byref(n)) == 0
        return t.value, n.value
    (tdp, ndp), (tbp, nbp) = total("dp_hsamples"), total("bipolar_quad_fused")
    span, pairs = ctypes.c_double(), ctypes.c_longlong()
    assert lib.ogg_timing_pair_span(kid["dp_hsamples"], kid["bipolar_quad_fused"], ctypes.byref(span), ctypes.byref(pairs)) == 0
    assert ndp == nbp == 3 and pairs.value == 6 and tdp > 0 and tbp > 0
    assert max(tdp, tbp) <= span.value * 1.001 and span.value <= (tdp + tbp) * 1.5 + 1.0
    lib.ogg_timing_reset()
    assert total("dp_hsamples") == (0.0, 0)

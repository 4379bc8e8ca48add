"""ctypes face of the C oracle (oracle/ogg_oracle.c) -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Function names and argument meaning follow /root/reference/ocean_grid_generator.py (cited as
OGG:n) so parity tests read like calls into the reference.  The 1-D Mercator latitude law
(OGG:292-441) is restated here in numpy on the portable-trig twin (oracle/numpypi).
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.
"""
import ctypes
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
if HERE not in sys.path:
    sys.path.insert(0, HERE)
import numpypi.numpypi_series as pnp  # noqa: E402  (portable-trig twin)

PI_180 = np.pi / 180.0
_default_Re = 6371.0e3
_LIB = None
_dp = ctypes.POINTER(ctypes.c_double)
_i64 = ctypes.c_int64
_d = ctypes.c_double
_int = ctypes.c_int


def build():
    subprocess.check_call(["make", "-s", "-C", HERE])


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(HERE, "libogg_oracle.so")
        if not os.path.exists(path):
            build()
        _LIB = ctypes.CDLL(path)
        _LIB.ogo_generate_bipolar_cap_mesh.restype = ctypes.c_double
    return _LIB


def _p(a):
    return a.ctypes.data_as(_dp) if a is not None else None


def _c(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def num_threads():
    return int(lib().ogo_num_threads())


def unary(name, x):
    x = _c(x)
    out = np.empty_like(x)
    getattr(lib(), "ogo_" + name)(_p(x), _p(out), _i64(x.size))
    return out


def arctan2(y, x):
    y, x = np.broadcast_arrays(_c(y), _c(x))
    y, x = _c(y), _c(x)
    out = np.empty_like(x)
    lib().ogo_arctan2(_p(y), _p(x), _p(out), _i64(x.size))
    return out


def mod360(x):
    return unary("mod360", x)


# ---- bipolar cap --------------------------------------------------------------------------
def bipolar_projection(lamg, phig, lon_bp, rp, metrics_only=False):
    lamg, phig = _c(lamg), _c(phig)
    hi, hj = np.empty_like(lamg), np.empty_like(lamg)
    lams = phis = None
    if not metrics_only:
        lams, phis = np.empty_like(lamg), np.empty_like(lamg)
    lib().ogo_bipolar_projection(_p(lamg), _p(phig), _i64(lamg.size), _d(lon_bp), _d(rp), _int(int(metrics_only)),
                                 _p(lams), _p(phis), _p(hi), _p(hj))
    if metrics_only:
        return hi, hj
    return lams, phis, hi, hj


def generate_bipolar_cap_mesh(Ni, Nj_ncap, lat0_bp, lon_bp, ensure_nj_even=True):
    if Nj_ncap % 2 != 0 and ensure_nj_even:
        Nj_ncap = Nj_ncap - 1
    lams = np.empty((Nj_ncap + 1, Ni + 1))
    phis = np.empty((Nj_ncap + 1, Ni + 1))
    hi = np.empty((Nj_ncap + 1, Ni))
    hj = np.empty((Nj_ncap, Ni + 1))
    lib().ogo_generate_bipolar_cap_mesh(_i64(Ni), _i64(Nj_ncap), _d(lat0_bp), _d(lon_bp), _p(lams), _p(phis), _p(hi), _p(hj))
    return lams, phis, hi, hj


def bipolar_rp(lat0_bp):
    """rp of OGG:117 / :1062."""
    return float(unary("tan", np.array([0.5 * (90 - lat0_bp) * PI_180]))[0])


def bipolar_cap_ij_array(i, j, Ni, Nj_ncap, lat0_bp, lon_bp, rp):
    i, j = _c(i), _c(j)
    hi = np.empty((j.size, i.size))
    hj = np.empty((j.size, i.size))
    lib().ogo_bipolar_cap_ij_array(_p(i), _i64(i.size), _p(j), _i64(j.size), _i64(Ni), _i64(Nj_ncap), _d(lat0_bp),
                                   _d(lon_bp), _d(rp), _p(hi), _p(hj))
    return hi, hj


def bipolar_cap_metrics_quad_fast(order, nx, ny, lat0_bp, lon_bp, rp, Re=_default_Re):
    dxq, dyq, daq = np.empty((ny + 1, nx)), np.empty((ny, nx + 1)), np.empty((ny, nx))
    rc = lib().ogo_bipolar_cap_metrics_quad_fast(_int(order), _i64(nx), _i64(ny), _d(lat0_bp), _d(lon_bp), _d(rp),
                                                 _d(Re), _p(dxq), _p(dyq), _p(daq))
    if rc:
        raise Exception("Uncoded order")
    return dxq, dyq, daq


def sample_index(order, c0, c1, total, fudge=False):
    """j1d / i1d of OGG:143-156, 573-581 for cells [c0, c1] (c1 = the ghost cell that closes the slab): order nodes per
    cell; with `fudge`, a node equal to `total` becomes total - 0.001 (OGG:146-147)."""
    a, b = quad_positions(order)
    c = np.arange(c0, c1 + 1, dtype=np.float64)
    v = (b[None, :] * c[:, None] + a[None, :] * (c[:, None] + 1)).reshape(-1)
    if fudge:
        last = v.reshape(-1, order)[:, order - 1]
        last[last == float(total)] = float(total) - 0.001
    return np.ascontiguousarray(v)


def quad_cells(order, nx, ny, dx, dy, Re=_default_Re):
    """Tail of OGG:174-187 / 585-599 on sample arrays [order*(ny+1)][order*(nx+1)] -> dxq [ny+1][nx], dyq [ny][nx+1], daq [ny][nx]."""
    dx, dy = _c(dx), _c(dy)
    dxq, dyq, daq = np.empty((ny + 1, nx)), np.empty((ny, nx + 1)), np.empty((ny, nx))
    lib().ogo_quad_cells(_int(order), _i64(nx), _i64(ny), _p(dx), _p(dy), _d(Re), _p(dxq), _p(dyq), _p(daq))
    return dxq, dyq, daq


def quad_positions(n=3):
    a, b = np.empty(5), np.empty(5)
    if lib().ogo_quad_positions(_int(n), _p(a), _p(b)):
        raise Exception("Uncoded order")
    return a[:n].copy(), b[:n].copy()


def quad_average(y):
    y = _c(y)
    if y.size not in (2, 3, 4, 5):
        raise Exception("Uncoded order")
    out = np.empty(1)
    lib().ogo_quad_average(_p(y), _int(y.size), _p(out))
    return out[0]


def quad_average_2d(y):
    y = _c(y)
    if y.shape[0] != y.shape[1]:
        raise Exception("Input array is not squared!")
    if y.shape[0] not in (2, 3, 4, 5):
        raise Exception("Uncoded order")
    out = np.empty(1)
    lib().ogo_quad_average_2d(_p(y), _int(y.shape[0]), _p(out))
    return out[0]


# ---- displaced-pole cap -------------------------------------------------------------------
def dp_constants(lat0, lam_pole, r_pole):
    rj, zr, zi = _d(), _d(), _d()
    lib().ogo_dp_constants(_d(lat0), _d(lam_pole), _d(r_pole), ctypes.byref(rj), ctypes.byref(zr), ctypes.byref(zi))
    return rj.value, complex(zr.value, zi.value)


def displacedPoleCap_projection(lon_grid, lat_grid, z_0, r_joint):
    lon_grid, lat_grid = _c(lon_grid), _c(lat_grid)
    nj, ni = lon_grid.shape
    lam, phi = np.empty_like(lon_grid), np.empty_like(lon_grid)
    lib().ogo_dp_projection(_p(lon_grid), _p(lat_grid), _i64(nj), _i64(ni), _d(z_0.real), _d(z_0.imag), _d(r_joint),
                            _p(lam), _p(phi))
    return lam, phi


def monotonic_bounding(x, x_0):
    assert x.dtype == np.float64 and x.flags.c_contiguous
    lib().ogo_monotonic_bounding(_p(x), _i64(x.shape[0]), _i64(x.shape[1]), _d(x_0))
    return x


def displacedPoleCap_mesh(i, j, ni, nj, lon0, lat0, lam_pole, r_pole, excluded_fraction=None):
    i, j = _c(i), _c(j)
    lams, phis = np.empty((j.size, i.size)), np.empty((j.size, i.size))
    lib().ogo_dp_mesh(_p(i), _i64(i.size), _p(j), _i64(j.size), _i64(ni), _i64(nj), _d(lon0), _d(lat0), _d(lam_pole),
                      _d(r_pole), _p(lams), _p(phis))
    londp, latdp = lams[0, 0], phis[0, 0]
    if excluded_fraction is not None:
        ny = lams.shape[0]
        jmin = np.ceil(excluded_fraction * ny)
        jmin = jmin + np.mod(jmin, 2)
        jmint = int(jmin)
        return lams[jmint:, :], phis[jmint:, :], londp, latdp
    return lams, phis, londp, latdp


def generate_displaced_pole_grid(Ni, Nj_scap, lon0, lat0, lon_dp, r_dp):
    return displacedPoleCap_mesh(np.arange(Ni + 1), np.arange(Nj_scap + 1), Ni, Nj_scap, lon0, lat0, lon_dp, r_dp)


def great_arc_distance(j0, i0, j1, i1, nx, ny, lon0, lat0, lon_dp, r_dp):
    j0, i0, j1, i1 = _c(j0), _c(i0), _c(j1), _c(i1)
    out = np.empty((j0.size, i0.size))
    lib().ogo_great_arc_distance(_p(j0), _p(i0), _p(j1), _p(i1), _i64(j0.size), _i64(i0.size), _i64(nx), _i64(ny),
                                 _d(lon0), _d(lat0), _d(lon_dp), _d(r_dp), _p(out))
    return out


def _numerical_h(direction, j, i, nx, ny, lon0, lat0, lon_dp, r_dp, eps, order):
    j, i = _c(j), _c(i)
    out = np.empty((j.size, i.size))
    rc = lib().ogo_numerical_h(_int(direction), _p(j), _i64(j.size), _p(i), _i64(i.size), _i64(nx), _i64(ny), _d(lon0),
                               _d(lat0), _d(lon_dp), _d(r_dp), _d(eps), _int(order), _p(out))
    if rc:
        raise Exception("order not coded")
    return out


def numerical_hi(j, i, nx, ny, lon0, lat0, lon_dp, r_dp, eps, order=6):
    return _numerical_h(0, j, i, nx, ny, lon0, lat0, lon_dp, r_dp, eps, order)


def numerical_hj(j, i, nx, ny, lon0, lat0, lon_dp, r_dp, eps, order=6):
    return _numerical_h(1, j, i, nx, ny, lon0, lat0, lon_dp, r_dp, eps, order)


def displacedPoleCap_metrics_quad(order, nx, ny, lon0, lat0, lon_dp, r_dp, Re=_default_Re):
    dxq, dyq, daq = np.empty((ny + 1, nx)), np.empty((ny, nx + 1)), np.empty((ny, nx))
    rc = lib().ogo_dp_metrics_quad(_int(order), _i64(nx), _i64(ny), _d(lon0), _d(lat0), _d(lon_dp), _d(r_dp), _d(Re),
                                   _p(dxq), _p(dyq), _p(daq))
    if rc == -1:
        raise Exception("Uncoded order")
    if rc == -2:
        raise Exception("order not coded")
    return dxq, dyq, daq


# ---- bands ----------------------------------------------------------------------------------
def generate_grid_metrics_MIDAS(x, y, axis_units="degrees", Re=_default_Re, latlon_areafix=True):
    x, y = _c(x), _c(y)
    nyp, nxp = x.shape
    dx, dy, area = np.empty((nyp, nxp - 1)), np.empty((nyp - 1, nxp)), np.empty((nyp - 1, nxp - 1))
    lib().ogo_metrics_midas(_p(x), _p(y), _i64(nyp), _i64(nxp), _d(Re), _int(int(latlon_areafix)), _p(dx), _p(dy), _p(area))
    return dx, dy, area


def angle_x(x, y):
    if x.shape != y.shape:
        raise Exception("Input arrays do not have the same shape!")
    x, y = _c(x), _c(y)
    out = np.empty_like(x)
    lib().ogo_angle_x(_p(x), _p(y), _i64(x.shape[0]), _i64(x.shape[1]), _p(out))
    return out


def generate_latlon_grid(lni, lnj, llon0, llen_lon, llat0, llen_lat, ensure_nj_even=True):
    lam, phi = np.empty((lnj + 1, lni + 1)), np.empty((lnj + 1, lni + 1))
    lib().ogo_latlon_grid(_i64(lni), _i64(lnj), _d(llon0), _d(llen_lon), _d(llat0), _d(llen_lat), _p(lam), _p(phi))
    if lam.shape[0] % 2 == 0 and ensure_nj_even:                                        # OGG:836-838
        lam, phi = lam[1:], phi[1:]
    return np.ascontiguousarray(lam), np.ascontiguousarray(phi)


# ---- 1-D Mercator latitude law (numpy on the portable twin) ---------------------------------
def y_mercator(Ni, phi):                                                               # OGG:292-295
    R = Ni / (2 * np.pi)
    return R * (pnp.log((1.0 + pnp.sin(phi)) / pnp.cos(phi)))


def phi_mercator(Ni, y):                                                               # OGG:298-301
    R = Ni / (2 * np.pi)
    return pnp.arctan(pnp.sinh(y / R)) * (180 / np.pi)


def y_mercator_rounded(Ni, phi):                                                       # OGG:309-311
    y_float = y_mercator(Ni, phi)
    return (np.sign(y_float) * np.round(np.abs(y_float))).astype(int)


def mercator_phi_1d(Ni, phi_s, phi_n, shift_equator_to_u_point=True, ensure_nj_even=True):
    """Plain (enhanced_equatorial=0) branch of OGG:314-347: the 1-D latitude vector."""
    y_star = y_mercator_rounded(Ni, np.array([phi_s * PI_180, phi_n * PI_180]))
    if y_star[0] % 2 == 0 and shift_equator_to_u_point:
        y_star[0] -= 1
        y_star[1] -= 1
    if (y_star[1] - y_star[0] + 1) % 2 == 0 and ensure_nj_even:
        y_star[1] -= 1
    return phi_mercator(Ni, np.arange(y_star[0], y_star[1] + 1))

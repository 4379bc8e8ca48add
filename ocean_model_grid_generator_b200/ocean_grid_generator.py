#!/usr/bin/env python
"""Drop-in mirror of the reference module `ocean_grid_generator` for the supergrid hot path.

Same function names, argument meaning, return shapes, exceptions and CLI flags as
/root/reference/ocean_grid_generator.py (cited as OGG:n); the arithmetic runs in the sm_100a kernels of
libogg_b200.so through the C ABI (include/ogg_b200.h).  Array-in functions accept numpy arrays (results come
back as numpy arrays, like the reference) or CUDA float64 torch tensors (results stay on the device).
Generator functions take an extra keyword `device=False`; with device=True they return CUDA tensors.
Host code here is 1-D / integer glue only; there is no CPU implementation of the 2-D math.
Out of scope: the matplotlib plotting helpers (OGG:625-679).
"""
from __future__ import print_function

import argparse
import sys

import numpy

from . import _cabi
from . import device as _dev
from . import devnp as np  # the reference's `np` is numpypi.numpypi_series; ours evaluates on the GPU
from . import netcdf3 as _netcdf3

# Constants (OGG:13-16)
PI_180 = np.pi / 180.0
_default_Re = 6371.0e3
HUGE = 1.0e30


def chksum(x, lbl):
    """sha256 + statistics line (OGG:19-30); host utility."""
    import hashlib

    if type(x) in (float, int, numpy.float64):
        y = numpy.array(x)
    else:
        y = numpy.zeros(x.shape)
        y[:] = x
    ymin, ymax, ymean = y.min(), y.max(), y.mean()
    ysd = numpy.sqrt(((y - ymean) ** 2).mean())
    print(hashlib.sha256(y).hexdigest(), "%10s" % lbl,
          "min = %.15f" % ymin, "max = %.15f" % ymax, "mean = %.15f" % ymean, "sd = %.15f" % ysd)


# --------------------------------------------------------------------------------------------------
# small plumbing
# --------------------------------------------------------------------------------------------------
def _out(t, on_dev):
    return t if on_dev else _dev.to_host(t)


def _any_dev(*arrays):
    return any(_dev.is_dev(a) for a in arrays)


# --------------------------------------------------------------------------------------------------
# bipolar cap (OGG:33-188)
# --------------------------------------------------------------------------------------------------
def bipolar_projection(lamg, phig, lon_bp, rp, metrics_only=False):
    """OGG:33-100."""
    on_dev = _any_dev(lamg, phig)
    tl, tp = _dev.to_dev(lamg), _dev.to_dev(phig)
    if tl.shape != tp.shape:
        tl, tp = [t.contiguous() for t in __import__("torch").broadcast_tensors(tl, tp)]
    hi, hj = _dev.empty(*tl.shape), _dev.empty(*tl.shape)
    lams = phis = None
    if not metrics_only:
        lams, phis = _dev.empty(*tl.shape), _dev.empty(*tl.shape)
    _cabi.call("ogg_bipolar_projection", _dev.ptr(tl), _dev.ptr(tp), tl.numel(), float(lon_bp), float(rp),
               int(bool(metrics_only)), _dev.ptr(lams), _dev.ptr(phis), _dev.ptr(hi), _dev.ptr(hj), _dev.stream())
    if metrics_only:
        return _out(hi, on_dev), _out(hj, on_dev)
    return _out(lams, on_dev), _out(phis, on_dev), _out(hi, on_dev), _out(hj, on_dev)


def bipolar_rp(lat0_bp):
    """rp = tan(0.5*(90-lat0_bp)*PI_180) (OGG:117, :1062), on the device."""
    return float(np.tan(0.5 * (90 - lat0_bp) * PI_180))


def generate_bipolar_cap_mesh(Ni, Nj_ncap, lat0_bp, lon_bp, ensure_nj_even=True, device=False):
    """OGG:103-122."""
    print("Generating bipolar grid bounded at latitude ", lat0_bp)
    if Nj_ncap % 2 != 0 and ensure_nj_even:
        print("   Supergrid has an odd number of area cells!")
        print("   The number of j's is not even. Fixing this by cutting one row.")
        Nj_ncap = Nj_ncap - 1
    Ni, Nj_ncap = int(Ni), int(Nj_ncap)
    rp = bipolar_rp(lat0_bp)
    lams, phis = _dev.empty(Nj_ncap + 1, Ni + 1), _dev.empty(Nj_ncap + 1, Ni + 1)
    h_i_inv, h_j_inv = _dev.empty(Nj_ncap + 1, Ni), _dev.empty(Nj_ncap, Ni + 1)
    _cabi.call("ogg_bipolar_cap_mesh", Ni, Nj_ncap, 0, Nj_ncap + 1, float(lat0_bp), float(lon_bp), rp,
               _dev.ptr(lams), _dev.ptr(phis), _dev.ptr(h_i_inv), _dev.ptr(h_j_inv), _dev.stream())
    print("   number of js=", phis.shape[0])
    return _out(lams, device), _out(phis, device), _out(h_i_inv, device), _out(h_j_inv, device)


def bipolar_cap_ij_array(i, j, Ni, Nj_ncap, lat0_bp, lon_bp, rp):
    """OGG:125-133."""
    on_dev = _any_dev(i, j)
    ti, tj = _dev.to_dev(i), _dev.to_dev(j)
    hi, hj = _dev.empty(tj.numel(), ti.numel()), _dev.empty(tj.numel(), ti.numel())
    _cabi.call("ogg_bipolar_cap_ij_array", _dev.ptr(ti), ti.numel(), _dev.ptr(tj), tj.numel(), int(Ni), int(Nj_ncap),
               float(lat0_bp), float(lon_bp), float(rp), _dev.ptr(hi), _dev.ptr(hj), _dev.stream())
    return _out(hi, on_dev), _out(hj, on_dev)


def bipolar_cap_metrics_quad_fast(order, nx, ny, lat0_bp, lon_bp, rp, Re=_default_Re, device=False):
    """OGG:136-188."""
    print("   Calculating bipolar cap metrics via quadrature ...")
    nx, ny = int(nx), int(ny)
    dxq, dyq, daq = _dev.empty(ny + 1, nx), _dev.empty(ny, nx + 1), _dev.empty(ny, nx)
    _cabi.call("ogg_bipolar_cap_metrics_quad", int(order), nx, ny, 0, ny, float(lat0_bp), float(lon_bp), float(rp),
               float(Re), 0, _dev.ptr(dxq), _dev.ptr(dyq), _dev.ptr(daq), _dev.stream())
    return _out(dxq, device), _out(dyq, device), _out(daq, device)


# --------------------------------------------------------------------------------------------------
# quadrature helpers (OGG:191-255)
# --------------------------------------------------------------------------------------------------
def quad_positions(n=3):
    """OGG:191-204."""
    import ctypes
    a = (ctypes.c_double * 5)()
    b = (ctypes.c_double * 5)()
    _cabi.check(_cabi.load().ogg_quad_positions(int(n), a, b), "ogg_quad_positions")
    return numpy.array(a[:n]), numpy.array(b[:n])


def quad_average(y):
    """OGG:207-222."""
    on_dev = _dev.is_dev(y)
    ty = _dev.to_dev(y).reshape(-1)
    if ty.numel() not in (2, 3, 4, 5):
        raise Exception("Uncoded order")
    out = _dev.empty(1)
    _cabi.call("ogg_quad_average", _dev.ptr(ty), 1, ty.numel(), _dev.ptr(out), _dev.stream())
    return out[0] if on_dev else numpy.float64(_dev.to_host(out)[0])


def quad_average_2d(y):
    """OGG:225-255."""
    on_dev = _dev.is_dev(y)
    if y.shape[0] != y.shape[1]:
        raise Exception("Input array is not squared!")
    if y.shape[0] not in (2, 3, 4, 5):
        raise Exception("Uncoded order")
    ty = _dev.to_dev(y)
    out = _dev.empty(1)
    _cabi.call("ogg_quad_average_2d", _dev.ptr(ty), 1, int(y.shape[0]), _dev.ptr(out), _dev.stream())
    return out[0] if on_dev else numpy.float64(_dev.to_host(out)[0])


# --------------------------------------------------------------------------------------------------
# Mercator band: 1-D latitude law on the host (O(Nj) values; trig on the device), 2-D tile on the device
# --------------------------------------------------------------------------------------------------
def _lagrange(x, y, q):
    """Lagrange interpolation through (x[k], y[k]) in the reference's operation order: each basis
    numerator/denominator is a left-to-right product, terms are paired (k, n-1-k) outside-in
    (OGG:258-289 for n=4 and n=6)."""
    n = len(x)
    terms = []
    for k in range(n):
        others = [m for m in range(n) if m != k]
        num = q - x[others[0]]
        den = x[k] - x[others[0]]
        for m in others[1:]:
            num = num * (q - x[m])
            den = den * (x[k] - x[m])
        terms.append((num / den) * y[k])
    acc = terms[0] + terms[n - 1]
    for k in range(1, n // 2):
        acc = acc + (terms[k] + terms[n - 1 - k])
    return acc


def lagrange_interp(x, y, q):
    """OGG:258-270 (four points)."""
    return _lagrange(list(x[:4]), list(y[:4]), q)


def lagrange_interp_6pt(x, y, q):
    """OGG:272-289 (six points)."""
    return _lagrange(list(x[:6]), list(y[:6]), q)


def _mercator_law(name, v, R, y_first=None, n=None):
    """One fused launch of a 1-D Mercator law (ogg_y_mercator / ogg_phi_mercator); numpy in -> numpy out."""
    if v is None:
        out = _dev.empty(int(n))
        _cabi.call(name, None, int(n), float(y_first), float(R), _dev.ptr(out), _dev.stream())
        return _dev.to_host(out)
    scalar = numpy.ndim(v) == 0 and not _dev.is_dev(v)
    t = _dev.to_dev(numpy.atleast_1d(v) if scalar else v)
    out = _dev.empty(*t.shape)
    if name == "ogg_y_mercator":
        _cabi.call(name, _dev.ptr(t), t.numel(), float(R), _dev.ptr(out), _dev.stream())
    else:
        _cabi.call(name, _dev.ptr(t), t.numel(), 0.0, float(R), _dev.ptr(out), _dev.stream())
    if _dev.is_dev(v):
        return out
    h = _dev.to_host(out)
    return numpy.float64(h[0]) if scalar else h


def y_mercator(Ni, phi):
    """Equation (1) (OGG:292-295): R * log((1 + sin(phi)) / cos(phi))."""
    R = Ni / (2 * np.pi)
    return _mercator_law("ogg_y_mercator", phi, R)


def phi_mercator(Ni, y):
    """Equation (2) (OGG:298-301), degrees: arctan(sinh(y / R)) * (180 / pi)."""
    R = Ni / (2 * np.pi)
    return _mercator_law("ogg_phi_mercator", y, R)


def d_phi_mercator_dy(Ni, y):
    """OGG:303-306."""
    R = Ni / (2 * np.pi)
    return np.cosh(y / R) / (1. + (np.sinh(y / R)) ** 2) / R


def y_mercator_rounded(Ni, phi):
    """OGG:309-311."""
    y_float = y_mercator(Ni, phi)
    return (numpy.sign(y_float) * numpy.round(numpy.abs(y_float))).astype(int)


def _last_below(v, bound):
    return numpy.where(v < bound)[0][-1]


def mercator_latitudes(Ni, phi_s, phi_n, refineR, shift_equator_to_u_point=True, ensure_nj_even=True,
                       enhanced_equatorial=0):
    """The 1-D latitude vector of generate_mercator_grid (OGG:316-428), before the final row-parity cut."""
    print("Requesting Mercator grid with phi range: phi_s,phi_n=", phi_s, phi_n)
    y_star = y_mercator_rounded(Ni, numpy.array([phi_s * PI_180, phi_n * PI_180]))
    print("   y*=", y_star, "nj=", y_star[1] - y_star[0] + 1)
    if y_star[0] % 2 == 0:                                            # equator must be a u-point (OGG:321-328)
        print("  *Equator may not be a u-point!")
        if shift_equator_to_u_point:
            print("  *Fixing this by shifting the bounds!")
            y_star = y_star - 1
            print("   y*=", y_star, "nj=", y_star[1] - y_star[0] + 1)
    if (y_star[1] - y_star[0] + 1) % 2 == 0:                          # OGG:329-333
        print("  *Supergrid has an odd number of area cells!")
        if ensure_nj_even:
            print("  *Fixing this by shifting the y_star[1] ")
            y_star[1] = y_star[1] - 1
    phi_M = _mercator_law("ogg_phi_mercator", None, Ni / (2 * np.pi), y_first=y_star[0], n=y_star[1] - y_star[0] + 1)  # OGG:336
    # OGG:335 prints phi_mercator(Ni, y_star): element-wise, so exactly the two end points of phi_M
    print("   Generating Mercator grid with phi range: phi_s,phi_n=", phi_M[[0, -1]])
    equator_index = numpy.searchsorted(phi_M, 0.0)
    if equator_index == 0:
        raise Exception("   Ooops: Equator is not in the grid")
    print("   Equator is at j=", equator_index)
    if equator_index % 2 == 0:
        print("  *Equator is not going to be a u-point of this grid patch.")
    if not enhanced_equatorial:
        return phi_M

    # Enhanced equatorial resolution (OGG:349-428): Mercator south part, Lagrange shoulder, uniform core, mirror.
    print("   Enhancing the equator region resolution")
    six_point = False
    phi_enh_d, phi_cub_d = -5.0, -30
    N_cub = 132 * refineR / 2
    dphi_e = 0.13 * 2 / refineR
    N_enh = 40 * refineR / 2
    if refineR == 1:                                                   # closest to SPEAR (OGG:367-372)
        phi_enh_d, phi_cub_d, N_cub, N_enh = -10, -20, 29, 55
        dphi_e = -phi_enh_d / N_enh / 0.981
    if refineR == 4 and enhanced_equatorial == 8:                      # OGG:374-382
        phi_enh_d, phi_cub_d, N_cub = -10, -20, 101
        N_enh = 2 * enhanced_equatorial * abs(phi_enh_d) + 1
        dphi_e = -phi_enh_d / N_enh
        six_point = True
    if refineR == 4 and enhanced_equatorial == 6:                      # OGG:384-390
        phi_enh_d, phi_cub_d, N_cub = -10, -20, 101
        N_enh = 2 * enhanced_equatorial * abs(phi_enh_d) + 1
        dphi_e = -phi_enh_d / N_enh
    jc = _last_below(phi_M, phi_cub_d)
    south = phi_M[0:jc]
    phi_e = phi_enh_d
    if not six_point:
        nodes = [0, 1, N_cub - 2, N_cub - 1]
        vals = [phi_M[jc - 1], phi_M[jc], phi_e - dphi_e, phi_e]
        shoulder = lagrange_interp(nodes, vals, numpy.arange(N_cub))
    else:
        N_cub = 111
        nodes = [0, 1, 2, N_cub - 3, N_cub - 2, N_cub - 1]
        vals = [phi_M[jc - 1], phi_M[jc], phi_M[jc + 1], phi_e - dphi_e, phi_e, phi_e + dphi_e]
        shoulder = lagrange_interp_6pt(nodes, vals, numpy.arange(N_cub))
    print("   Meridional range of pure Mercator=(", south[0], ",", south[-2], ") U (", -south[-2], ",", -south[0], ").")
    print("   Meridional range of cubic interpolation=(", shoulder[0], ",", shoulder[-2], ") U (", -shoulder[-2], ",",
          -shoulder[0], ").")
    joined = numpy.concatenate((south[0:-1], shoulder))
    core = numpy.linspace(joined[-1], 0, int(N_enh))
    print("   Meridional range of enhanced resolution=(", core[0], ",", -core[0], ").")
    print("   Meridional value of enhanced resolution=", core[1] - core[0])
    half = numpy.concatenate((joined[0:-1], core))
    phi_M = numpy.concatenate((half[0:-1], -half[::-1]))               # symmetric about the equator
    return phi_M[0:_last_below(phi_M, phi_n)]                          # limit by the requested phi_n (OGG:426-427)


def generate_mercator_grid(Ni, phi_s, phi_n, lon0_M, lenlon_M, refineR, shift_equator_to_u_point=True,
                           ensure_nj_even=True, enhanced_equatorial=0, device=False):
    """OGG:314-441."""
    phi_M = mercator_latitudes(Ni, phi_s, phi_n, refineR, shift_equator_to_u_point, ensure_nj_even,
                               enhanced_equatorial)
    lam_M = lon0_M + numpy.arange(Ni + 1) * lenlon_M / float(Ni)
    if phi_M.shape[0] % 2 == 0 and ensure_nj_even:                     # OGG:434-437
        print("   The number of j's is not even. Fixing this by cutting one row at south.")
        phi_M = phi_M[1:]
    x, y = _tile_band(lam_M, phi_M)
    print("   Final Mercator grid range=", float(phi_M[0]), float(phi_M[-1]))
    print("   number of js=", phi_M.shape[0])
    return _out(x, device), _out(y, device)


def _tile_band(lam_1d, phi_1d):
    """np.tile of the two 1-D laws (OGG:430-432, 840-841) as a device fill."""
    tl, tp = _dev.to_dev(lam_1d), _dev.to_dev(phi_1d)
    nyp, nxp = tp.numel(), tl.numel()
    x, y = _dev.empty(nyp, nxp), _dev.empty(nyp, nxp)
    _cabi.call("ogg_band_fill_metrics", _dev.ptr(tl), _dev.ptr(tp), nyp, nxp, _default_Re, _cabi.OGG_BAND_NO_METRICS,
               _dev.ptr(x), _dev.ptr(y), None, None, None, None, _dev.stream())
    return x, y


def latlon_vectors(lni, lnj, llon0, llen_lon, llat0, llen_lat, ensure_nj_even=True):
    """1-D laws of generate_latlon_grid (OGG:834-838)."""
    llonSP = llon0 + numpy.arange(lni + 1) * llen_lon / float(lni)
    llatSP = llat0 + numpy.arange(lnj + 1) * llen_lat / float(lnj)
    if llatSP.shape[0] % 2 == 0 and ensure_nj_even:
        print("   The number of j's is not even. Fixing this by cutting one row at south.")
        llatSP = numpy.delete(llatSP, 0, 0)
    return llonSP, llatSP


def generate_latlon_grid(lni, lnj, llon0, llen_lon, llat0, llen_lat, ensure_nj_even=True, device=False):
    """OGG:832-846."""
    print("Generating regular lat-lon grid between latitudes ", llat0, llat0 + llen_lat)
    lam1, phi1 = latlon_vectors(lni, lnj, llon0, llen_lon, llat0, llen_lat, ensure_nj_even)
    x, y = _tile_band(lam1, phi1)
    print("   generated regular lat-lon grid between latitudes ", float(phi1[0]), float(phi1[-1]))
    print("   number of js=", phi1.shape[0])
    return _out(x, device), _out(y, device)


# --------------------------------------------------------------------------------------------------
# displaced-pole cap (OGG:447-601)
# --------------------------------------------------------------------------------------------------
def displacedPoleCap_projection(lon_grid, lat_grid, z_0, r_joint):
    """OGG:447-467."""
    on_dev = _any_dev(lon_grid, lat_grid)
    tl, tp = _dev.to_dev(lon_grid), _dev.to_dev(lat_grid)
    lam, phi = _dev.empty(*tl.shape), _dev.empty(*tl.shape)
    z_0 = complex(z_0)
    _cabi.call("ogg_dp_projection", _dev.ptr(tl), _dev.ptr(tp), tl.shape[0], tl.shape[1], z_0.real, z_0.imag,
               float(r_joint), _dev.ptr(lam), _dev.ptr(phi), _dev.stream())
    return _out(lam, on_dev), _out(phi, on_dev)


def monotonic_bounding(x, x_0):
    """OGG:470-475; in place, like the reference."""
    if _dev.is_dev(x):
        if x.is_contiguous() and x.dtype == _dev.torch.float64:
            _cabi.call("ogg_monotonic_bounding", _dev.ptr(x), x.shape[0], x.shape[1], float(x_0), _dev.stream())
        else:   # a strided view (e.g. a column slice): scan a contiguous copy, write it back through the view
            t = _dev.to_dev(x).clone()
            _cabi.call("ogg_monotonic_bounding", _dev.ptr(t), t.shape[0], t.shape[1], float(x_0), _dev.stream())
            x.copy_(t)
        return x
    t = _dev.to_dev(x)
    _cabi.call("ogg_monotonic_bounding", _dev.ptr(t), t.shape[0], t.shape[1], float(x_0), _dev.stream())
    x[...] = _dev.to_host(t)
    return x


def displacedPoleCap_baseGrid(i, j, ni, nj, lon0, lat0):
    """OGG:478-485 (1-D host glue)."""
    u = lon0 + i * 360.0 / float(ni)
    a = -90.0
    b = lat0
    v = a + j * (b - a) / float(nj)
    du = numpy.roll(u, shift=-1, axis=0) - u
    dv = numpy.roll(v, shift=-1, axis=0) - v
    return u, v, du, dv


def dp_constants(lat0, lam_pole, r_pole):
    """(r_joint, z_0) of OGG:494-495, evaluated on the device."""
    out = _dev.empty(3)
    _cabi.call("ogg_dp_constants", float(lat0), float(lam_pole), float(r_pole), _dev.ptr(out), _dev.stream())
    h = _dev.to_host(out)
    return float(h[0]), complex(h[1], h[2])


def displacedPoleCap_mesh(i, j, ni, nj, lon0, lat0, lam_pole, r_pole, excluded_fraction=None, device=False):
    """OGG:488-506."""
    on_dev = device or _any_dev(i, j)
    ti, tj = _dev.to_dev(i), _dev.to_dev(j)
    lams, phis = _dev.empty(tj.numel(), ti.numel()), _dev.empty(tj.numel(), ti.numel())
    _cabi.call("ogg_dp_mesh", _dev.ptr(ti), ti.numel(), _dev.ptr(tj), tj.numel(), int(ni), int(nj), float(lon0),
               float(lat0), float(lam_pole), float(r_pole), _dev.ptr(lams), _dev.ptr(phis), _dev.stream())
    londp, latdp = float(lams[0, 0]), float(phis[0, 0])
    if excluded_fraction is not None:
        ny, nx = lams.shape
        jmin = numpy.ceil(excluded_fraction * ny)
        jmin = jmin + numpy.mod(jmin, 2)
        jmint = int(jmin)
        lams, phis = lams[jmint:, :], phis[jmint:, :]
    return _out(lams, on_dev), _out(phis, on_dev), numpy.float64(londp), numpy.float64(latdp)


def generate_displaced_pole_grid(Ni, Nj_scap, lon0, lat0, lon_dp, r_dp, device=False):
    """OGG:509-518."""
    print("Generating displaced pole grid bounded at latitude ", lat0)
    print("   requested displaced pole lon,rdp=", lon_dp, r_dp)
    i_s = numpy.arange(Ni + 1)
    j_s = numpy.arange(Nj_scap + 1)
    x, y, londp, latdp = displacedPoleCap_mesh(i_s, j_s, Ni, Nj_scap, lon0, lat0, lon_dp, r_dp, device=device)
    print("   generated displaced pole lon,lat=", londp, latdp)
    return x, y, londp, latdp


def great_arc_distance(j0, i0, j1, i1, nx, ny, lon0, lat0, lon_dp, r_dp):
    """OGG:522-532."""
    on_dev = _any_dev(j0, i0, j1, i1)
    tj0, ti0, tj1, ti1 = [_dev.to_dev(a) for a in (j0, i0, j1, i1)]
    out = _dev.empty(tj0.numel(), ti0.numel())
    _cabi.call("ogg_great_arc_distance", _dev.ptr(tj0), _dev.ptr(ti0), _dev.ptr(tj1), _dev.ptr(ti1), tj0.numel(),
               ti0.numel(), int(nx), int(ny), float(lon0), float(lat0), float(lon_dp), float(r_dp), _dev.ptr(out),
               _dev.stream())
    return _out(out, on_dev)


def _numerical_h(direction, j, i, nx, ny, lon0, lat0, lon_dp, r_dp, eps, order):
    on_dev = _any_dev(j, i)
    tj, ti = _dev.to_dev(j), _dev.to_dev(i)
    out = _dev.empty(tj.numel(), ti.numel())
    _cabi.call("ogg_numerical_h", direction, _dev.ptr(tj), tj.numel(), _dev.ptr(ti), ti.numel(), int(nx), int(ny),
               float(lon0), float(lat0), float(lon_dp), float(r_dp), float(eps), int(order), _dev.ptr(out), _dev.stream())
    return _out(out, on_dev)


def numerical_hi(j, i, nx, ny, lon0, lat0, lon_dp, r_dp, eps, order=6):
    """OGG:535-547."""
    return _numerical_h(0, j, i, nx, ny, lon0, lat0, lon_dp, r_dp, eps, order)


def numerical_hj(j, i, nx, ny, lon0, lat0, lon_dp, r_dp, eps, order=6):
    """OGG:550-562."""
    return _numerical_h(1, j, i, nx, ny, lon0, lat0, lon_dp, r_dp, eps, order)


def displacedPoleCap_metrics_quad(order, nx, ny, lon0, lat0, lon_dp, r_dp, Re=_default_Re, device=False):
    """OGG:565-601."""
    print("   Calculating displaced pole cap metrics via quadrature ...")
    nx, ny = int(nx), int(ny)
    dxq, dyq, daq = _dev.empty(ny + 1, nx), _dev.empty(ny, nx + 1), _dev.empty(ny, nx)
    _cabi.call("ogg_dp_metrics_quad", int(order), nx, ny, 0, ny, float(lon0), float(lat0), float(lon_dp), float(r_dp),
               float(Re), 0, _dev.ptr(dxq), _dev.ptr(dyq), _dev.ptr(daq), _dev.stream())
    return _out(dxq, device), _out(dyq, device), _out(daq, device)


def cut_below(lam, phi, lowerlat):
    """OGG:604-611."""
    nj = lam.shape[0]
    jmin = nj - 1
    for j in range(0, nj):
        if phi[j, 0] > lowerlat:
            jmin = j
            break
    return lam[jmin:, :], phi[jmin:, :]


def cut_above(lam, phi, upperlat):
    """OGG:614-621."""
    nj = lam.shape[0]
    jmax = nj - 1
    for j in range(0, nj):
        if phi[j, 0] > upperlat:
            jmax = j
            break
    return lam[0:jmax, :], phi[0:jmax, :]


# --------------------------------------------------------------------------------------------------
# spherical metrics (OGG:682-770)
# --------------------------------------------------------------------------------------------------
def mdist(x1, x2):
    """Positive distance modulo 360 (OGG:682-684)."""
    if _any_dev(x1, x2):
        import torch
        a, b = np.mod360(x1 - x2), np.mod360(x2 - x1)
        return torch.minimum(a, b)
    x1, x2 = numpy.asarray(x1, dtype=numpy.float64), numpy.asarray(x2, dtype=numpy.float64)
    return numpy.minimum(np.mod360(x1 - x2), np.mod360(x2 - x1))


def generate_grid_metrics_MIDAS(x, y, axis_units="degrees", Re=_default_Re, latlon_areafix=True, halo=None):
    """OGG:687-716.  halo=(x_row, y_row): the mesh row owned by the next row-slab (dy, area then get one more row)."""
    on_dev = _any_dev(x, y)
    tx, ty = _dev.to_dev(x), _dev.to_dev(y)
    nyp, nxp = tx.shape
    hx = hy = None
    extra = 0
    if halo is not None:
        hx, hy = _dev.to_dev(halo[0]).reshape(-1), _dev.to_dev(halo[1]).reshape(-1)
        extra = 1
    dx, dy, area = _dev.empty(nyp, nxp - 1), _dev.empty(nyp - 1 + extra, nxp), _dev.empty(nyp - 1 + extra, nxp - 1)
    _cabi.call("ogg_metrics_midas", _dev.ptr(tx), _dev.ptr(ty), nyp, nxp, float(Re), int(bool(latlon_areafix)),
               _dev.ptr(hx), _dev.ptr(hy), _dev.ptr(dx), _dev.ptr(dy), _dev.ptr(area), _dev.stream())
    return _out(dx, on_dev), _out(dy, on_dev), _out(area, on_dev)


def angle_x(x, y):
    """Orientation angle of the grid box (OGG:719-729)."""
    if tuple(x.shape) != tuple(y.shape):
        raise Exception("Input arrays do not have the same shape!")
    on_dev = _any_dev(x, y)
    tx, ty = _dev.to_dev(x), _dev.to_dev(y)
    out = _dev.empty(*tx.shape)
    _cabi.call("ogg_angle_x", _dev.ptr(tx), _dev.ptr(ty), tx.shape[0], tx.shape[1], _dev.ptr(out), _dev.stream())
    return _out(out, on_dev)


def device_sum(a, n=None, stride=1, offset=0):
    """Sum of a.flat[offset + k*stride], k < n, on the device (fixed-shape tree)."""
    t = _dev.to_dev(a)
    flat = t.reshape(-1)
    if n is None:
        n = (flat.numel() - offset + stride - 1) // stride
    out = _dev.empty(1)
    _cabi.call("ogg_sum_strided", _dev.ptr_off(flat, offset), int(n), int(stride), _dev.ptr(out), _dev.stream())
    return float(out[0])


def metrics_error(dx_, dy_, area_, Ni, lat1, lat2=90, Re=_default_Re, bipolar=False, displaced_pole=-999,
                  excluded_fraction=None):
    """OGG:732-770: %% errors of the summed area / arc lengths against the exact spherical values."""
    def col_sum(a, c):
        return device_sum(a, n=a.shape[0], stride=a.shape[1], offset=c)

    def row_sum(a, r):
        r = r % a.shape[0]
        return device_sum(a, n=a.shape[1], stride=1, offset=r * a.shape[1])

    sums = {"area": device_sum(area_), "dy_quarter": col_sum(dy_, Ni // 4), "dx_first": row_sum(dx_, 0)}
    if lat1 > lat2 or bipolar:
        sums["dx_last"] = row_sum(dx_, -1)
    if displaced_pole != -999:
        antipole = displaced_pole + Ni // 2
        if displaced_pole > Ni // 2:
            antipole = displaced_pole - Ni // 2
        sums["dy_pole"], sums["dy_antipole"] = col_sum(dy_, displaced_pole), col_sum(dy_, antipole)
    return metrics_error_from_sums(sums, lat1, lat2, Re, bipolar, displaced_pole, excluded_fraction)


def metrics_error_from_sums(sums, lat1, lat2=90, Re=_default_Re, bipolar=False, displaced_pole=-999, excluded_fraction=None):
    """The arithmetic of OGG:743-770 on the sums OGG:747-756 takes from the arrays: sums = {"area": sum(area_),
    "dy_quarter": sum(dy_[:, Ni//4]), "dx_first": sum(dx_[0]), "dx_last": sum(dx_[-1]), "dy_pole" / "dy_antipole":
    sum(dy_[:, displaced_pole]) / sum(dy_[:, antipole])}.  Row-slabs / row chunks add their partial sums first."""
    exact_area = 2 * np.pi * (Re ** 2) * numpy.abs(np.sin(lat2 * PI_180) - np.sin(lat1 * PI_180))
    exact_lat_arc_length = numpy.abs(lat2 - lat1) * PI_180 * Re
    exact_lon_arc_length = np.cos(lat1 * PI_180) * 2 * np.pi * Re
    grid_lat_arc_length = sums["dy_quarter"]
    grid_lon_arc_length = sums["dx_first"]
    if lat1 > lat2:
        grid_lon_arc_length = sums["dx_last"]
    if bipolar:
        grid_lon_arc_length2 = sums["dx_last"]
        lon_arc2_error = 100 * (grid_lon_arc_length2 / 4 - exact_lat_arc_length) / exact_lat_arc_length
    area_error = 100 * (sums["area"] - exact_area) / exact_area
    lat_arc_error = 100 * (grid_lat_arc_length - exact_lat_arc_length) / exact_lat_arc_length
    lon_arc_error = 100 * (grid_lon_arc_length - exact_lon_arc_length) / exact_lon_arc_length
    if displaced_pole != -999:
        grid_lat_arc_length = sums["dy_pole"] + sums["dy_antipole"]
        lat_arc_error = 100 * (grid_lat_arc_length - 2.0 * exact_lat_arc_length) / exact_lat_arc_length
    if excluded_fraction:
        print("   Cannot estimate area and dy accuracies with excluded_fraction (doughnut)! ")
    if bipolar:
        return area_error, lat_arc_error, lon_arc_error, lon_arc2_error
    return area_error, lat_arc_error, lon_arc_error


def displacedPoleCap_metrics_sums(order, nx, ny, lon0, lat0, lon_dp, r_dp, displaced_pole, Re=_default_Re, max_rows=64):
    """The sums metrics_error (OGG:747-756, 763-765) takes from displacedPoleCap_metrics_quad's arrays over the WHOLE cap,
    accumulated over chunks of `max_rows` cell rows in bounded scratch -- no array of the full cap is ever held.  This is
    how main()'s `CHECK_metrics_hquad` line for the southern cap (OGG:1170-1174) is evaluated when only the rows kept
    after the doughnut cut are generated."""
    nx, ny = int(nx), int(ny)
    antipole = displaced_pole + nx // 2
    if displaced_pole > nx // 2:
        antipole = displaced_pole - nx // 2
    rows = min(int(max_rows), ny)
    dxq, dyq, daq = _dev.empty(rows + 1, nx), _dev.empty(rows, nx + 1), _dev.empty(rows, nx)
    sums = {"area": 0.0, "dy_quarter": 0.0, "dx_first": 0.0, "dx_last": 0.0, "dy_pole": 0.0, "dy_antipole": 0.0}
    for j0 in range(0, ny, rows):
        j1 = min(ny, j0 + rows)
        n = j1 - j0
        _cabi.call("ogg_dp_metrics_quad", int(order), nx, ny, j0, j1, float(lon0), float(lat0), float(lon_dp), float(r_dp),
                   float(Re), 0, _dev.ptr(dxq), _dev.ptr(dyq), _dev.ptr(daq), _dev.stream())
        sums["area"] += device_sum(daq, n=n * nx)
        for key, col in (("dy_quarter", nx // 4), ("dy_pole", displaced_pole), ("dy_antipole", antipole)):
            sums[key] += device_sum(dyq, n=n, stride=nx + 1, offset=col)
        if j0 == 0:
            sums["dx_first"] = device_sum(dxq, n=nx)
        if j1 == ny:
            sums["dx_last"] = device_sum(dxq, n=nx, offset=n * nx)
    return sums


# --------------------------------------------------------------------------------------------------
# file output (OGG:773-829)
# --------------------------------------------------------------------------------------------------
_STAGING = {}


class _DeviceVariable(object):
    """One output variable resident in HBM, handed to the CDF-2 writer as a producer of big-endian bytes.

    The byte order is reversed on the GPU (ogg_bswap64) with the kernel's stores landing directly in pinned host memory,
    in chunks through two staging buffers: the device fills one while the host writes the other to the file.  Replaces
    the reference path's host copy + netCDF-C's per-element byte swap (write_nc, OGG:795-821)."""
    CHUNK = 1 << 23     # elements per staging buffer (64 MiB)

    def __init__(self, t):
        self.t = t.contiguous()
        self.shape = tuple(t.shape)

    def bigendian_chunks(self):
        import torch
        flat = self.t.reshape(-1)
        n = flat.numel()
        key = torch.cuda.current_device()
        if key not in _STAGING:
            _STAGING[key] = ([torch.empty(self.CHUNK, dtype=torch.int64, pin_memory=True) for _ in range(2)],
                             [torch.cuda.Event() for _ in range(2)], torch.cuda.Stream())
        bufs, events, side = _STAGING[key]
        side.wait_stream(torch.cuda.current_stream())      # the producers of self.t
        starts = list(range(0, n, self.CHUNK))

        def launch(k):
            e0 = starts[k]
            m = min(self.CHUNK, n - e0)
            with torch.cuda.stream(side):
                _cabi.call("ogg_bswap64", _dev.ptr_off(flat, e0), _dev.ptr(bufs[k % 2]), m, _dev.stream())
                events[k % 2].record(side)
            return m

        sizes = {0: launch(0)} if starts else {}
        for k in range(len(starts)):
            if k + 1 < len(starts):
                sizes[k + 1] = launch(k + 1)    # the other buffer: its previous contents were written out one step ago
            events[k % 2].synchronize()
            yield memoryview(bufs[k % 2].numpy()).cast("B")[:8 * sizes[k]]


def write_nc(x, y, dx, dy, area, angle_dx, axis_units="degrees", fnam=None, format="NETCDF3_64BIT", description=None,
             history=None, source=None, no_changing_meta=None, debug=False):
    """Same variables, dimensions, attributes and CDF-2 container as the reference's netCDF4 call sequence
    (OGG:773-829).  Device tensors are streamed out big-endian by the GPU (see _DeviceVariable); numpy arrays are
    converted on the host."""
    if fnam is None:
        fnam = "supergrid.nc"
    arrs = [x, y, dx, dy, area, angle_dx]
    if debug:
        for a, lbl in zip(arrs, ("x", "y", "dx", "dy", "area", "angle_dx")):
            chksum(_dev.to_host(a) if _dev.is_dev(a) else numpy.asarray(a), lbl)
    arrs = [_DeviceVariable(a) if _dev.is_dev(a) else numpy.asarray(a) for a in arrs]
    print("   Writing netcdf file with ny,nx= ", arrs[4].shape[0], arrs[4].shape[1])
    attrs = None
    if not no_changing_meta:
        attrs = [("history", history), ("description", description), ("source", source)]
    _netcdf3.write_supergrid(fnam, *arrs, attrs, format=format)


def usage():
    print("ocean_grid_generator.py -f <output_grid_filename> -r <inverse_degrees_resolution> [--rdp=<displacement_factor/0.2> --exfracdp=0.5 --south_cutoff_ang=<degrees_south_to_start> --south_cutoff_row=<rows_south_to_cut> --match_dy bp so --even_j --plot --write_subgrid_files --enhanced_equatorial --no-metrics --grids=sc]")


def main(inverse_resolution, gridfilename="ocean_hgrid.nc", r_dp=0.0, lon_dp=80.0, lat_dp=-99.0, exfracdp=None,
         south_cutoff_row=0, south_cutoff_ang=-90.0, reproduce_MIDAS_grids=False, write_subgrid_files=False,
         plotem=False, no_changing_meta=False, enhanced_equatorial=0, debug=False,
         grids=["bipolar", "mercator", "so", " sc", "all"], match_dy=["bp", "so", " sc", "p125sc", ""],
         skip_metrics=False, ensure_nj_even=False, shift_equator_to_u_point=True, bipolar_lower_lat=-99.0,
         mercator_lower_lat=-99.0, mercator_upper_lat=-99.0, south_ocean_lower_lat=-99.0,
         south_ocean_upper_lat=-99.0, no_south_cap=False):
    """OGG:855-1449: same keyword interface; the work is planned on the host and built in place on the device."""
    from . import supergrid
    return supergrid.run_main(dict(locals()))


def _build_parser():
    """The reference's 23 flags (OGG:1454-1524)."""
    p = argparse.ArgumentParser(description="create ocean hgrid")
    p.add_argument("-r", "--inverse_resolution", type=float, required=True,
                   help="inverse of the horizontal resolution (e.g. 4 for 1/4 degree)")
    p.add_argument("-f", "--gridfilename", type=str, required=False, default="ocean_hgrid.nc",
                   help="name for output grid file")
    p.add_argument("--r_dp", type=float, required=False, default=0.0,
                   help="displacement factor/0.2 for the displaced south pole, do not specify both r_dp and lat_dp!")
    p.add_argument("--exfracdp", type=float, required=False, default=0.49,
                   help="exclusion factor that determines the size of the hole arount SP!")
    p.add_argument("--lon_dp", type=float, required=False, default=80.0, help="longitude of the displaced south pole")
    p.add_argument("--lat_dp", type=float, required=False, default=-99.0,
                   help="latitude of the displaced south pole, do not specify both r_dp and lat_dp!")
    p.add_argument("--south_cutoff_ang", type=float, required=False, default=-90.0, help="degrees south to start")
    p.add_argument("--south_cutoff_row", type=int, required=False, default=0, help="rows to cut from the grid at south")
    p.add_argument("--bipolar_lower_lat", type=float, required=False, default=-90.0,
                   help="starting (lower) latitude of Northern Bipolar sub grid")
    p.add_argument("--mercator_lower_lat", type=float, required=False, default=-90.0,
                   help="starting (lower) latitude of Mercator sub grid")
    p.add_argument("--mercator_upper_lat", type=float, required=False, default=-99.0,
                   help="ending (upper) latitude of Mercator sub grid")
    p.add_argument("--south_ocean_lower_lat", type=float, required=False, default=-90.0,
                   help="starting (lower) latitude of SO sub grid")
    p.add_argument("--south_ocean_upper_lat", type=float, required=False, default=-99.0,
                   help="ending (upper) latitude of SO sub grid")
    p.add_argument("--no_south_cap", action="store_true", help="do not generate a southern cap sub grid")
    p.add_argument("--match_dy", type=str, nargs="+", required=False, default=[],
                   help="set the number of j-points of subgrid such that latitude resolution (dy) becomes continous "
                        "at the joints: bp, so, p125sc")
    p.add_argument("--ensure_nj_even", action="store_true", required=False, default=False,
                   help="error out if the number of j-points in any subgrid is not even")
    p.add_argument("--plotem", action="store_true", help="make a rudimentary plot of the subgrids")
    p.add_argument("--skip_metrics", action="store_true",
                   help="skip generating the metrics, only for fast debugging purposes")
    p.add_argument("--write_subgrid_files", action="store_true", help="write subgrids to separate files ")
    p.add_argument("--no_changing_meta", action="store_true",
                   help="do not write meta data to netcdf files that might change")
    p.add_argument("--enhanced_equatorial", type=int, required=False, default=0,
                   help="generate a subgrid that has an enhanced resolution around the equator")
    p.add_argument("--shift_equator_to_u_point", action="store_false", required=False, default=True,
                   help="if the equator is not a u point shift the Mercator subgrid by 1 j-point to make it a u point")
    p.add_argument("--grids", type=str, nargs="+", required=False, default="all",
                   help="specify the subgrids to generate, choices are bipolar, mercator, so, sc, all. Default is all")
    return p


def cli(argv=None):
    args = vars(_build_parser().parse_args(argv))
    main(**args)


if __name__ == "__main__":
    cli()

"""Portable (bit-reproducible) elementary functions in pure numpy -- ORACLE / TEST INFRASTRUCTURE.

Stand-in for `numpypi.numpypi_series`, which the reference imports as `np`
(/root/reference/ocean_grid_generator.py:5) but which is not vendored, not installed and not
version-pinned (setup.cfg:22 names a branch).  PARITY UNPINNED against upstream numpypi: the
polynomials below are this repo's own specification (DESIGN.md "Portable-trig spec"), NOT
upstream's, so the reference's whole-file md5/sha256 pins cannot be reproduced.

Contract (the same one the C oracle `oracle/ogg_oracle.c` and the CUDA device header
`ocean_model_grid_generator_b200/csrc/ogg_trig.cuh` implement, function by function, bit for bit):
  * IEEE-754 binary64, round-to-nearest-even, every + - * / sqrt individually rounded
    (no FMA contraction), evaluation order exactly as written below;
  * rint = round-half-even, selects by comparison only;
  * coefficient sets are the classic minimax sets published with Sun's freely distributable
    fdlibm (1993); reduction / evaluation order / branch structure are simplified and fixed here.
Everything that is not overridden is re-exported from numpy, as the reference expects
(np.tile, np.where, np.float64, np.pi ... are reached through this module).

Overridden beyond the seven trig functions, so that the oracle is deterministic on any box:
  log, sinh, cosh, exp (used by the 1-D Mercator latitude law, reference :292-311),
  angle and absolute for complex input (used by displacedPoleCap_projection, reference :457,:465).
"""
import numpy as _np
from numpy import *  # noqa: F401,F403  (re-export, see module docstring)

_f8 = _np.float64

# ---- constants (hex literals are the definition) -----------------------------------------
_TWO_OVER_PI = float.fromhex("0x1.45f306dc9c883p-1")   # fl(2/pi)
_P1 = float.fromhex("0x1.921fb54400000p+0")            # pi/2, leading 33 bits
_P2 = float.fromhex("0x1.0b4611a600000p-34")           # next 33 bits
_P3 = float.fromhex("0x1.3198a2e037073p-69")           # fl(pi/2 - P1 - P2)
_PIO2_HI = float.fromhex("0x1.921fb54442d18p+0")       # fl(pi/2)
_PIO2_LO = 6.12323399573676603587e-17                  # fl(pi/2 - PIO2_HI)
_PI = 3.141592653589793                                # fl(pi) == numpy.pi
_PI_LO = 1.2246467991473532e-16                        # fl(pi - PI)
_S = (-1.66666666666666324348e-01, 8.33333333332248946124e-03, -1.98412698298579493134e-04,
      2.75573137070700676789e-06, -2.50507602534068634195e-08, 1.58969099521155010221e-10)
_C = (4.16666666666666019037e-02, -1.38888888888741095749e-03, 2.48015872894767294178e-05,
      -2.75573143513906633035e-07, 2.08757232129817482790e-09, -1.13596475577881948265e-11)
_ATANHI = (4.63647609000806093515e-01, 7.85398163397448278999e-01,
           9.82793723247329054082e-01, 1.57079632679489655800e+00)
_ATANLO = (2.26987774529616870924e-17, 3.06161699786838301793e-17,
           1.39033110312309984516e-17, 6.12323399573676603587e-17)
_AT = (3.33333333333329318027e-01, -1.99999999998764832476e-01, 1.42857142725034663711e-01,
       -1.11111104054623557880e-01, 9.09088713343650656196e-02, -7.69187620504482999495e-02,
       6.66107313738753120669e-02, -5.83357013379057348645e-02, 4.97687799461593236017e-02,
       -3.65315727442169155270e-02, 1.62858201153657823623e-02)
_PS = (1.66666666666666657415e-01, -3.25565818622400915405e-01, 2.01212532134862925881e-01,
       -4.00555345006794114027e-02, 7.91534994289814532176e-04, 3.47933107596021167570e-05)
_QS = (-2.40339491173441421878e+00, 2.02094576023350569471e+00,
       -6.88283971605453293030e-01, 7.70381505559019352791e-02)
_LN2_HI = 6.93147180369123816490e-01
_LN2_LO = 1.90821492927058770002e-10
_INV_LN2 = 1.44269504088896338700e+00
_EP = (1.66666666666666019037e-01, -2.77777777770155933842e-03, 6.61375632143793436117e-05,
       -1.65339022054652515390e-06, 4.13813679705723846039e-08)
_LG = (6.666666666666735130e-01, 3.999999999940941908e-01, 2.857142874366239149e-01,
       2.222219843214978396e-01, 1.818357216161805012e-01, 1.531383769920937332e-01,
       1.479819860511658591e-01)
_SQRT_HALF = 0.70710678118654752440


def _arr(x):
    return _np.asarray(x, dtype=_f8)


def _ret(y, x):
    """numpy ufuncs return scalars for scalar input; keep that behaviour."""
    if _np.ndim(x) == 0:
        return _f8(y)
    return y


# ---- sin / cos / tan ---------------------------------------------------------------------
def _reduce(x):
    k = _np.rint(x * _TWO_OVER_PI)
    r = ((x - k * _P1) - k * _P2) - k * _P3
    return k, r


def _sin_k(r, z):
    p = _S[1] + z * (_S[2] + z * (_S[3] + z * (_S[4] + z * _S[5])))
    return r + (z * r) * (_S[0] + z * p)


def _cos_k(r, z):
    rr = z * (_C[0] + z * (_C[1] + z * (_C[2] + z * (_C[3] + z * (_C[4] + z * _C[5])))))
    hz = 0.5 * z
    w = 1.0 - hz
    return w + (((1.0 - w) - hz) + z * rr)


def _sincos(x):
    x = _arr(x)
    k, r = _reduce(x)
    z = r * r
    s = _sin_k(r, z)
    c = _cos_k(r, z)
    n = k.astype(_np.int64) & 3
    return n, s, c


def sin(x):
    n, s, c = _sincos(x)
    y = _np.where(n == 0, s, _np.where(n == 1, c, _np.where(n == 2, -s, -c)))
    return _ret(y, x)


def cos(x):
    n, s, c = _sincos(x)
    y = _np.where(n == 0, c, _np.where(n == 1, -s, _np.where(n == 2, -c, s)))
    return _ret(y, x)


def tan(x):
    n, s, c = _sincos(x)
    odd = (n & 1) == 1
    num = _np.where(odd, -c, s)
    den = _np.where(odd, s, c)
    with _np.errstate(divide="ignore", invalid="ignore"):
        y = num / den
    return _ret(y, x)


# ---- arctan / arctan2 --------------------------------------------------------------------
def _atan_pos(ax):
    """arctan for ax >= 0 (ax may be +inf)."""
    i0 = ax < 0.4375
    i1 = ax < 0.6875
    i2 = ax < 1.1875
    i3 = ax < 2.4375
    with _np.errstate(divide="ignore", invalid="ignore", over="ignore"):
        num = _np.where(i0, ax, _np.where(i1, 2.0 * ax - 1.0, _np.where(i2, ax - 1.0,
                        _np.where(i3, ax - 1.5, -1.0))))
        den = _np.where(i0, 1.0, _np.where(i1, 2.0 + ax, _np.where(i2, ax + 1.0,
                        _np.where(i3, 1.0 + 1.5 * ax, ax))))
        t = num / den
    hi = _np.where(i1, _ATANHI[0], _np.where(i2, _ATANHI[1], _np.where(i3, _ATANHI[2], _ATANHI[3])))
    lo = _np.where(i1, _ATANLO[0], _np.where(i2, _ATANLO[1], _np.where(i3, _ATANLO[2], _ATANLO[3])))
    z = t * t
    w = z * z
    s1 = z * (_AT[0] + w * (_AT[2] + w * (_AT[4] + w * (_AT[6] + w * (_AT[8] + w * _AT[10])))))
    s2 = w * (_AT[1] + w * (_AT[3] + w * (_AT[5] + w * (_AT[7] + w * _AT[9]))))
    ts = t * (s1 + s2)
    return _np.where(i0, t - ts, hi - ((ts - lo) - t))


def arctan(x):
    xa = _arr(x)
    y = _np.copysign(_atan_pos(_np.abs(xa)), xa)
    return _ret(y, x)


def arctan2(y, x):
    ya, xa = _np.broadcast_arrays(_arr(y), _arr(x))
    ay = _np.abs(ya)
    ax = _np.abs(xa)
    with _np.errstate(divide="ignore", invalid="ignore"):
        q = ay / ax
    q = _np.where((ay == 0.0) & (ax == 0.0), 0.0, q)
    a = _atan_pos(q)
    a = _np.where(xa < 0.0, _PI - (a - _PI_LO), a)
    out = _np.copysign(a, ya)
    if _np.ndim(y) == 0 and _np.ndim(x) == 0:
        return _f8(out)
    return out


# ---- arcsin / arccos ---------------------------------------------------------------------
def _asin_r(t):
    p = t * (_PS[0] + t * (_PS[1] + t * (_PS[2] + t * (_PS[3] + t * (_PS[4] + t * _PS[5])))))
    q = 1.0 + t * (_QS[0] + t * (_QS[1] + t * (_QS[2] + t * _QS[3])))
    return p / q


def arcsin(x):
    xa = _arr(x)
    ax = _np.abs(xa)
    small = ax < 0.5
    with _np.errstate(invalid="ignore"):
        t = _np.where(small, ax * ax, (1.0 - ax) * 0.5)
        w = _asin_r(t)
        s = _np.sqrt(t)
        big = _PIO2_HI - (2.0 * (s + s * w) - _PIO2_LO)
    y = _np.where(small, ax + ax * w, big)
    y = _np.where(ax > 1.0, _np.nan, y)
    return _ret(_np.copysign(y, xa), x)


def arccos(x):
    xa = _arr(x)
    ax = _np.abs(xa)
    small = ax < 0.5
    with _np.errstate(invalid="ignore"):
        t = _np.where(small, xa * xa, (1.0 - ax) * 0.5)
        w = _asin_r(t)
        s = _np.sqrt(t)
        y_small = _PIO2_HI - (xa - (_PIO2_LO - xa * w))
        y_pos = 2.0 * (s + s * w)
        y_neg = _PI - 2.0 * (s + (s * w - _PIO2_LO))
    y = _np.where(small, y_small, _np.where(xa > 0.0, y_pos, y_neg))
    y = _np.where(ax > 1.0, _np.nan, y)
    return _ret(y, x)


# ---- exp / log / sinh / cosh (1-D Mercator law only) ---------------------------------------
def exp(x):
    xa = _arr(x)
    k = _np.rint(xa * _INV_LN2)
    hi = xa - k * _LN2_HI
    lo = k * _LN2_LO
    r = hi - lo
    t = r * r
    c = r - t * (_EP[0] + t * (_EP[1] + t * (_EP[2] + t * (_EP[3] + t * _EP[4]))))
    y = 1.0 - ((lo - (r * c) / (2.0 - c)) - hi)
    return _ret(_np.ldexp(y, k.astype(_np.int64)), x)


def log(x):
    xa = _arr(x)
    with _np.errstate(invalid="ignore", divide="ignore"):
        m, e = _np.frexp(xa)                       # xa = m * 2**e, 0.5 <= m < 1
        lowm = m < _SQRT_HALF
        m = _np.where(lowm, 2.0 * m, m)            # sqrt(1/2) <= m < sqrt(2)
        k = (e - lowm.astype(e.dtype)).astype(_f8)
        f = m - 1.0
        s = f / (2.0 + f)
        z = s * s
        w = z * z
        t1 = w * (_LG[1] + w * (_LG[3] + w * _LG[5]))
        t2 = z * (_LG[0] + w * (_LG[2] + w * (_LG[4] + w * _LG[6])))
        R = t2 + t1
        hfsq = 0.5 * (f * f)
        y = k * _LN2_HI - ((hfsq - (s * (hfsq + R) + k * _LN2_LO)) - f)
        y = _np.where(xa == 0.0, -_np.inf, _np.where(xa < 0.0, _np.nan, y))
    return _ret(y, x)


_SH = (1.0 / 6.0, 1.0 / 120.0, 1.0 / 5040.0, 1.0 / 362880.0, 1.0 / 39916800.0, 1.0 / 6227020800.0)


def sinh(x):
    xa = _arr(x)
    ax = _np.abs(xa)
    z = ax * ax
    p = _SH[0] + z * (_SH[1] + z * (_SH[2] + z * (_SH[3] + z * (_SH[4] + z * _SH[5]))))
    small = ax + (ax * z) * p
    t = exp(ax)
    big = 0.5 * (t - 1.0 / t)
    y = _np.where(ax < 0.5, small, big)
    return _ret(_np.copysign(y, xa), x)


def cosh(x):
    xa = _arr(x)
    t = exp(_np.abs(xa))
    return _ret(0.5 * (t + 1.0 / t), x)


# ---- complex helpers used by the displaced-pole projection -----------------------------------
def absolute(x, *args, **kwargs):
    """|x|; for complex input the canonical sqrt(re*re + im*im) (no scaling: |w| = O(1) here)."""
    if _np.iscomplexobj(x) and not args and not kwargs:
        xa = _np.asarray(x)
        re, im = xa.real, xa.imag
        return _ret(_np.sqrt(re * re + im * im), x)
    return _np.absolute(x, *args, **kwargs)


abs = absolute  # noqa: A001  (numpy exports abs as an alias of absolute)


def angle(z, deg=False):
    """Same structure as numpy.angle (arctan2 then *180/pi) on the portable arctan2."""
    za = _np.asarray(z)
    if _np.iscomplexobj(za):
        zim, zre = za.imag, za.real
    else:
        zim, zre = _np.zeros_like(za, dtype=_f8), za
    a = arctan2(zim, zre)
    if deg:
        a = a * (180.0 / _PI)
    return a

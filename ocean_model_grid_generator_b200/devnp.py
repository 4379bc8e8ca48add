"""`numpypi.numpypi_series`-shaped module whose elementary functions run on the GPU.

The reference does `import numpypi.numpypi_series as np` (ocean_grid_generator.py:5) and reaches both the
portable trig and the rest of numpy through that name.  This module keeps that shape for the drop-in's
host glue: sin/cos/tan/arcsin/arccos/arctan/arctan2 (+ exp/log/sinh/cosh, which the 1-D Mercator law
uses) are evaluated by the C-ABI kernels (ogg_sin ... in include/ogg_b200.h) -- the same device code the
big kernels inline -- and everything else is numpy.  There is no host implementation of these functions.
"""
import numpy as _np
from numpy import *  # noqa: F401,F403

from . import _cabi
from . import device as _dev


def _unary(name, x):
    scalar = _np.ndim(x) == 0 and not _dev.is_dev(x)
    t = _dev.to_dev(_np.atleast_1d(x) if scalar else x)
    out = _dev.empty(*t.shape)
    _cabi.call(name, _dev.ptr(t), _dev.ptr(out), t.numel(), _dev.stream())
    if _dev.is_dev(x):
        return out
    h = _dev.to_host(out)
    return _np.float64(h[0]) if scalar else h


def sin(x): return _unary("ogg_sin", x)            # noqa: E704
def cos(x): return _unary("ogg_cos", x)            # noqa: E704
def tan(x): return _unary("ogg_tan", x)            # noqa: E704
def arcsin(x): return _unary("ogg_arcsin", x)      # noqa: E704
def arccos(x): return _unary("ogg_arccos", x)      # noqa: E704
def arctan(x): return _unary("ogg_arctan", x)      # noqa: E704
def exp(x): return _unary("ogg_exp", x)            # noqa: E704
def log(x): return _unary("ogg_log", x)            # noqa: E704
def sinh(x): return _unary("ogg_sinh", x)          # noqa: E704
def cosh(x): return _unary("ogg_cosh", x)          # noqa: E704
def mod360(x): return _unary("ogg_mod360", x)      # noqa: E704


def arctan2(y, x):
    on_dev = _dev.is_dev(y) or _dev.is_dev(x)
    scalar = (not on_dev) and _np.ndim(y) == 0 and _np.ndim(x) == 0
    if on_dev:
        ty, tx = _dev.to_dev(y), _dev.to_dev(x)
    else:
        by, bx = _np.broadcast_arrays(_np.atleast_1d(_np.asarray(y, dtype=_np.float64)),
                                      _np.atleast_1d(_np.asarray(x, dtype=_np.float64)))
        ty, tx = _dev.to_dev(by), _dev.to_dev(bx)
    out = _dev.empty(*ty.shape)
    _cabi.call("ogg_arctan2", _dev.ptr(ty), _dev.ptr(tx), _dev.ptr(out), ty.numel(), _dev.stream())
    if on_dev:
        return out
    h = _dev.to_host(out)
    return _np.float64(h[0]) if scalar else h

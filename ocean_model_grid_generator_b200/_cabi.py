"""ctypes binding of libogg_b200.so (C ABI declared in include/ogg_b200.h).

No CPU fallback lives here or anywhere in the package: a missing library, a missing symbol or a
missing GPU raises immediately.
"""
import ctypes
import os

from . import build as _build

_LIB = None
_dp = ctypes.c_void_p
_i64 = ctypes.c_int64
_d = ctypes.c_double
_int = ctypes.c_int
_vp = ctypes.c_void_p



class OggBand(ctypes.Structure):
    """ogg_band_t of include/ogg_b200.h: one band of ogg_band_fill_metrics_batch (device pointers)."""
    _fields_ = [("lam", ctypes.c_void_p), ("phi", ctypes.c_void_p), ("nyp", ctypes.c_int64), ("nxp", ctypes.c_int64),
                ("flags", ctypes.c_int), ("x", ctypes.c_void_p), ("y", ctypes.c_void_p), ("dx", ctypes.c_void_p),
                ("dy", ctypes.c_void_p), ("area", ctypes.c_void_p), ("angle", ctypes.c_void_p)]


OGG_BAND_SKIP_TOP_VERTEX_ROW = 1
OGG_QUAD_NO_TOP_DX_ROW = 1
OGG_BAND_NO_METRICS = 2

# name -> argtypes, exactly as declared in include/ogg_b200.h
SIGNATURES = {
    "ogg_sin": [_dp, _dp, _i64, _vp], "ogg_cos": [_dp, _dp, _i64, _vp], "ogg_tan": [_dp, _dp, _i64, _vp],
    "ogg_arcsin": [_dp, _dp, _i64, _vp], "ogg_arccos": [_dp, _dp, _i64, _vp], "ogg_arctan": [_dp, _dp, _i64, _vp],
    "ogg_exp": [_dp, _dp, _i64, _vp], "ogg_log": [_dp, _dp, _i64, _vp], "ogg_sinh": [_dp, _dp, _i64, _vp],
    "ogg_cosh": [_dp, _dp, _i64, _vp], "ogg_mod360": [_dp, _dp, _i64, _vp],
    "ogg_arctan2": [_dp, _dp, _dp, _i64, _vp],
    "ogg_y_mercator": [_dp, _i64, _d, _dp, _vp], "ogg_phi_mercator": [_dp, _i64, _d, _d, _dp, _vp],
    "ogg_band_fill_metrics": [_dp, _dp, _i64, _i64, _d, _int, _dp, _dp, _dp, _dp, _dp, _dp, _vp],
    "ogg_band_fill_metrics_batch": [ctypes.POINTER(OggBand), _int, _d, _vp],
    "ogg_metrics_midas": [_dp, _dp, _i64, _i64, _d, _int, _dp, _dp, _dp, _dp, _dp, _vp],
    "ogg_angle_x": [_dp, _dp, _i64, _i64, _dp, _vp],
    "ogg_bipolar_projection": [_dp, _dp, _i64, _d, _d, _int, _dp, _dp, _dp, _dp, _vp],
    "ogg_bipolar_cap_mesh": [_i64, _i64, _i64, _i64, _d, _d, _d, _dp, _dp, _dp, _dp, _vp],
    "ogg_bipolar_cap_ij_array": [_dp, _i64, _dp, _i64, _i64, _i64, _d, _d, _d, _dp, _dp, _vp],
    "ogg_bipolar_cap_metrics_quad": [_int, _i64, _i64, _i64, _i64, _d, _d, _d, _d, _int, _dp, _dp, _dp, _vp],
    "ogg_quad_positions": [_int, ctypes.POINTER(_d), ctypes.POINTER(_d)],
    "ogg_quad_average": [_dp, _i64, _int, _dp, _vp],
    "ogg_quad_average_2d": [_dp, _i64, _int, _dp, _vp],
    "ogg_dp_constants": [_d, _d, _d, _dp, _vp],
    "ogg_dp_projection": [_dp, _dp, _i64, _i64, _d, _d, _d, _dp, _dp, _vp],
    "ogg_monotonic_bounding": [_dp, _i64, _i64, _d, _vp],
    "ogg_dp_mesh": [_dp, _i64, _dp, _i64, _i64, _i64, _d, _d, _d, _d, _dp, _dp, _vp],
    "ogg_dp_cap_mesh": [_i64, _i64, _i64, _i64, _d, _d, _d, _d, _dp, _dp, _vp],
    "ogg_dp_cap_mesh_noalloc": [_i64, _i64, _i64, _i64, _d, _d, _d, _d, _dp, _dp, _vp],
    "ogg_great_arc_distance": [_dp, _dp, _dp, _dp, _i64, _i64, _i64, _i64, _d, _d, _d, _d, _dp, _vp],
    "ogg_numerical_h": [_int, _dp, _i64, _dp, _i64, _i64, _i64, _d, _d, _d, _d, _d, _int, _dp, _vp],
    "ogg_dp_metrics_quad": [_int, _i64, _i64, _i64, _i64, _d, _d, _d, _d, _d, _int, _dp, _dp, _dp, _vp],
    "ogg_sum_strided": [_dp, _i64, _i64, _dp, _vp],
    "ogg_checksum_accumulate": [_dp, _i64, _vp, _vp],
    "ogg_bswap64": [_vp, _vp, _i64, _vp],
    "ogg_fp64_issue_rate": [_i64, ctypes.POINTER(_d)],
    "ogg_selftest_divsqrt": [_i64, ctypes.c_uint64, ctypes.c_int, ctypes.POINTER(ctypes.c_uint64)],
    "ogg_guard_audit": [ctypes.POINTER(ctypes.c_uint64), ctypes.c_int],
    "ogg_launch_count": [], "ogg_timing_enable": [_int], "ogg_timing_reset": [], "ogg_timing_kernels": [],
    "ogg_timing_name": [_int], "ogg_timing_get": [_int, ctypes.POINTER(_d), ctypes.POINTER(ctypes.c_longlong)],
    "ogg_timing_pair_span": [_int, _int, ctypes.POINTER(_d), ctypes.POINTER(ctypes.c_longlong)],
    "ogg_device_count": [],
    "ogg_version": [],
}

OGG_ERR_UNCODED_ORDER = -1
OGG_ERR_ORDER_NOT_CODED = -2
OGG_ERR_SHAPE = -3
OGG_ERR_ARG = -4
OGG_ERR_TOO_LARGE = -5
OGG_ERR_CUDA = -1000
OGG_NOALLOC_SCRATCH_DOUBLES = 262144   # include/ogg_b200.h


class OggError(RuntimeError):
    pass


def library_path():
    # OGG_B200_LIB: developer knob for A/B-timing another build of the same sources (tools/kbench.py); still a CUDA library
    return os.environ.get("OGG_B200_LIB") or _build.LIB


def load(build_if_missing=True):
    """dlopen the library and set every prototype; raises if it cannot be had."""
    global _LIB
    if _LIB is not None:
        return _LIB
    path = library_path()
    if not os.path.exists(path):
        if not build_if_missing:
            raise OggError("libogg_b200.so is missing (%s); run __graft_entry__.build()" % path)
        _build.build()
    lib = ctypes.CDLL(path)
    skipped = []
    for name, argtypes in SIGNATURES.items():
        if os.environ.get("OGG_B200_LIB") and not hasattr(lib, name):
            skipped.append(name)  # developer A/B against an older build of the same sources (tools/kbench.py)
            continue
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.argtypes = argtypes
        fn.restype = (ctypes.c_char_p if name in ("ogg_version", "ogg_timing_name")
                      else ctypes.c_longlong if name == "ogg_launch_count" else _int)
    if skipped:
        import warnings
        warnings.warn("OGG_B200_LIB=%s does not export %s; calling any of them will fail with AttributeError"
                      % (path, ", ".join(skipped)))
    _LIB = lib
    return lib


def check(rc, what=""):
    """Map a status to the reference's exception vocabulary (SURVEY 8(b) error conventions)."""
    if rc == 0:
        return
    if rc == OGG_ERR_UNCODED_ORDER:
        raise Exception("Uncoded order")
    if rc == OGG_ERR_ORDER_NOT_CODED:
        raise Exception("order not coded")
    if rc == OGG_ERR_SHAPE:
        raise Exception("Input arrays do not have the same shape!")
    if rc == OGG_ERR_ARG:
        raise OggError("invalid argument to %s" % what)
    if rc == OGG_ERR_TOO_LARGE:
        raise OggError("%s: the request does not fit the library's fixed %d-double work array (no-allocation entry "
                       "points never grow it); split the rows into smaller ranges or use the allocating entry point"
                       % (what, OGG_NOALLOC_SCRATCH_DOUBLES))
    if rc <= OGG_ERR_CUDA:
        raise OggError("CUDA error %d in %s" % (OGG_ERR_CUDA - rc, what))
    raise OggError("status %d from %s" % (rc, what))


def call(name, *args):
    lib = load()
    check(getattr(lib, name)(*args), name)


def require_gpu():
    lib = load()
    if lib.ogg_device_count() < 1:
        raise OggError("no CUDA device: this package has no CPU path (see oracle/ for the test-only CPU oracle)")

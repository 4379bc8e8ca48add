import hashlib

import numpy as np


def bits_equal(a, b):
    a = np.ascontiguousarray(a, dtype=np.float64)
    b = np.ascontiguousarray(b, dtype=np.float64)
    if a.shape != b.shape:
        return False
    return bool(((a.view(np.int64) == b.view(np.int64)) | (np.isnan(a) & np.isnan(b))).all())


def assert_bits(a, b, what=""):
    a = np.ascontiguousarray(a, dtype=np.float64)
    b = np.ascontiguousarray(b, dtype=np.float64)
    assert a.shape == b.shape, "%s: shape %s vs %s" % (what, a.shape, b.shape)
    bad = ~((a.view(np.int64) == b.view(np.int64)) | (np.isnan(a) & np.isnan(b)))
    if bad.any():
        idx = np.argwhere(bad)[0]
        raise AssertionError("%s: %d of %d values differ; first at %s: %r vs %r" % (
            what, bad.sum(), a.size, idx.tolist(), a[tuple(idx)], b[tuple(idx)]))


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a, dtype=np.float64).tobytes()).hexdigest()

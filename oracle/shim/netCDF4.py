"""Minimal in-memory stand-in for netCDF4-python (TEST INFRASTRUCTURE).

Covers exactly the surface the reference's write_nc uses (ocean_grid_generator.py:773-829):
Dataset(fnam, "w", clobber=, format=), createDimension, createVariable (len(), [:] =, .units),
global attributes, sync, close, stringtochar.  close() leaves the variables in
`netCDF4.LAST[fnam]` (dict of numpy arrays) and, if OGG_SHIM_NPZ=1, in `<fnam>.npz`.
"""
import os
import numpy as _np

LAST = {}


class _Dim:
    def __init__(self, name, size):
        self.name, self.size = name, size

    def __len__(self):
        return self.size


class _Var:
    def __init__(self, ds, name, dtype, dims):
        object.__setattr__(self, "_ds", ds)
        object.__setattr__(self, "_name", name)
        if isinstance(dims, str):
            dims = (dims,)
        shape = tuple(ds._dims[d].size for d in dims)
        object.__setattr__(self, "_shape", shape)
        object.__setattr__(self, "_data", _np.zeros(shape, dtype=dtype))
        object.__setattr__(self, "attrs", {})

    def __len__(self):
        return self._shape[0]

    def __setitem__(self, key, value):
        self._data[key] = value

    def __setattr__(self, k, v):
        self.attrs[k] = v


class Dataset:
    def __init__(self, fnam, mode="r", clobber=True, format="NETCDF4"):
        object.__setattr__(self, "_fnam", str(fnam))
        object.__setattr__(self, "_format", format)
        object.__setattr__(self, "_dims", {})
        object.__setattr__(self, "_vars", {})
        object.__setattr__(self, "attrs", {})

    def createDimension(self, name, size):
        d = _Dim(name, size)
        self._dims[name] = d
        return d

    def createVariable(self, name, dtype, dims):
        v = _Var(self, name, dtype, dims)
        self._vars[name] = v
        return v

    def __setattr__(self, k, v):
        self.attrs[k] = v

    def sync(self):
        pass

    def close(self):
        out = {k: v._data for k, v in self._vars.items()}
        LAST[self._fnam] = out
        if os.environ.get("OGG_SHIM_NPZ") == "1":
            _np.savez(self._fnam + ".npz", **out)


def stringtochar(a):
    a = _np.asarray(a)
    n = a.dtype.itemsize
    return a.view("S1").reshape(a.shape + (n,))

"""CDF-2 ("NETCDF3_64BIT" = classic format, 64-bit offsets) writer for the supergrid file.

Reproduces the container the reference produces through netCDF4-python (ocean_grid_generator.py:773-829):
dimensions nyp, nxp, ny, nx, string(255); variables in creation order tile(S1; string), y, x (units "degrees"),
dy (ny,nxp; "meters"), dx (nyp,nx; "meters"), area (ny,nx; "m2"), angle_dx (nyp,nxp; "degrees"); optional global
attributes history / description / source.  Layout follows the published classic-format grammar (header, then
the fixed-size variables contiguous in definition order, 4-byte aligned, big-endian).  netCDF-C is absent from
this image, so byte-identity with a netCDF-C written file is by construction from the grammar, not by comparison.
"""
import struct

import numpy as np

NC_CHAR, NC_DOUBLE = 2, 6
NC_DIMENSION, NC_VARIABLE, NC_ATTRIBUTE = 0x0A, 0x0B, 0x0C


def _pad4(n):
    return (4 - n % 4) % 4


def _name(s):
    b = s.encode("ascii")
    return struct.pack(">i", len(b)) + b + b"\x00" * _pad4(len(b))


def _char_attr(name, text):
    b = text.encode("utf-8")
    return _name(name) + struct.pack(">ii", NC_CHAR, len(b)) + b + b"\x00" * _pad4(len(b))


def _att_list(attrs):
    if not attrs:
        return struct.pack(">ii", 0, 0)
    return struct.pack(">ii", NC_ATTRIBUTE, len(attrs)) + b"".join(_char_attr(k, v) for k, v in attrs)


def write_supergrid(fnam, x, y, dx, dy, area, angle_dx, global_attrs=None, format="NETCDF3_64BIT"):
    if format not in ("NETCDF3_64BIT", "NETCDF3_64BIT_OFFSET"):
        raise ValueError("only the reference's NETCDF3_64BIT container is written")
    ny, nx = area.shape
    nyp, nxp = ny + 1, nx + 1
    assert x.shape == (nyp, nxp) and y.shape == (nyp, nxp) and angle_dx.shape == (nyp, nxp)
    assert dx.shape == (nyp, nx) and dy.shape == (ny, nxp)
    dims = [("nyp", nyp), ("nxp", nxp), ("ny", ny), ("nx", nx), ("string", 255)]
    dim_id = {n: k for k, (n, _) in enumerate(dims)}
    tile = np.zeros(255, dtype="S1")
    tile[:5] = [c.encode() for c in "tile1"]
    variables = [  # name, dims, type, attrs, data
        ("tile", ("string",), NC_CHAR, [], tile),
        ("y", ("nyp", "nxp"), NC_DOUBLE, [("units", "degrees")], y),
        ("x", ("nyp", "nxp"), NC_DOUBLE, [("units", "degrees")], x),
        ("dy", ("ny", "nxp"), NC_DOUBLE, [("units", "meters")], dy),
        ("dx", ("nyp", "nx"), NC_DOUBLE, [("units", "meters")], dx),
        ("area", ("ny", "nx"), NC_DOUBLE, [("units", "m2")], area),
        ("angle_dx", ("nyp", "nxp"), NC_DOUBLE, [("units", "degrees")], angle_dx),
    ]

    def var_bytes(v):
        _, vdims, vtype, _, _ = v
        n = 1
        for d in vdims:
            n *= dict(dims)[d]
        raw = n * (1 if vtype == NC_CHAR else 8)
        return raw + _pad4(raw)

    def header(begins):
        h = b"CDF\x02" + struct.pack(">i", 0)
        h += struct.pack(">ii", NC_DIMENSION, len(dims)) + b"".join(_name(n) + struct.pack(">i", s) for n, s in dims)
        h += _att_list(global_attrs)
        h += struct.pack(">ii", NC_VARIABLE, len(variables))
        for v, begin in zip(variables, begins):
            vname, vdims, vtype, vatts, _ = v
            vsize = var_bytes(v)
            h += _name(vname) + struct.pack(">i", len(vdims)) + b"".join(struct.pack(">i", dim_id[d]) for d in vdims)
            h += _att_list(vatts) + struct.pack(">i", vtype)
            h += struct.pack(">I", vsize if vsize < 2 ** 32 else 2 ** 32 - 1) + struct.pack(">q", begin)
        return h

    hlen = len(header([0] * len(variables)))
    begins, off = [], hlen
    for v in variables:
        begins.append(off)
        off += var_bytes(v)
    with open(str(fnam), "wb") as f:
        f.write(header(begins))
        for v in variables:
            _, _, vtype, _, data = v
            if vtype == NC_CHAR:
                raw = data.tobytes()
                f.write(raw + b"\x00" * _pad4(len(raw)))
            elif hasattr(data, "bigendian_chunks"):
                # a producer of ready-made big-endian bytes (the device path: byte order reversed on the GPU, written
                # straight into pinned host staging buffers): nothing is converted or copied on the host
                for chunk in data.bigendian_chunks():
                    f.write(chunk)
            else:
                a = np.ascontiguousarray(data, dtype=np.float64)
                step = max(1, (1 << 24) // max(1, a.shape[1]))
                for r in range(0, a.shape[0], step):
                    f.write(a[r:r + step].astype(">f8").tobytes())

"""Repo-owned stand-in for the un-vendored `numpypi` package (TEST INFRASTRUCTURE ONLY).

The reference does `import numpypi.numpypi_series as np` (ocean_grid_generator.py:5).  Upstream
numpypi is absent from the image and unpinned, so this package supplies the portable-intrinsics
module the oracle runs the unmodified reference with.  See numpypi_series.py.
"""

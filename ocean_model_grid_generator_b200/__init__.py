"""B200-native supergrid construction hot path of ocean_model_grid_generator.

Layout:
  csrc/                 CUDA kernels (sm_100a, FP64, -fmad=false) + the C ABI of include/ogg_b200.h
  _cabi.py              ctypes binding of libogg_b200.so (fails loudly when the library or a GPU is missing)
  devnp.py              numpypi.numpypi_series-shaped module whose functions run on the device
  ocean_grid_generator  drop-in mirror of the reference module (same names, arguments, errors, CLI)
  supergrid.py          plan + in-place device build of the stitched supergrid, row-slab sharding
  netcdf3.py            CDF-2 ("NETCDF3_64BIT") writer with the reference's variable layout
"""
__version__ = "0.1.0"

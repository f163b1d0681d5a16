"""Host-side mirror of the reference's plugin surface for the hot path.

Module names follow the reference files they stand in for:
  volume_func / bitfield_func / multivol_func / trunc_exp   <- arcnerf/ops/*.py
  hashgrid_encoder / sh_encoder                             <- arcnerf/models/base_modules/encoding/
  tcnn_fusedmlp_module                                      <- arcnerf/models/base_modules/geo_rad_model/
  ray_helper                                                <- arcnerf/render/ray_helper.py
Every function keeps the reference's name, argument order and error behaviour and reaches the CUDA
kernels only through the C ABI (include/arcb200.h).
"""

"""d4est_b200: B200 (sm_100a) matrix-free SIPG/DG Laplacian path behind d4est's operator API.

Python here is only the host-side mirror of the C ABI (capi) used by tests and bench.py;
the product is libd4est_b200.so (CUDA) + the d4est-facing C glue (glue/).
"""
from .capi import MeshLevel, D4BError, lib, LIB_PATH  # noqa: F401

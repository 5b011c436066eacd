"""xash3d-fwgs_b200: B200-native `ref/cuda` renderer plugin for Xash3D FWGS + a headless engine host.

Python is only the bench / verification driver (ctypes over the C ABI); the product is
`libref_cuda.so` (C host code + hand-written sm_100a CUDA kernels) which exports the same
`GetRefAPI` / `ref_interface_t` surface as the reference's ref_soft (engine/ref_api.h:660).
"""
from .hostlib import Host, REF_CUDA_PATH, REF_SOFT_PATH, HOST_LIB_PATH, pixels_to_rgb  # noqa: F401

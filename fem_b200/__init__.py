"""fem_b200 -- B200-native implementation of the `FEM map` hot path (haowenz/FEM).

csrc/   CUDA kernels (sm_100a) + the C ABI (include/femgpu.h) -> lib/libfemgpu.so
host/   the `FEM` command line and libfemhost.so (plain C): readers, index, batch engine, SAM writer
api.py  ctypes mirror of the reference's call sites for tests and bench.py
"""
from .api import FemGpuError, Mapper, ReadBatch, MAPPING_DTYPE, load_library, sort_mappings  # noqa: F401

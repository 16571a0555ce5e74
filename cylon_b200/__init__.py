"""cylon_b200 -- B200 (sm_100a) kernels behind Cylon's distributed relational hot path.

The product is the CUDA library cylon_b200/lib/libcylon_b200.so (C-ABI: include/cylon_b200.h; C++ host API
mirroring cylon::Table / DistributedJoin / JoinConfig / GroupBy: include/cylon/). This Python package is a thin
ctypes binding over it for tests and benchmarks.
"""
__version__ = "0.1.0"

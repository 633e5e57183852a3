"""B200-native path-tracing core behind the ray-tracing-cuda scene API — Python host bindings.

The compute lives in csrc/ (hand-written CUDA for sm_100a behind the C-ABI of include/rtcore.h,
built into librtcore.so).  There is no CPU fallback: `lib()` raises if the library is missing.
"""
from . import abi, scenes  # noqa: F401
from .core import Context, RtError, lib, nccl_unique_id, write_image, write_pfm  # noqa: F401

"""b200ops - B200-native (sm_100a) decode hot path behind the vLLM operator surface.

Python here is host glue only (device memory, streams): every op is hand-written CUDA in
libb200ops.so, reached through the C ABI of include/b200ops.h. No CPU fallback exists.
"""
from . import _lib, ops, fused_moe  # noqa: F401

__all__ = ["_lib", "ops", "fused_moe"]

"""B200-native paged-decode attention hot path (drop-in for the reference's
``ops.paged_attention_v1/v2``, ``ops.reshape_and_cache[_flash]``,
``ops.convert_fp8`` and the ``AttentionBackend`` that calls them).

The compute path is hand-written CUDA for sm_100a behind the C ABI declared in
``include/b200_paged_attn.h``; there is no CPU or library fallback -- importing
the op modules fails loudly when the native libraries are missing.
"""
from . import _native  # noqa: F401
from ._native import load, native_lib_paths  # noqa: F401

__all__ = ["load", "native_lib_paths"]

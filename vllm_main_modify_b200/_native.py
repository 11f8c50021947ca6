"""Loader of the native libraries.  No fallback: a missing library is an error."""
from __future__ import annotations

import ctypes
import os
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libb200pa.so")
SHIM_PATH = os.path.join(_HERE, "_b200pa_torch.so")

# every symbol include/b200_paged_attn.h declares
C_ABI_SYMBOLS = (
    "b200pa_abi_version", "b200pa_last_error", "b200pa_kv_dtype_from_string",
    "b200pa_launch_count",
    "b200pa_paged_attention_workspace_bytes",
    "b200pa_paged_attention_v1", "b200pa_paged_attention_v1_builtin_scratch",
    "b200pa_paged_attention_v2", "b200pa_paged_attention_lse",
    "b200pa_decode_step", "b200pa_paged_attention_flash", "b200pa_prefill_attention",
    "b200pa_merge_attn_states", "b200pa_correct_attn_out",
    "b200pa_swap_blocks", "b200pa_copy_blocks", "b200pa_gather_and_maybe_dequant_cache",
    "b200pa_cp_gather_cache",
    "b200pa_reshape_and_cache", "b200pa_reshape_and_cache_flash",
    "b200pa_convert_fp8",
    "b200pa_get_max_shared_memory_per_block_device_attribute",
)

# test hooks (include/b200_paged_attn_testing.h): present when built with -DB200PA_TESTING=1
TESTING_SYMBOLS = ("b200pa_set_impl", "b200pa_get_impl", "b200pa_testing_set_stream_tune",
                   "b200pa_testing_set_trace")

IMPL_AUTO, IMPL_GENERIC, IMPL_STREAM = 0, 1, 2

_lock = threading.Lock()
_clib = None
_shim_loaded = False
_vllm_ns = False


def native_lib_paths():
    return LIB_PATH, SHIM_PATH


def c_lib() -> ctypes.CDLL:
    """dlopen libb200pa.so (the C ABI).  Works without a GPU; compute calls don't."""
    global _clib
    with _lock:
        if _clib is None:
            if not os.path.exists(LIB_PATH):
                raise ImportError(
                    f"{LIB_PATH} is missing: build it with "
                    "`python -m vllm_main_modify_b200.build` (there is no fallback path)")
            lib = ctypes.CDLL(LIB_PATH, mode=ctypes.RTLD_GLOBAL)
            lib.b200pa_last_error.restype = ctypes.c_char_p
            lib.b200pa_launch_count.restype = ctypes.c_uint64
            lib.b200pa_kv_dtype_from_string.argtypes = [ctypes.c_char_p]
            _clib = lib
        return _clib


def load(register_vllm_namespaces: bool = True):
    """Load the C ABI and the torch shim; returns ``torch.ops.b200pa``.

    With ``register_vllm_namespaces`` the ops are also defined as
    ``torch.ops._C.paged_attention_v1/v2``, ``torch.ops._C_cache_ops.*`` and
    ``torch.ops._C_cuda_utils.*`` unless another extension (the reference's own
    ``vllm._C``) already defined them in this process.
    """
    global _shim_loaded, _vllm_ns
    import torch
    c_lib()
    with _lock:
        if not _shim_loaded:
            if not os.path.exists(SHIM_PATH):
                raise ImportError(
                    f"{SHIM_PATH} is missing: build it with "
                    "`python -m vllm_main_modify_b200.build` (there is no fallback path)")
            torch.ops.load_library(SHIM_PATH)
            _shim_loaded = True
        if register_vllm_namespaces and not _vllm_ns:
            already = hasattr(torch.ops._C, "paged_attention_v1") if hasattr(torch.ops, "_C") else False
            if not already:
                shim = ctypes.CDLL(SHIM_PATH)
                shim.b200pa_register_vllm_namespaces()
            _vllm_ns = True
    return torch.ops.b200pa


def has_testing_hooks() -> bool:
    return hasattr(c_lib(), "b200pa_set_impl")


def set_impl(impl: int) -> None:
    """Test hook: force one kernel family (needs a -DB200PA_TESTING=1 build)."""
    c_lib().b200pa_set_impl(int(impl))


def get_impl() -> int:
    return int(c_lib().b200pa_get_impl()) if has_testing_hooks() else IMPL_AUTO


def set_stream_tune(tune: int) -> None:
    """Test hook: alternative launch shape of the stream kernel (0 = default)."""
    c_lib().b200pa_testing_set_stream_tune(int(tune))


def set_trace(device_ptr: int, nbytes: int) -> None:
    """Test hook: arm (or, with 0, disarm) the in-kernel phase trace of the stream kernel."""
    lib = c_lib()
    lib.b200pa_testing_set_trace.argtypes = [ctypes.c_void_p, ctypes.c_int64]
    lib.b200pa_testing_set_trace(ctypes.c_void_p(device_ptr or None), int(nbytes))


def launch_count() -> int:
    return int(c_lib().b200pa_launch_count())

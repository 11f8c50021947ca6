"""TEST INFRASTRUCTURE ONLY -- loader for the reference's own CUDA kernels (oracle tier O1).

``oracle/build_ref_cuda.sh`` compiles ``csrc/attention/paged_attention_v1.cu``,
``csrc/attention/paged_attention_v2.cu``, ``csrc/attention/merge_attn_states.cu`` and
``csrc/cache_kernels.cu`` of the reference,
unmodified, for sm_100a into ``oracle/_ref/libvllm_ref_cuda.so``; the ops register under the
private namespace ``vllm_ref_cuda`` (``oracle/ref_cuda_binding.cpp``).  Used by ``tests/`` as the
parity target north_star names ("the reference's own paged_attention implementation") and by
``tools/kbench.py --ref-cuda`` as the same-box kernel to beat.  The product never loads it.
"""
from __future__ import annotations

import os
from typing import Optional

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_ref", "libvllm_ref_cuda.so")
_loaded: Optional[str] = None


def available() -> bool:
    return os.path.exists(LIB_PATH) and torch.cuda.is_available()


def load() -> str:
    global _loaded
    if _loaded is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError("oracle/_ref/libvllm_ref_cuda.so is not built (run "
                               "oracle/build_ref_cuda.sh where /root/reference exists)")
        torch.ops.load_library(LIB_PATH)
        _loaded = LIB_PATH
    return _loaded


def _ops():
    load()
    return torch.ops.vllm_ref_cuda


def paged_attention_v1(*args) -> None:
    _ops().paged_attention_v1(*args)


def paged_attention_v2(*args) -> None:
    _ops().paged_attention_v2(*args)


def reshape_and_cache(*args) -> None:
    _ops().reshape_and_cache(*args)


def reshape_and_cache_flash(*args) -> None:
    _ops().reshape_and_cache_flash(*args)


def convert_fp8(*args) -> None:
    _ops().convert_fp8(*args)


def merge_attn_states(*args) -> None:
    _ops().merge_attn_states(*args)


def swap_blocks(*args) -> None:
    _ops().swap_blocks(*args)


def copy_blocks(*args) -> None:
    _ops().copy_blocks(*args)


def gather_and_maybe_dequant_cache(*args) -> None:
    _ops().gather_and_maybe_dequant_cache(*args)


def cp_gather_cache(*args) -> None:
    _ops().cp_gather_cache(*args)

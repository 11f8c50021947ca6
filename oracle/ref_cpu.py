"""TEST INFRASTRUCTURE ONLY -- loader for the reference's own CPU kernels.

``oracle/build_ref.sh`` compiles ``csrc/cpu/attention.cpp`` and
``csrc/cpu/cache.cpp`` of the reference, unmodified, into
``oracle/_ref/libvllm_ref_cpu_<isa>.so``; the ops register under the private
namespace ``vllm_ref_cpu`` (``oracle/ref_binding.cpp``).  Used as (a) the live
check of the C restatement, (b) the timed CPU baseline (``kind: "reference"``).
"""
from __future__ import annotations

import os
from typing import Optional

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_REF_DIR = os.path.join(_HERE, "_ref")
_loaded: Optional[str] = None


def _cpu_flags() -> set:
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("flags"):
                    return set(line.split(":", 1)[1].split())
    except OSError:
        pass
    return set()


def _pick_variant() -> Optional[str]:
    flags = _cpu_flags()
    order = []
    avx512 = {"avx512f", "avx512vl", "avx512bw", "avx512dq"} <= flags
    if avx512 and "avx512_bf16" in flags and "avx512_vnni" in flags:
        order.append("avx512bf16")
    if avx512:
        order.append("avx512")
    if "avx2" in flags and "fma" in flags and "f16c" in flags:
        order.append("avx2")
    for name in order:
        path = os.path.join(_REF_DIR, f"libvllm_ref_cpu_{name}.so")
        if os.path.exists(path):
            return path
    return None


def available() -> bool:
    return _pick_variant() is not None


def load() -> str:
    """Load the best ISA variant this host supports; returns its path."""
    global _loaded
    if _loaded is None:
        path = _pick_variant()
        if path is None:
            raise RuntimeError(
                "oracle/_ref is not built (run oracle/build_ref.sh where "
                "/root/reference exists)")
        torch.ops.load_library(path)
        _loaded = path
    return _loaded


def variant() -> str:
    return os.path.basename(load()).replace("libvllm_ref_cpu_", "").replace(".so", "")


def paged_attention_v1(*args) -> None:
    load()
    torch.ops.vllm_ref_cpu.paged_attention_v1(*args)


def paged_attention_v2(*args) -> None:
    load()
    torch.ops.vllm_ref_cpu.paged_attention_v2(*args)


def reshape_and_cache(*args) -> None:
    load()
    torch.ops.vllm_ref_cpu.reshape_and_cache(*args)

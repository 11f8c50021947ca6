"""TEST INFRASTRUCTURE ONLY -- CPU restatements of the block-management cache ops.

Nothing here is imported by the product.  Each function follows the reference kernel it names
(csrc/cache_kernels.cu) with plain indexing on CPU tensors; fp8 dequantisation goes through the C
oracle's ``convert_fp8`` (pinned on tests/golden/fp8_golden.npz).  Pinning of THESE functions: on
the GPU box they are compared bit for bit with the reference's own CUDA ops compiled unmodified
(oracle/_ref/libvllm_ref_cuda.so, tests/test_gpu_block_ops.py); the reference has no CPU
implementation and no stored vectors for them, so without a GPU they are checked through
round-trip properties only (tests/test_block_ops_oracle.py).
"""
from __future__ import annotations

from typing import List, Optional

import torch

from . import oracle as _o


def swap_blocks(src: torch.Tensor, dst: torch.Tensor, block_mapping: torch.Tensor) -> None:
    """cache_kernels.cu:27-70: dst block d <- src block s, whole block (stride(0) elements)."""
    for s, d in block_mapping.tolist():
        dst[d].copy_(src[s])


def copy_blocks(key_caches: List[torch.Tensor], value_caches: List[torch.Tensor],
                block_mapping: torch.Tensor) -> None:
    """cache_kernels.cu:74-101, 125-172: in every layer block d <- block s, keys then values.
    Pairs are applied in order here; the kernel runs them concurrently, so (like the reference's
    own test, tests/kernels/attention/test_cache.py:56-120) mappings must not chain."""
    for kc, vc in zip(key_caches, value_caches):
        for s, d in block_mapping.tolist():
            kc[d].copy_(kc[s])
            vc[d].copy_(vc[s])


def _dequant(rows: torch.Tensor, dtype: torch.dtype, kv_cache_dtype: str, scale: float):
    out = torch.empty(rows.shape, dtype=dtype)
    _o.convert_fp8(out, rows.contiguous(), scale, kv_cache_dtype)
    return out


def gather_and_maybe_dequant_cache(src_cache: torch.Tensor, dst: torch.Tensor,
                                   block_table: torch.Tensor, cu_seq_lens: torch.Tensor,
                                   batch_size: int, kv_cache_dtype: str, scale,
                                   seq_starts: Optional[torch.Tensor] = None) -> None:
    """cache_kernels.cu:628-712: token p of sequence b comes from block
    block_table[b, seq_starts[b] // BS + p // BS], slot p % BS."""
    BS = src_cache.shape[1]
    flat = src_cache.flatten(2, -1)
    sc = float(scale.item()) if isinstance(scale, torch.Tensor) else float(scale)
    for b in range(batch_size):
        lo, hi = int(cu_seq_lens[b]), int(cu_seq_lens[b + 1])
        off = int(seq_starts[b]) // BS if seq_starts is not None else 0      # :663-668
        for p in range(hi - lo):
            row = flat[int(block_table[b, off + p // BS]), p % BS]
            if kv_cache_dtype == "auto":
                dst[lo + p].copy_(row)                                         # :683
            else:
                dst[lo + p].copy_(_dequant(row, dst.dtype, kv_cache_dtype, sc))  # :685-686


def cp_gather_cache(src_cache: torch.Tensor, dst: torch.Tensor, block_table: torch.Tensor,
                    cu_seq_lens: torch.Tensor, batch_size: int,
                    seq_starts: Optional[torch.Tensor] = None) -> None:
    """cache_kernels.cu:785-849: like the above without conversion; seq_starts counts tokens."""
    BS = src_cache.shape[1]
    flat = src_cache.flatten(2, -1)
    for b in range(batch_size):
        lo, hi = int(cu_seq_lens[b]), int(cu_seq_lens[b + 1])
        start = int(seq_starts[b]) if seq_starts is not None else 0           # :816-823
        for p in range(hi - lo):
            t = start + p
            dst[lo + p].copy_(flat[int(block_table[b, t // BS]), t % BS])


def xpacked_to_flash(key_cache: torch.Tensor, value_cache: torch.Tensor):
    """x-packed caches ``key_cache[NB,Hkv,D/x,BS,x]``, ``value_cache[NB,Hkv,D,BS]``
    (what ``reshape_and_cache`` writes, ``csrc/cache_kernels.cu:212-264`` of the reference) ->
    the token-major NHD caches ``[NB,BS,Hkv,D]`` that ``reshape_and_cache_flash`` writes
    (``:280-353``).  Pure re-indexing, no arithmetic: element (block, head, d, token) of one
    layout is element (block, token, head, d) of the other."""
    NB, Hkv, Dx, BS, x = key_cache.shape
    kf = key_cache.permute(0, 3, 1, 2, 4).reshape(NB, BS, Hkv, Dx * x).contiguous()
    vf = value_cache.permute(0, 3, 1, 2).contiguous()
    return kf, vf


def flash_to_xpacked(key_cache: torch.Tensor, value_cache: torch.Tensor):
    """Inverse of :func:`xpacked_to_flash` for ``[NB,BS,Hkv,D]`` caches (any strides)."""
    NB, BS, Hkv, D = key_cache.shape
    x = 16 // key_cache.element_size()
    kx = key_cache.reshape(NB, BS, Hkv, D // x, x).permute(0, 2, 3, 1, 4).contiguous()
    vx = value_cache.permute(0, 2, 3, 1).contiguous()
    return kx, vx

"""TEST INFRASTRUCTURE ONLY -- ctypes front-end of ``paged_attn_oracle.c``.

Every function mirrors the signature of the reference op it checks
(``vllm/_custom_ops.py:39-97, 1590-1618, 1650-1654`` of the reference) but takes
CPU torch tensors.  Nothing here is imported by the product.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from typing import Optional

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libpaged_attn_oracle.so")
_SRC_PATH = os.path.join(_HERE, "paged_attn_oracle.c")

F32, F16, BF16 = 0, 1, 2
KV_AUTO, KV_E4M3, KV_E5M2 = 0, 1, 2

_SCALAR = {torch.float32: F32, torch.float16: F16, torch.bfloat16: BF16}


def kv_kind(kv_cache_dtype: str) -> int:
    if kv_cache_dtype == "auto":
        return KV_AUTO
    if kv_cache_dtype in ("fp8", "fp8_e4m3"):
        return KV_E4M3
    if kv_cache_dtype == "fp8_e5m2":
        return KV_E5M2
    raise RuntimeError(f"Unsupported data type of kv cache: {kv_cache_dtype}")


def build(force: bool = False) -> str:
    """Compile the C oracle (system gcc; the image's $CC has no libgomp)."""
    if (not force and os.path.exists(_LIB_PATH)
            and os.path.getmtime(_LIB_PATH) >= os.path.getmtime(_SRC_PATH)):
        return _LIB_PATH
    subprocess.check_call([
        "/usr/bin/gcc", "-O2", "-fopenmp", "-shared", "-fPIC",
        "-ffp-contract=off", "-o", _LIB_PATH, _SRC_PATH, "-lm"
    ])
    return _LIB_PATH


_lib = None


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(build())
        _lib.ora_f32_to_fp8.restype = ctypes.c_uint8
        _lib.ora_f32_to_fp8.argtypes = [ctypes.c_float, ctypes.c_int]
        _lib.ora_fp8_to_f32.restype = ctypes.c_float
        _lib.ora_fp8_to_f32.argtypes = [ctypes.c_uint8, ctypes.c_int]
        _lib.ora_f32_to_bf16.restype = ctypes.c_uint16
        _lib.ora_f32_to_bf16.argtypes = [ctypes.c_float]
        _lib.ora_f32_to_f16.restype = ctypes.c_uint16
        _lib.ora_f32_to_f16.argtypes = [ctypes.c_float]
        _lib.ora_f16_to_f32.restype = ctypes.c_float
        _lib.ora_f16_to_f32.argtypes = [ctypes.c_uint16]
        _lib.ora_num_threads.restype = ctypes.c_int
    return _lib


def _p(t: Optional[torch.Tensor]):
    return ctypes.c_void_p(0 if t is None else t.data_ptr())


def _cpu(t: torch.Tensor) -> torch.Tensor:
    assert t.device.type == "cpu", "the oracle runs on CPU tensors"
    return t


def _scale(s) -> float:
    return float(s.item()) if isinstance(s, torch.Tensor) else float(s)


def reshape_and_cache(key, value, key_cache, value_cache, slot_mapping,
                      kv_cache_dtype: str, k_scale, v_scale) -> None:
    """cache_kernels.cu:416-442 of the reference."""
    _cpu(key)
    assert key_cache.is_contiguous() and value_cache.is_contiguous()
    num_tokens = slot_mapping.shape[0]
    num_heads, head_size = key.shape[1], key.shape[2]
    block_size, x = key_cache.shape[3], key_cache.shape[4]
    rc = lib().ora_reshape_and_cache(
        _p(key), _p(value), _p(key_cache), _p(value_cache),
        _p(slot_mapping.contiguous()), ctypes.c_int(num_tokens),
        ctypes.c_int64(key.stride(0)), ctypes.c_int64(value.stride(0)),
        ctypes.c_int(num_heads), ctypes.c_int(head_size),
        ctypes.c_int(block_size), ctypes.c_int(x),
        ctypes.c_int(_SCALAR[key.dtype]), ctypes.c_int(kv_kind(kv_cache_dtype)),
        ctypes.c_float(_scale(k_scale)), ctypes.c_float(_scale(v_scale)))
    assert rc == 0


def reshape_and_cache_flash(key, value, key_cache, value_cache, slot_mapping,
                            kv_cache_dtype: str, k_scale, v_scale) -> None:
    """cache_kernels.cu:459-497 of the reference (NHD or HND by strides)."""
    _cpu(key)
    assert key_cache.stride(0) == value_cache.stride(0)
    num_tokens = slot_mapping.shape[0]
    num_heads, head_size = key.shape[1], key.shape[2]
    block_size = key_cache.shape[1]
    rc = lib().ora_reshape_and_cache_flash(
        _p(key), _p(value), _p(key_cache), _p(value_cache),
        _p(slot_mapping.contiguous()), ctypes.c_int(num_tokens),
        ctypes.c_int64(key.stride(0)), ctypes.c_int64(value.stride(0)),
        ctypes.c_int64(key_cache.stride(0)), ctypes.c_int64(key_cache.stride(1)),
        ctypes.c_int64(key_cache.stride(2)), ctypes.c_int(num_heads),
        ctypes.c_int(head_size), ctypes.c_int(block_size),
        ctypes.c_int(_SCALAR[key.dtype]), ctypes.c_int(kv_kind(kv_cache_dtype)),
        ctypes.c_float(_scale(k_scale)), ctypes.c_float(_scale(v_scale)))
    assert rc == 0


def convert_fp8(output, input, scale: float = 1.0, kv_dtype: str = "fp8") -> None:
    """cache_kernels.cu:573-623 of the reference."""
    assert output.is_contiguous() and input.is_contiguous()
    to_fp8 = output.dtype == torch.uint8
    st = _SCALAR[input.dtype if to_fp8 else output.dtype]
    rc = lib().ora_convert_fp8(_p(output), _p(input), ctypes.c_int64(input.numel()),
                               ctypes.c_float(scale), ctypes.c_int(st),
                               ctypes.c_int(kv_kind(kv_dtype)),
                               ctypes.c_int(1 if to_fp8 else 0))
    assert rc == 0


def _attention(out, exp_sums, max_logits, tmp_out, query, key_cache,
               value_cache, num_kv_heads, scale, block_tables, seq_lens,
               block_size, max_seq_len, alibi_slopes, kv_cache_dtype, k_scale,
               v_scale, tp_rank, bs_local, bs_vert, bs_bsize, bs_step,
               partition_size, faithful, out_f32):
    _cpu(query)
    num_seqs, num_heads, head_size = query.shape
    assert query.stride(1) == head_size and query.stride(2) == 1
    x = key_cache.shape[4]
    assert block_size == value_cache.shape[3]
    block_tables = block_tables.to(torch.int32).contiguous()
    seq_lens = seq_lens.to(torch.int32).contiguous()
    if alibi_slopes is not None:
        alibi_slopes = alibi_slopes.to(torch.float32).contiguous()
    max_parts = 0 if exp_sums is None else exp_sums.shape[-1]
    rc = lib().ora_paged_attention(
        _p(out), _p(out_f32), _p(query), ctypes.c_int64(query.stride(0)),
        _p(key_cache), _p(value_cache), ctypes.c_int64(key_cache.stride(0)),
        ctypes.c_int64(key_cache.stride(1)), ctypes.c_int(num_seqs),
        ctypes.c_int(num_heads), ctypes.c_int(num_kv_heads),
        ctypes.c_int(head_size), ctypes.c_int(block_size), ctypes.c_int(x),
        ctypes.c_float(scale), _p(block_tables),
        ctypes.c_int(block_tables.shape[1]), _p(seq_lens), _p(alibi_slopes),
        ctypes.c_int(_SCALAR[query.dtype]), ctypes.c_int(kv_kind(kv_cache_dtype)),
        ctypes.c_float(_scale(k_scale)), ctypes.c_float(_scale(v_scale)),
        ctypes.c_int(partition_size), ctypes.c_int(1 if faithful else 0),
        _p(exp_sums), _p(max_logits), _p(tmp_out), ctypes.c_int(max_parts),
        ctypes.c_int(tp_rank), ctypes.c_int(bs_local), ctypes.c_int(bs_vert),
        ctypes.c_int(bs_bsize), ctypes.c_int(bs_step))
    assert rc == 0


def paged_attention_v1(out, query, key_cache, value_cache, num_kv_heads, scale,
                       block_tables, seq_lens, block_size, max_seq_len,
                       alibi_slopes, kv_cache_dtype, k_scale, v_scale,
                       tp_rank=0, blocksparse_local_blocks=0,
                       blocksparse_vert_stride=0, blocksparse_block_size=64,
                       blocksparse_head_sliding_step=0, *, faithful=True,
                       out_f32=None) -> None:
    """paged_attention_v1.cu:160-182 + attention_kernels.cuh:85-522."""
    _attention(out, None, None, None, query, key_cache, value_cache,
               num_kv_heads, scale, block_tables, seq_lens, block_size,
               max_seq_len, alibi_slopes, kv_cache_dtype, k_scale, v_scale,
               tp_rank, blocksparse_local_blocks, blocksparse_vert_stride,
               blocksparse_block_size, blocksparse_head_sliding_step, 0,
               faithful, out_f32)


def paged_attention_v2(out, exp_sums, max_logits, tmp_out, query, key_cache,
                       value_cache, num_kv_heads, scale, block_tables, seq_lens,
                       block_size, max_seq_len, alibi_slopes, kv_cache_dtype,
                       k_scale, v_scale, tp_rank=0, blocksparse_local_blocks=0,
                       blocksparse_vert_stride=0, blocksparse_block_size=64,
                       blocksparse_head_sliding_step=0, *, faithful=True,
                       out_f32=None) -> None:
    """paged_attention_v2.cu:167-192 + attention_kernels.cuh:525-664."""
    _attention(out, exp_sums, max_logits, tmp_out, query, key_cache,
               value_cache, num_kv_heads, scale, block_tables, seq_lens,
               block_size, max_seq_len, alibi_slopes, kv_cache_dtype, k_scale,
               v_scale, tp_rank, blocksparse_local_blocks,
               blocksparse_vert_stride, blocksparse_block_size,
               blocksparse_head_sliding_step, 512, faithful, out_f32)


def num_threads() -> int:
    return int(lib().ora_num_threads())

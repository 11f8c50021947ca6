"""Python op shim -- same names, argument order, defaults and return values as
the five paged-decode entries of the reference's ``vllm/_custom_ops.py``
(:39-65 paged_attention_v1, :68-97 paged_attention_v2, :1590-1602
reshape_and_cache, :1605-1618 reshape_and_cache_flash, :1650-1654 convert_fp8).

They dispatch to ``torch.ops.b200pa.*`` (the TORCH_LIBRARY shim over the C ABI);
the same kernels are also reachable as ``torch.ops._C.*`` / ``_C_cache_ops.*`` so
that the reference's own ``_custom_ops`` works unmodified on top of this library.
"""
from typing import Optional

import torch

from ._native import load

_ops = load()


def paged_attention_v1(
    out: torch.Tensor,
    query: torch.Tensor,
    key_cache: torch.Tensor,
    value_cache: torch.Tensor,
    num_kv_heads: int,
    scale: float,
    block_tables: torch.Tensor,
    seq_lens: torch.Tensor,
    block_size: int,
    max_seq_len: int,
    alibi_slopes: Optional[torch.Tensor],
    kv_cache_dtype: str,
    k_scale: torch.Tensor,
    v_scale: torch.Tensor,
    tp_rank: int = 0,
    blocksparse_local_blocks: int = 0,
    blocksparse_vert_stride: int = 0,
    blocksparse_block_size: int = 64,
    blocksparse_head_sliding_step: int = 0,
) -> None:
    _ops.paged_attention_v1(
        out, query, key_cache, value_cache, num_kv_heads, scale, block_tables,
        seq_lens, block_size, max_seq_len, alibi_slopes, kv_cache_dtype,
        k_scale, v_scale, tp_rank, blocksparse_local_blocks,
        blocksparse_vert_stride, blocksparse_block_size,
        blocksparse_head_sliding_step)


def paged_attention_v2(
    out: torch.Tensor,
    exp_sum: torch.Tensor,
    max_logits: torch.Tensor,
    tmp_out: torch.Tensor,
    query: torch.Tensor,
    key_cache: torch.Tensor,
    value_cache: torch.Tensor,
    num_kv_heads: int,
    scale: float,
    block_tables: torch.Tensor,
    seq_lens: torch.Tensor,
    block_size: int,
    max_seq_len: int,
    alibi_slopes: Optional[torch.Tensor],
    kv_cache_dtype: str,
    k_scale: torch.Tensor,
    v_scale: torch.Tensor,
    tp_rank: int = 0,
    blocksparse_local_blocks: int = 0,
    blocksparse_vert_stride: int = 0,
    blocksparse_block_size: int = 64,
    blocksparse_head_sliding_step: int = 0,
) -> None:
    _ops.paged_attention_v2(
        out, exp_sum, max_logits, tmp_out, query, key_cache, value_cache,
        num_kv_heads, scale, block_tables, seq_lens, block_size, max_seq_len,
        alibi_slopes, kv_cache_dtype, k_scale, v_scale, tp_rank,
        blocksparse_local_blocks, blocksparse_vert_stride,
        blocksparse_block_size, blocksparse_head_sliding_step)


def paged_attention_v2_partials(out, exp_sum, max_logits, tmp_out, query, key_cache, value_cache,
                                num_kv_heads, scale, block_tables, seq_lens, block_size,
                                max_seq_len, alibi_slopes, kv_cache_dtype, k_scale, v_scale,
                                tp_rank: int = 0, blocksparse_local_blocks: int = 0,
                                blocksparse_vert_stride: int = 0, blocksparse_block_size: int = 64,
                                blocksparse_head_sliding_step: int = 0) -> None:
    """``paged_attention_v2`` for a caller that READS the scratch: on return ``exp_sum`` /
    ``max_logits`` / ``tmp_out`` hold the reference's per-partition partials
    (``attention_kernels.cuh:345-351, 487``).  The plain op leaves them untouched on its fast path;
    this per-call form (no reference counterpart) runs the slower generic kernel that fills them.
    The environment variable ``B200PA_V2_REFERENCE_SCRATCH=1`` does the same process-wide for
    callers that cannot be changed."""
    _ops.paged_attention_v2_partials(
        out, exp_sum, max_logits, tmp_out, query, key_cache, value_cache,
        num_kv_heads, scale, block_tables, seq_lens, block_size, max_seq_len,
        alibi_slopes, kv_cache_dtype, k_scale, v_scale, tp_rank,
        blocksparse_local_blocks, blocksparse_vert_stride,
        blocksparse_block_size, blocksparse_head_sliding_step)


def merge_attn_states(output: torch.Tensor,
                      prefix_output: torch.Tensor,
                      prefix_lse: torch.Tensor,
                      suffix_output: torch.Tensor,
                      suffix_lse: torch.Tensor,
                      output_lse: Optional[torch.Tensor] = None) -> None:
    """``vllm/_custom_ops.py:143-150``: output [n,h,d], lse tensors [h,n] fp32."""
    _ops.merge_attn_states(output, output_lse, prefix_output, prefix_lse,
                           suffix_output, suffix_lse)


def paged_attention_lse(
    out: torch.Tensor,
    lse: torch.Tensor,
    query: torch.Tensor,
    key_cache: torch.Tensor,
    value_cache: torch.Tensor,
    num_kv_heads: int,
    scale: float,
    block_tables: torch.Tensor,
    seq_lens: torch.Tensor,
    block_size: int,
    max_seq_len: int,
    alibi_slopes: Optional[torch.Tensor],
    kv_cache_dtype: str,
    k_scale: torch.Tensor,
    v_scale: torch.Tensor,
) -> None:
    """``paged_attention_v1`` that also writes ``lse[num_seqs, num_heads]`` (fp32, any strides):
    the natural-log sum-exp of the scaled logits (``-inf`` for an empty sequence).  No twin in
    the reference's ``_custom_ops``; it is what ``can_return_lse_for_decode``
    (``abstract.py:115-121``) needs from a decode backend."""
    _ops.paged_attention_lse(
        out, lse, query, key_cache, value_cache, num_kv_heads, scale,
        block_tables, seq_lens, block_size, max_seq_len, alibi_slopes,
        kv_cache_dtype, k_scale, v_scale)


def decode_step(
    out: torch.Tensor,
    lse: Optional[torch.Tensor],
    query: torch.Tensor,
    key: torch.Tensor,
    value: torch.Tensor,
    key_cache: torch.Tensor,
    value_cache: torch.Tensor,
    slot_mapping: torch.Tensor,
    num_kv_heads: int,
    scale: float,
    block_tables: torch.Tensor,
    seq_lens: torch.Tensor,
    block_size: int,
    max_seq_len: int,
    alibi_slopes: Optional[torch.Tensor],
    kv_cache_dtype: str,
    k_scale: torch.Tensor,
    v_scale: torch.Tensor,
) -> None:
    """``reshape_and_cache`` of one new token per sequence + ``paged_attention_v1`` (+ ``lse``
    when given) in ONE launch: what ``PagedAttention.write_to_paged_cache`` followed by
    ``forward_decode`` (``paged_attn.py:73-198``) do for a pure decode batch.  Token ``b`` must be
    position ``seq_lens[b] - 1`` of sequence ``b`` (``slot_mapping[b]`` its slot).  Same cache
    bytes and same output as the two ops; no twin in the reference's ``_custom_ops``."""
    _ops.decode_step(
        out, lse, query, key, value, key_cache, value_cache, slot_mapping, num_kv_heads, scale,
        block_tables, seq_lens, block_size, max_seq_len, alibi_slopes, kv_cache_dtype,
        k_scale, v_scale)


def paged_attention_flash(
    out: torch.Tensor,
    lse: Optional[torch.Tensor],
    query: torch.Tensor,
    key: Optional[torch.Tensor],
    value: Optional[torch.Tensor],
    key_cache: torch.Tensor,
    value_cache: torch.Tensor,
    slot_mapping: Optional[torch.Tensor],
    scale: float,
    block_tables: torch.Tensor,
    seq_lens: torch.Tensor,
    max_seq_len: int,
    alibi_slopes: Optional[torch.Tensor],
    kv_cache_dtype: str,
    k_scale: torch.Tensor,
    v_scale: torch.Tensor,
) -> None:
    """Paged decode attention over the token-major cache ``reshape_and_cache_flash`` writes:
    ``key_cache`` / ``value_cache`` ``[num_blocks, block_size, num_kv_heads, head_size]``, NHD or
    HND by strides (``cache_kernels.cu:280-353`` of the reference).  Same math as
    ``paged_attention_v1`` (+ ``lse`` when given).  With ``key`` / ``value`` / ``slot_mapping`` the
    new token of every sequence is quantised and stored in the same launch (the effect of
    ``reshape_and_cache_flash`` before the attention).  No twin in the reference's
    ``_custom_ops``: its paged_attention ops only read the x-packed layout (SURVEY note 1)."""
    _ops.paged_attention_flash(
        out, lse, query, key, value, key_cache, value_cache, slot_mapping, scale,
        block_tables, seq_lens, max_seq_len, alibi_slopes, kv_cache_dtype, k_scale, v_scale)


def prefill_attention(
    out: torch.Tensor,
    lse: Optional[torch.Tensor],
    query: torch.Tensor,
    key_cache: torch.Tensor,
    value_cache: torch.Tensor,
    scale: float,
    block_tables: torch.Tensor,
    query_start_loc: torch.Tensor,
    seq_lens: torch.Tensor,
    max_query_len: int,
    alibi_slopes: Optional[torch.Tensor],
    kv_cache_dtype: str,
    k_scale: torch.Tensor,
    v_scale: torch.Tensor,
    heads_per_tile: int = 0,
) -> None:
    """Causal attention of query chunks (prefill / chunked prefill / multi-token decode) over the
    paged cache, whose K/V already hold the chunk -- what the reference's paged backends call
    ``flash_attn_varlen_func(..., block_table)`` for (``cpu_attn.py:584-612``).  ``query`` /
    ``out`` ``[num_tokens, num_heads, head_size]``, ``query_start_loc`` ``[num_seqs + 1]`` int32,
    ``seq_lens`` ``[num_seqs]`` int32 (context + new tokens), ``lse`` optional
    ``[num_tokens, num_heads]`` fp32.

    Limits: at most 65535 sequences per call.  With an fp8 cache the kernel computes in f16, so a
    bf16 query must stay inside the f16 range (|q| <= 65504; larger values give inf / NaN, tiny
    ones flush to zero) -- the reference dequantises to the query dtype instead.  A sequence whose
    ``seq_lens`` entry is smaller than its number of query tokens (inconsistent metadata) gets
    zero rows and ``lse = -inf``."""
    _ops.prefill_attention(out, lse, query, key_cache, value_cache, scale, block_tables,
                           query_start_loc, seq_lens, max_query_len, alibi_slopes,
                           kv_cache_dtype, k_scale, v_scale, heads_per_tile)


def correct_attn_out(out: torch.Tensor, lses: torch.Tensor, cp_rank: int,
                     ctx=None) -> "tuple[torch.Tensor, torch.Tensor]":
    """``vllm/attention/ops/common.py:86-116`` (``ctx`` is the reference's Triton cache, unused):
    rescales ``out [B,H,D]`` in place by ``exp(lses[cp_rank] - logsumexp_n(lses))`` and returns
    ``(out, lse_total)``."""
    lse = _ops.correct_attn_out(out, lses, cp_rank)
    return out, lse


def reshape_and_cache(
    key: torch.Tensor,
    value: torch.Tensor,
    key_cache: torch.Tensor,
    value_cache: torch.Tensor,
    slot_mapping: torch.Tensor,
    kv_cache_dtype: str,
    k_scale: torch.Tensor,
    v_scale: torch.Tensor,
) -> None:
    _ops.reshape_and_cache(key, value, key_cache, value_cache, slot_mapping,
                           kv_cache_dtype, k_scale, v_scale)


def reshape_and_cache_flash(
    key: torch.Tensor,
    value: torch.Tensor,
    key_cache: torch.Tensor,
    value_cache: torch.Tensor,
    slot_mapping: torch.Tensor,
    kv_cache_dtype: str,
    k_scale: torch.Tensor,
    v_scale: torch.Tensor,
) -> None:
    _ops.reshape_and_cache_flash(key, value, key_cache, value_cache,
                                 slot_mapping, kv_cache_dtype, k_scale, v_scale)


def convert_fp8(output: torch.Tensor,
                input: torch.Tensor,
                scale: float = 1.0,
                kv_dtype: str = "fp8") -> None:
    _ops.convert_fp8(output, input, scale, kv_dtype)


def get_max_shared_memory_per_block_device_attribute(device: int) -> int:
    """vllm/_custom_ops.py ``get_max_shared_memory_per_block_device_attribute``."""
    return _ops.get_max_shared_memory_per_block_device_attribute(device)


def copy_blocks(key_caches: "list[torch.Tensor]", value_caches: "list[torch.Tensor]",
                block_mapping: torch.Tensor) -> None:
    """``vllm/_custom_ops.py:1634-1637``; ``block_mapping`` [num_pairs, 2] int64 on the GPU."""
    _ops.copy_blocks(key_caches, value_caches, block_mapping)


def swap_blocks(src: torch.Tensor, dst: torch.Tensor, block_mapping: torch.Tensor) -> None:
    """``vllm/_custom_ops.py:1645-1647``; ``block_mapping`` [num_pairs, 2] int64 on the CPU;
    src / dst: GPU<->GPU (same device), GPU->CPU or CPU->GPU."""
    _ops.swap_blocks(src, dst, block_mapping)


def gather_and_maybe_dequant_cache(src_cache: torch.Tensor,
                                   dst: torch.Tensor,
                                   block_table: torch.Tensor,
                                   cu_seq_lens: torch.Tensor,
                                   batch_size: int,
                                   kv_cache_dtype: str,
                                   scale: torch.Tensor,
                                   seq_starts: Optional[torch.Tensor] = None) -> None:
    """``vllm/_custom_ops.py:1657-1668``."""
    _ops.gather_and_maybe_dequant_cache(src_cache, dst, block_table, cu_seq_lens, batch_size,
                                        kv_cache_dtype, scale, seq_starts)


def cp_gather_cache(src_cache: torch.Tensor,
                    dst: torch.Tensor,
                    block_table: torch.Tensor,
                    cu_seq_lens: torch.Tensor,
                    batch_size: int,
                    seq_starts: Optional[torch.Tensor] = None) -> None:
    """``vllm/_custom_ops.py:1671-1678``."""
    _ops.cp_gather_cache(src_cache, dst, block_table, cu_seq_lens, batch_size, seq_starts)

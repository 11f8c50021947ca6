"""Paged-attention helper: same class, static methods, argument meaning and
v1/v2 heuristic as the reference's ``vllm/attention/ops/paged_attn.py:24-198``,
on top of this package's ops."""
from dataclasses import dataclass
from typing import List, Optional, Tuple

import torch

from . import _custom_ops as ops

# Same as PARTITION_SIZE of the split-K kernel (paged_attn.py:21 of the reference).
_PARTITION_SIZE = 512
# head sizes the one-launch decode kernels serve (csrc/paged_attention_stream.cu, stream_supported)
_STREAM_HEAD_SIZES = (32, 64, 80, 96, 112, 128, 192, 256)


@dataclass
class PagedAttentionMetadata:
    """Metadata for PagedAttention (paged_attn.py:24-38 of the reference)."""
    # (batch_size,) total tokens seen so far per sequence.
    seq_lens_tensor: Optional[torch.Tensor]
    # Maximum sequence length in the batch. 0 if it is a prefill-only batch.
    max_decode_seq_len: int
    # (batch_size, max_blocks_per_seq) physical block numbers per sequence.
    block_tables: Optional[torch.Tensor]


class PagedAttention:

    @staticmethod
    def get_supported_head_sizes() -> List[int]:
        return [32, 64, 80, 96, 112, 120, 128, 192, 256]

    @staticmethod
    def get_kv_cache_shape(num_blocks: int, block_size: int, num_kv_heads: int,
                           head_size: int) -> Tuple[int, ...]:
        return (2, num_blocks, block_size * num_kv_heads * head_size)

    @staticmethod
    def split_kv_cache(kv_cache: torch.Tensor, num_kv_heads: int,
                       head_size: int) -> Tuple[torch.Tensor, torch.Tensor]:
        x = 16 // kv_cache.element_size()
        num_blocks = kv_cache.shape[1]
        key_cache = kv_cache[0].view(num_blocks, num_kv_heads, head_size // x, -1, x)
        value_cache = kv_cache[1].view(num_blocks, num_kv_heads, head_size, -1)
        return key_cache, value_cache

    @staticmethod
    def write_to_paged_cache(key: torch.Tensor, value: torch.Tensor,
                             key_cache: torch.Tensor, value_cache: torch.Tensor,
                             slot_mapping: torch.Tensor, kv_cache_dtype: str,
                             k_scale: torch.Tensor, v_scale: torch.Tensor) -> None:
        ops.reshape_and_cache(key, value, key_cache, value_cache,
                              slot_mapping.flatten(), kv_cache_dtype, k_scale, v_scale)

    @staticmethod
    def use_v1(max_seq_len: int, num_seqs: int, num_heads: int) -> bool:
        """The reference heuristic (paged_attn.py:134-135)."""
        max_num_partitions = (max_seq_len + _PARTITION_SIZE - 1) // _PARTITION_SIZE
        return (max_seq_len <= 8192
                and (max_num_partitions == 1 or num_seqs * num_heads > 512))

    @staticmethod
    def forward_decode(
        query: torch.Tensor,
        key_cache: torch.Tensor,
        value_cache: torch.Tensor,
        block_tables: torch.Tensor,
        seq_lens: torch.Tensor,
        max_seq_len: int,
        kv_cache_dtype: str,
        num_kv_heads: int,
        scale: float,
        alibi_slopes: Optional[torch.Tensor],
        k_scale: torch.Tensor,
        v_scale: torch.Tensor,
        tp_rank: int = 0,
        blocksparse_local_blocks: int = 0,
        blocksparse_vert_stride: int = 0,
        blocksparse_block_size: int = 64,
        blocksparse_head_sliding_step: int = 0,
        output: Optional[torch.Tensor] = None,
    ) -> torch.Tensor:
        if blocksparse_vert_stride is not None and blocksparse_vert_stride > 1:
            block_size = value_cache.size(-1)
            assert (blocksparse_block_size > 0 and
                    blocksparse_block_size % block_size == 0), \
                (f"{blocksparse_block_size=} needs to be a multiple of"
                 f"{block_size=} used in block_tables.")

        if output is None:
            output = torch.empty_like(query)
        block_size = value_cache.shape[3]
        num_seqs, num_heads, head_size = query.shape
        # The decode kernels split long contexts by their own schedule inside ONE launch and keep
        # their partials in the library's workspace, so where they serve the shape there is no
        # v1 / v2 choice to make and no per-call scratch to allocate (the reference's
        # [num_seqs, num_heads, partitions, head_size] tmp_output is 128 MiB for 128 x 64 heads at
        # 32K context).  Only the shapes left to the SIMT kernel (head size 120, 8-token blocks,
        # fp32) keep the reference's heuristic and its scratch.
        one_launch = (block_size in (16, 32) and head_size in _STREAM_HEAD_SIZES
                      and query.dtype in (torch.float16, torch.bfloat16))
        if one_launch or PagedAttention.use_v1(max_seq_len, num_seqs, num_heads):
            ops.paged_attention_v1(
                output, query, key_cache, value_cache, num_kv_heads, scale,
                block_tables, seq_lens, block_size, max_seq_len, alibi_slopes,
                kv_cache_dtype, k_scale, v_scale, tp_rank,
                blocksparse_local_blocks, blocksparse_vert_stride,
                blocksparse_block_size, blocksparse_head_sliding_step)
        else:
            assert _PARTITION_SIZE % block_size == 0
            max_num_partitions = (max_seq_len + _PARTITION_SIZE - 1) // _PARTITION_SIZE
            tmp_output = torch.empty(
                size=(num_seqs, num_heads, max_num_partitions, head_size),
                dtype=output.dtype, device=output.device)
            exp_sums = torch.empty(
                size=(num_seqs, num_heads, max_num_partitions),
                dtype=torch.float32, device=output.device)
            max_logits = torch.empty_like(exp_sums)
            ops.paged_attention_v2(
                output, exp_sums, max_logits, tmp_output, query, key_cache,
                value_cache, num_kv_heads, scale, block_tables, seq_lens,
                block_size, max_seq_len, alibi_slopes, kv_cache_dtype, k_scale,
                v_scale, tp_rank, blocksparse_local_blocks,
                blocksparse_vert_stride, blocksparse_block_size,
                blocksparse_head_sliding_step)
        return output

    @staticmethod
    def write_and_decode(
        query: torch.Tensor,
        key: torch.Tensor,
        value: torch.Tensor,
        key_cache: torch.Tensor,
        value_cache: torch.Tensor,
        slot_mapping: torch.Tensor,
        block_tables: torch.Tensor,
        seq_lens: torch.Tensor,
        max_seq_len: int,
        kv_cache_dtype: str,
        num_kv_heads: int,
        scale: float,
        alibi_slopes: Optional[torch.Tensor],
        k_scale: torch.Tensor,
        v_scale: torch.Tensor,
        output: Optional[torch.Tensor] = None,
        lse: Optional[torch.Tensor] = None,
    ) -> torch.Tensor:
        """``write_to_paged_cache`` + ``forward_decode`` (:73-92, :94-198 of the reference's
        ``paged_attn.py``) of a pure decode batch (one new token per sequence, the sequence's
        last) in one launch (``ops.decode_step``)."""
        if output is None:
            output = torch.empty_like(query)
        ops.decode_step(output, lse, query, key, value, key_cache, value_cache,
                        slot_mapping.flatten(), num_kv_heads, scale, block_tables, seq_lens,
                        value_cache.shape[3], max_seq_len, alibi_slopes, kv_cache_dtype,
                        k_scale, v_scale)
        return output

    @staticmethod
    def forward_prefix(
        query: torch.Tensor,
        key: Optional[torch.Tensor],
        value: Optional[torch.Tensor],
        kv_cache_dtype: str,
        key_cache: torch.Tensor,
        value_cache: torch.Tensor,
        block_tables: torch.Tensor,
        query_start_loc: torch.Tensor,
        seq_lens_tensor: torch.Tensor,
        max_query_len: int,
        alibi_slopes: Optional[torch.Tensor],
        sliding_window: Optional[int],
        k_scale: torch.Tensor,
        v_scale: torch.Tensor,
        scale: Optional[float] = None,
        output: Optional[torch.Tensor] = None,
        lse: Optional[torch.Tensor] = None,
    ) -> torch.Tensor:
        """``PagedAttention.forward_prefix`` (:200-242 of the reference's ``paged_attn.py``):
        attention of each sequence's new tokens over its context AND the new tokens, causal.
        The reference's Triton kernel takes the new tokens' K/V from ``key`` / ``value`` and the
        context from the cache; here everything is read from the cache, which the backends have
        written before this call (``cpu_attn.py:548-612`` does the same with its
        ``flash_attn_varlen_func``) -- ``key`` / ``value`` are accepted for signature parity and
        not read.  ``scale`` defaults to ``head_size ** -0.5`` like the Triton kernel's."""
        if sliding_window is not None and sliding_window > 0:
            raise NotImplementedError("sliding_window is not supported by the paged_attention ops")
        if output is None:
            output = torch.empty_like(query)
        if scale is None:
            scale = float(query.shape[-1]) ** -0.5
        ops.prefill_attention(output, lse, query, key_cache, value_cache, scale, block_tables,
                              query_start_loc.to(torch.int32), seq_lens_tensor.to(torch.int32),
                              max_query_len, alibi_slopes, kv_cache_dtype, k_scale, v_scale)
        return output

    @staticmethod
    def forward_decode_with_lse(
        query: torch.Tensor,
        key_cache: torch.Tensor,
        value_cache: torch.Tensor,
        block_tables: torch.Tensor,
        seq_lens: torch.Tensor,
        max_seq_len: int,
        kv_cache_dtype: str,
        num_kv_heads: int,
        scale: float,
        alibi_slopes: Optional[torch.Tensor],
        k_scale: torch.Tensor,
        v_scale: torch.Tensor,
        output: Optional[torch.Tensor] = None,
        lse: Optional[torch.Tensor] = None,
    ) -> Tuple[torch.Tensor, torch.Tensor]:
        """Decode + softmax log-sum-exp ``[num_seqs, num_heads]`` fp32 (no reference twin: it is
        what ``AttentionImpl.can_return_lse_for_decode``, ``abstract.py:115-138``, asks for)."""
        if output is None:
            output = torch.empty_like(query)
        if lse is None:
            lse = torch.empty(query.shape[:2], dtype=torch.float32, device=query.device)
        ops.paged_attention_lse(output, lse, query, key_cache, value_cache, num_kv_heads, scale,
                                block_tables, seq_lens, value_cache.shape[3], max_seq_len,
                                alibi_slopes, kv_cache_dtype, k_scale, v_scale)
        return output, lse

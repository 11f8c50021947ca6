"""Attention backend for the decode hot path -- the model-facing contract.

Mirrors the reference's ``AttentionBackend`` / ``AttentionImpl`` /
``AttentionLayer`` contract (``vllm/attention/backends/abstract.py:27-181``) and
the metadata-builder shape of its V1 backends (``vllm/v1/attention/backends/
cpu_attn.py:324-440``, ``rocm_attn.py:279-426``): same method names, argument
meaning and error behaviour, so that registering this class as
``Attention(..., attn_backend=B200PagedAttentionBackend)``
(``vllm/attention/layer.py:101,184-192``) or through a platform plugin's
``get_attn_backend_cls`` is a pure wiring change.

It is deliberately free of ``import vllm`` at import time: the classes below are structural
twins of the ABCs, so the package works (and is tested) without the reference installed;
:func:`bind_abcs` / :func:`bind_vllm` return real subclasses of ``AttentionBackend`` /
``AttentionImpl`` when the reference's (or an installed vLLM's) ``abstract`` module is
importable.  Batches are split like the reference's paged backends split them
(``cpu_attn.py:584-638``): single-token sequences go to the decode kernel
(``decode_step`` / ``paged_attention_v1``), sequences with longer queries (prefill, chunked
prefill) to the causal chunk kernel over the paged cache (``ops.prefill_attention``,
csrc/paged_prefill.cu); the metadata builder reorders the batch decode-first
(``reorder_batch``, ``v1/attention/backends/utils.py:774-834``) and records the split.
Head sizes / block sizes the chunk kernel does not serve fall back to one single-query
attention per token (:func:`expand_queries_to_sequences`): correct, O(T x context).
"""
from __future__ import annotations

import os

from dataclasses import dataclass
from typing import Any, List, Optional, Tuple, Type

import torch

from .paged_attn import PagedAttention


class AttentionType:
    """``abstract.py:12-24`` of the reference."""
    DECODER = "decoder"
    ENCODER = "encoder"
    ENCODER_ONLY = "encoder_only"
    ENCODER_DECODER = "encoder_decoder"


@dataclass
class B200PagedAttentionMetadata:
    """Per-step tensors the runner hands to ``AttentionImpl.forward``.

    Field meaning follows the reference's V1 metadata classes
    (``cpu_attn.py:268-322``): everything lives on the device already.
    """
    num_actual_tokens: int            # tokens in the batch, excluding CUDA-graph padding
    max_query_len: int                # 1 for a pure decode batch
    seq_lens: torch.Tensor            # [num_seqs] int32, tokens seen so far (incl. the new one)
    max_seq_len: int                  # host copy of seq_lens.max()
    block_table: torch.Tensor         # [num_seqs, max_blocks_per_seq] int32
    slot_mapping: torch.Tensor        # [num_actual_tokens] int64, <0 = skip
    query_start_loc: Optional[torch.Tensor] = None  # [num_seqs + 1] int32
    # the decode-first split of a reordered batch (``split_decodes_and_prefills``,
    # v1/attention/backends/utils.py:721-771): sequences [0, num_decodes) have one query token
    # each, the remaining num_prefills longer chunks.  None: unknown (not reordered).
    num_decodes: Optional[int] = None
    num_prefills: Optional[int] = None
    num_prefill_tokens: Optional[int] = None
    max_decode_seq_len: Optional[int] = None        # host copy of seq_lens[:num_decodes].max()

    @property
    def num_decode_tokens(self) -> int:
        if self.max_query_len <= 1:
            return self.num_actual_tokens
        return self.num_decodes if self.num_decodes is not None else 0


def expand_queries_to_sequences(meta: "B200PagedAttentionMetadata"):
    """Prefill / chunked-prefill / mixed batches on the decode kernel: every query token becomes
    a single-query "sequence" that shares its sequence's block-table row and whose length is its
    causal context (``seq_len - query_len + position + 1``), so the kernel's ``seq_lens`` mask IS
    the causal mask.  The reference's paged backends delegate such batches to a second kernel
    (``cpu_attn.py:584-612``, ``rocm_attn.py:360-426``); this keeps them on the one sm_100a
    kernel -- O(T x context) KV reads, i.e. a correct fallback, not a prefill kernel.
    Device-side indexing only (no host synchronisation): returns
    ``(block_table[T, max_blocks], seq_lens[T])``."""
    n = meta.num_actual_tokens
    qsl = meta.query_start_loc
    if qsl is None:
        raise ValueError("batches with max_query_len > 1 need attn_metadata.query_start_loc")
    num_seqs = qsl.shape[0] - 1
    qsl = qsl.to(torch.int64)
    qlens = qsl[1:] - qsl[:-1]
    seqs = torch.arange(num_seqs, device=qsl.device)
    seq_of_tok = torch.repeat_interleave(seqs, qlens, output_size=n)
    pos = torch.arange(n, device=qsl.device) - qsl[seq_of_tok]
    ctx = meta.seq_lens[:num_seqs].to(torch.int64)[seq_of_tok] - qlens[seq_of_tok] + pos + 1
    return meta.block_table.index_select(0, seq_of_tok).contiguous(), ctx.to(torch.int32)


class B200PagedAttentionMetadataBuilder:
    """Builds :class:`B200PagedAttentionMetadata` from the runner's
    ``CommonAttentionMetadata`` (``vllm/v1/attention/backends/utils.py``), which is
    duck-typed here: ``num_actual_tokens, max_query_len, seq_lens, max_seq_len,
    block_table_tensor, slot_mapping, query_start_loc``."""

    # decode-only batches can be captured in a full CUDA graph: every op of the
    # path is asynchronous on the current stream and allocation-free.
    cudagraph_support = "UNIFORM_SINGLE_TOKEN_DECODE"
    reorder_batch_threshold = 1

    def __init__(self, kv_cache_spec: Any = None, layer_names: Optional[List[str]] = None,
                 vllm_config: Any = None, device: Optional[torch.device] = None) -> None:
        self.kv_cache_spec = kv_cache_spec
        self.layer_names = layer_names
        self.vllm_config = vllm_config
        self.device = device

    def reorder_batch(self, input_batch: Any, scheduler_output: Any) -> bool:
        """Decode-first order with the fewest swaps: what the runner does for a builder with a
        ``reorder_batch_threshold`` (``reorder_batch_to_split_decodes_and_prefills``,
        ``v1/attention/backends/utils.py:774-834``; the CPU backend's own ``reorder_batch``,
        ``cpu_attn.py:346-388``, is the mirror image).  A request is a decode when it has at most
        ``reorder_batch_threshold`` scheduled tokens.  Returns True if anything moved."""
        sched = scheduler_output.num_scheduled_tokens
        is_decode = [sched[rid] <= self.reorder_batch_threshold for rid in input_batch.req_ids]
        num_decodes = sum(is_decode)
        # decodes sitting behind the decode region, last first / prefills inside it, first first
        late_decodes = [i for i in range(len(is_decode) - 1, num_decodes - 1, -1) if is_decode[i]]
        early_prefills = [i for i in range(num_decodes) if not is_decode[i]]
        assert len(late_decodes) == len(early_prefills)
        for src, dst in zip(late_decodes, early_prefills):
            input_batch.swap_states(dst, src)
        return bool(late_decodes)

    def build(self, common_prefix_len: int, common_attn_metadata: Any,
              fast_build: bool = False) -> B200PagedAttentionMetadata:
        m = common_attn_metadata
        max_seq_len = getattr(m, "max_seq_len", None)
        seq_lens_cpu = getattr(m, "seq_lens_cpu", None)
        if max_seq_len is None:
            max_seq_len = int(seq_lens_cpu.max()) if seq_lens_cpu is not None else int(m.seq_lens.max())
        meta = B200PagedAttentionMetadata(
            num_actual_tokens=int(m.num_actual_tokens),
            max_query_len=int(m.max_query_len),
            seq_lens=m.seq_lens,
            max_seq_len=int(max_seq_len),
            block_table=m.block_table_tensor,
            slot_mapping=m.slot_mapping,
            query_start_loc=getattr(m, "query_start_loc", None),
        )
        qsl_cpu = getattr(m, "query_start_loc_cpu", None)
        if meta.max_query_len > 1 and qsl_cpu is not None:
            # split_decodes_and_prefills (utils.py:721-771) on the host copy: the leading run of
            # single-token sequences is the decode part.  The decode call takes sequence b to be
            # token b, so the run ends at the first sequence that does not have EXACTLY one token
            # (a zero-length entry would shift every later token onto its neighbour's block
            # table); the chunk kernel serves whatever follows, zero-length entries included.
            qlens = (qsl_cpu[1:] - qsl_cpu[:-1]).tolist()
            nd = 0
            while nd < len(qlens) and qlens[nd] == 1:
                nd += 1
            meta.num_decodes = nd
            meta.num_prefills = len(qlens) - nd
            meta.num_prefill_tokens = meta.num_actual_tokens - int(qsl_cpu[nd])
            if nd and seq_lens_cpu is not None:
                meta.max_decode_seq_len = int(seq_lens_cpu[:nd].max())
        return meta

    def build_for_cudagraph_capture(self, common_attn_metadata: Any) -> B200PagedAttentionMetadata:
        return self.build(0, common_attn_metadata)


class B200PagedAttentionImpl:
    """``AttentionImpl`` (``abstract.py:113-181``) for paged decode on B200."""

    # abstract.py:115-138 of the reference: the impl can hand back the softmax LSE of the decode
    # (ops.paged_attention_lse); the runner sets dcp_world_size / dcp_rank and the flag below
    # becomes true when decode context parallelism is on.
    can_return_lse_for_decode: bool = True
    need_to_return_lse_for_decode: bool = False
    dcp_world_size: int = 1
    dcp_rank: int = 0
    dcp_group: Any = None             # torch.distributed group of the DCP ranks
    # pure decode batches: cache write + attention in one launch (ops.decode_step);
    # B200PA_FUSED_STEP=0 keeps the reference's two calls (A/B runs)
    fused_decode_step: bool = os.environ.get("B200PA_FUSED_STEP", "1") != "0"

    def __init__(
        self,
        num_heads: int,
        head_size: int,
        scale: float,
        num_kv_heads: Optional[int] = None,
        alibi_slopes: Optional[List[float]] = None,
        sliding_window: Optional[int] = None,
        kv_cache_dtype: str = "auto",
        logits_soft_cap: Optional[float] = None,
        attn_type: str = AttentionType.DECODER,
        kv_sharing_target_layer_name: Optional[str] = None,
        **extra: Any,
    ) -> None:
        if kv_sharing_target_layer_name is not None:
            raise NotImplementedError("KV sharing is not supported by B200PagedAttentionImpl")
        if logits_soft_cap is not None:
            raise NotImplementedError("logits_soft_cap is not supported by the paged_attention ops")
        if sliding_window is not None:
            raise NotImplementedError("sliding_window is not supported by the paged_attention ops")
        if attn_type != AttentionType.DECODER:
            raise NotImplementedError("Encoder self-attention and encoder/decoder cross-attention "
                                      "are not implemented for B200PagedAttentionImpl")
        self.num_heads = num_heads
        self.head_size = head_size
        self.scale = float(scale)
        self.num_kv_heads = num_heads if num_kv_heads is None else num_kv_heads
        if alibi_slopes is not None:
            alibi_slopes = torch.tensor(alibi_slopes, dtype=torch.float32)
        self.alibi_slopes = alibi_slopes
        self.kv_cache_dtype = kv_cache_dtype
        assert self.num_heads % self.num_kv_heads == 0
        self.num_queries_per_kv = self.num_heads // self.num_kv_heads
        B200PagedAttentionBackend.validate_head_size(head_size)

    def forward(
        self,
        layer: Any,                      # AttentionLayer: supplies _k_scale / _v_scale
        query: torch.Tensor,             # [num_tokens, num_heads * head_size] or [T, H, D]
        key: Optional[torch.Tensor],     # [num_tokens, num_kv_heads * head_size] or [T, Hkv, D]
        value: Optional[torch.Tensor],
        kv_cache: torch.Tensor,          # get_kv_cache_shape(...): [2, num_blocks, BS*Hkv*D]
        attn_metadata: Optional[B200PagedAttentionMetadata],
        output: Optional[torch.Tensor] = None,
        output_scale: Optional[torch.Tensor] = None,
        output_block_scale: Optional[torch.Tensor] = None,
    ) -> torch.Tensor:
        if output_scale is not None or output_block_scale is not None:
            raise NotImplementedError("fused output quantization is not yet supported "
                                      "for B200PagedAttentionImpl")
        if output is None:
            output = torch.empty_like(query)
        if attn_metadata is None:        # profiling run
            return output
        if attn_metadata.max_query_len > 1:
            if self.need_to_return_lse_for_decode or self.dcp_world_size > 1:
                raise NotImplementedError("decode context parallelism serves decode batches only")
            return self._forward_multi_query(layer, query, key, value, kv_cache, attn_metadata,
                                             output)
        n = attn_metadata.num_actual_tokens
        q = query.view(-1, self.num_heads, self.head_size)
        out = output.view(-1, self.num_heads, self.head_size)
        key_cache, value_cache = PagedAttention.split_kv_cache(kv_cache, self.num_kv_heads,
                                                               self.head_size)
        if self.alibi_slopes is not None and self.alibi_slopes.device != q.device:
            self.alibi_slopes = self.alibi_slopes.to(q.device)
        want_lse = self.need_to_return_lse_for_decode or self.dcp_world_size > 1
        k = v = None
        if key is not None and value is not None:
            k = key.view(-1, self.num_kv_heads, self.head_size)
            v = value.view(-1, self.num_kv_heads, self.head_size)
        # pure decode batch, one new token per sequence: cache write + attention in one launch
        fused = (k is not None and self.fused_decode_step and not want_lse
                 and k.shape[0] == n and attn_metadata.slot_mapping.numel() == n
                 and attn_metadata.seq_lens.shape[0] == n)
        if fused:
            PagedAttention.write_and_decode(
                q[:n], k, v, key_cache, value_cache, attn_metadata.slot_mapping,
                attn_metadata.block_table, attn_metadata.seq_lens, attn_metadata.max_seq_len,
                self.kv_cache_dtype, self.num_kv_heads, self.scale, self.alibi_slopes,
                layer._k_scale, layer._v_scale, output=out[:n])
            return output
        if k is not None:
            PagedAttention.write_to_paged_cache(k, v, key_cache, value_cache,
                                                attn_metadata.slot_mapping, self.kv_cache_dtype,
                                                layer._k_scale, layer._v_scale)
        if want_lse:
            # decode context parallelism: this rank's KV slice with ALL query heads, then the
            # reference's cp_lse_ag_out_rs (common.py:119-140): gather LSEs, rescale, reduce-scatter
            from .dcp import cp_lse_ag_out_rs
            part = torch.empty_like(q[:n])
            part, lse = PagedAttention.forward_decode_with_lse(
                q[:n], key_cache, value_cache, attn_metadata.block_table,
                attn_metadata.seq_lens, attn_metadata.max_seq_len, self.kv_cache_dtype,
                self.num_kv_heads, self.scale, self.alibi_slopes, layer._k_scale, layer._v_scale,
                output=part)
            return cp_lse_ag_out_rs(part, lse, self.dcp_group)
        PagedAttention.forward_decode(
            q[:n], key_cache, value_cache, attn_metadata.block_table, attn_metadata.seq_lens,
            attn_metadata.max_seq_len, self.kv_cache_dtype, self.num_kv_heads, self.scale,
            self.alibi_slopes, layer._k_scale, layer._v_scale, output=out[:n])
        return output

    # shapes the chunk kernel (csrc/paged_prefill.cu) serves; others fall back to the expansion
    _PREFILL_HEAD_SIZES = (32, 64, 96, 128, 192, 256)

    def _chunk_kernel_serves(self, query: torch.Tensor, block_size: int, num_seqs: int = 0) -> bool:
        # num_seqs: the chunk kernel puts the sequences on grid.z (at most 65535)
        return (self.head_size in self._PREFILL_HEAD_SIZES and block_size in (16, 32)
                and query.dtype in (torch.float16, torch.bfloat16) and num_seqs <= 65535)

    def _forward_multi_query(self, layer, query, key, value, kv_cache, attn_metadata, output):
        """Batches with more than one query token per sequence (prefill, chunked prefill, mixed),
        split the way the reference's paged backends split them (``cpu_attn.py:584-638``): write
        every new token, then the leading single-token sequences (a decode-first batch,
        :meth:`B200PagedAttentionMetadataBuilder.reorder_batch`) on the decode kernel and the
        chunks on the causal chunk kernel over the paged cache.  Without the split (a batch that
        was not reordered) the chunk kernel takes every sequence: it is correct for any mix."""
        m = attn_metadata
        n = m.num_actual_tokens
        q = query.view(-1, self.num_heads, self.head_size)
        out = output.view(-1, self.num_heads, self.head_size)
        key_cache, value_cache = PagedAttention.split_kv_cache(kv_cache, self.num_kv_heads,
                                                               self.head_size)
        if key is not None and value is not None:
            PagedAttention.write_to_paged_cache(
                key.view(-1, self.num_kv_heads, self.head_size),
                value.view(-1, self.num_kv_heads, self.head_size), key_cache, value_cache,
                m.slot_mapping, self.kv_cache_dtype, layer._k_scale, layer._v_scale)
        if self.alibi_slopes is not None and self.alibi_slopes.device != q.device:
            self.alibi_slopes = self.alibi_slopes.to(q.device)
        if m.query_start_loc is None:
            raise ValueError("batches with max_query_len > 1 need attn_metadata.query_start_loc")
        if not self._chunk_kernel_serves(q, value_cache.shape[3], m.query_start_loc.shape[0] - 1):
            bt_tok, sl_tok = expand_queries_to_sequences(m)
            PagedAttention.forward_decode(
                q[:n], key_cache, value_cache, bt_tok, sl_tok, m.max_seq_len,
                self.kv_cache_dtype, self.num_kv_heads, self.scale, self.alibi_slopes,
                layer._k_scale, layer._v_scale, output=out[:n])
            return output
        num_seqs = m.query_start_loc.shape[0] - 1
        nd = m.num_decodes or 0
        if nd > 0:
            # decode part: sequence b < nd is token b (a decode-first batch)
            PagedAttention.forward_decode(
                q[:nd], key_cache, value_cache, m.block_table[:nd], m.seq_lens[:nd],
                m.max_decode_seq_len or m.max_seq_len, self.kv_cache_dtype, self.num_kv_heads,
                self.scale, self.alibi_slopes, layer._k_scale, layer._v_scale, output=out[:nd])
        if nd < num_seqs:
            # chunk part: the kernel indexes tokens through query_start_loc, so the whole q / out
            # go in with the tail of the per-sequence tensors (views, no copies)
            from . import _custom_ops as ops
            ops.prefill_attention(out[:n], None, q[:n], key_cache, value_cache, self.scale,
                                  m.block_table[nd:], m.query_start_loc[nd:],
                                  m.seq_lens[nd:num_seqs], m.max_query_len, self.alibi_slopes,
                                  self.kv_cache_dtype, layer._k_scale, layer._v_scale)
        return output

    def fused_output_quant_supported(self, quant_key: Any) -> bool:
        return False


class B200PagedAttentionBackend:
    """``AttentionBackend`` (``abstract.py:27-82``) twin."""

    accept_output_buffer: bool = True
    supports_quant_query_input: bool = False

    @staticmethod
    def get_name() -> str:
        return "B200_PAGED_ATTN"

    @staticmethod
    def get_impl_cls() -> Type[B200PagedAttentionImpl]:
        return B200PagedAttentionImpl

    @staticmethod
    def get_metadata_cls() -> Type[B200PagedAttentionMetadata]:
        return B200PagedAttentionMetadata

    @classmethod
    def make_metadata(cls, *args, **kwargs) -> B200PagedAttentionMetadata:
        return cls.get_metadata_cls()(*args, **kwargs)

    @staticmethod
    def get_builder_cls() -> Type[B200PagedAttentionMetadataBuilder]:
        return B200PagedAttentionMetadataBuilder

    @staticmethod
    def get_kv_cache_shape(num_blocks: int, block_size: int, num_kv_heads: int,
                           head_size: int) -> Tuple[int, ...]:
        return PagedAttention.get_kv_cache_shape(num_blocks, block_size, num_kv_heads, head_size)

    @staticmethod
    def get_kv_cache_stride_order() -> Tuple[int, ...]:
        # (2, num_blocks, block_size * num_kv_heads * head_size) is allocated as it is indexed
        return (0, 1, 2)

    @classmethod
    def full_cls_name(cls) -> Tuple[str, str]:
        return (cls.__module__, cls.__qualname__)

    @classmethod
    def get_supported_dtypes(cls) -> List[torch.dtype]:
        return [torch.float16, torch.bfloat16, torch.float32]

    @staticmethod
    def get_supported_head_sizes() -> List[int]:
        return PagedAttention.get_supported_head_sizes()

    @classmethod
    def validate_head_size(cls, head_size: int) -> None:
        supported = cls.get_supported_head_sizes()
        if head_size not in supported:
            raise ValueError(f"Head size {head_size} is not supported by {cls.get_name()}. "
                             f"Supported head sizes are: {supported}.")

    @staticmethod
    def use_cascade_attention(*args, **kwargs) -> bool:
        return False


# --------------------------------------------------------------------------------------------
# The same decode path over the TOKEN-MAJOR cache: the layout every flash-
# attention style backend of the reference allocates (``FlashAttentionBackend.get_kv_cache_shape``
# = ``(2, num_blocks, block_size, num_kv_heads, head_size)``), ``reshape_and_cache_flash`` writes
# (``csrc/cache_kernels.cu:280-353``) and the KV connectors move
# (``distributed/kv_transfer/kv_connector/utils.py:91``).  One launch per decode step:
# ``ops.paged_attention_flash`` with the new token's key / value / slot.
# --------------------------------------------------------------------------------------------
class B200FlashLayoutAttentionImpl(B200PagedAttentionImpl):
    """``AttentionImpl`` for decode on a ``(2, num_blocks, block_size, num_kv_heads, head_size)``
    cache (NHD, or HND when the runner permutes it: strides are honoured)."""

    def forward(self, layer: Any, query: torch.Tensor, key: Optional[torch.Tensor],
                value: Optional[torch.Tensor], kv_cache: torch.Tensor,
                attn_metadata: Optional[B200PagedAttentionMetadata],
                output: Optional[torch.Tensor] = None,
                output_scale: Optional[torch.Tensor] = None,
                output_block_scale: Optional[torch.Tensor] = None) -> torch.Tensor:
        from . import _custom_ops as ops
        if output_scale is not None or output_block_scale is not None:
            raise NotImplementedError("fused output quantization is not yet supported "
                                      "for B200FlashLayoutAttentionImpl")
        if output is None:
            output = torch.empty_like(query)
        if attn_metadata is None:        # profiling run
            return output
        multi = attn_metadata.max_query_len > 1
        n = attn_metadata.num_actual_tokens
        q = query.view(-1, self.num_heads, self.head_size)
        out = output.view(-1, self.num_heads, self.head_size)
        key_cache, value_cache = kv_cache.unbind(0)
        if self.alibi_slopes is not None and self.alibi_slopes.device != q.device:
            self.alibi_slopes = self.alibi_slopes.to(q.device)
        want_lse = self.need_to_return_lse_for_decode or self.dcp_world_size > 1
        if multi and want_lse:
            raise NotImplementedError("decode context parallelism serves decode batches only")
        block_table, seq_lens = attn_metadata.block_table, attn_metadata.seq_lens
        if multi:   # prefill / chunked prefill / mixed: one single-query attention per token
            block_table, seq_lens = expand_queries_to_sequences(attn_metadata)
        k = v = slots = None
        if key is not None and value is not None:
            k = key.view(-1, self.num_kv_heads, self.head_size)
            v = value.view(-1, self.num_kv_heads, self.head_size)
            slots = attn_metadata.slot_mapping.flatten()
            fused = (self.fused_decode_step and not want_lse and not multi and k.shape[0] == n
                     and slots.numel() == n and attn_metadata.seq_lens.shape[0] == n)
            if not fused:
                ops.reshape_and_cache_flash(k, v, key_cache, value_cache, slots,
                                            self.kv_cache_dtype, layer._k_scale, layer._v_scale)
                k = v = slots = None
        lse = None
        part = out[:n]
        if want_lse:
            part = torch.empty_like(q[:n])
            lse = torch.empty(n, self.num_heads, dtype=torch.float32, device=q.device)
        ops.paged_attention_flash(part, lse, q[:n], k, v, key_cache, value_cache, slots, self.scale,
                                  block_table, seq_lens, attn_metadata.max_seq_len, self.alibi_slopes,
                                  self.kv_cache_dtype, layer._k_scale, layer._v_scale)
        if want_lse:
            from .dcp import cp_lse_ag_out_rs
            return cp_lse_ag_out_rs(part, lse, self.dcp_group)
        return output


class B200FlashLayoutAttentionBackend(B200PagedAttentionBackend):
    """``AttentionBackend`` twin whose cache is the flash layout."""

    @staticmethod
    def get_name() -> str:
        return "B200_PAGED_ATTN_FLASH_LAYOUT"

    @staticmethod
    def get_impl_cls() -> Type[B200FlashLayoutAttentionImpl]:
        return B200FlashLayoutAttentionImpl

    @staticmethod
    def get_kv_cache_shape(num_blocks: int, block_size: int, num_kv_heads: int,
                           head_size: int, cache_dtype_str: str = "auto") -> Tuple[int, ...]:
        if block_size not in (16, 32):
            raise ValueError("B200_PAGED_ATTN_FLASH_LAYOUT serves block_size 16 and 32")
        return (2, num_blocks, block_size, num_kv_heads, head_size)

    @staticmethod
    def get_kv_cache_stride_order() -> Tuple[int, ...]:
        # NHD physical order; a runner that wants HND permutes dims 2 and 3 (the kernel reads strides)
        return (0, 1, 2, 3, 4)

    @staticmethod
    def get_supported_head_sizes() -> List[int]:
        return [32, 64, 80, 96, 112, 128, 192, 256]

    @classmethod
    def get_supported_dtypes(cls) -> List[torch.dtype]:
        return [torch.float16, torch.bfloat16]


# --------------------------------------------------------------------------------------------
# Real subclasses of the reference's ABCs, on demand
# --------------------------------------------------------------------------------------------
def bind_abcs(abstract_module: Any, flash_layout: bool = False):
    """``(Backend, Impl)`` as subclasses of ``abstract_module.AttentionBackend`` /
    ``.AttentionImpl`` -- the reference's ``vllm/attention/backends/abstract.py:27-181`` (or an
    installed vLLM's equivalent): what ``Attention(..., attn_backend=Backend)``
    (``vllm/attention/layer.py:101,184-198``) and ``get_attn_backend_cls`` expect to receive.
    The twins above implement every abstract method, so the subclasses add nothing but the base."""
    backend_twin = B200FlashLayoutAttentionBackend if flash_layout else B200PagedAttentionBackend
    impl_twin = backend_twin.get_impl_cls()
    impl = type(impl_twin.__name__ + "Bound", (impl_twin, abstract_module.AttentionImpl), {})
    backend = type(backend_twin.__name__ + "Bound", (backend_twin, abstract_module.AttentionBackend),
                   {"get_impl_cls": staticmethod(lambda: impl)})
    return backend, impl


def bind_vllm(flash_layout: bool = False):
    """:func:`bind_abcs` on whichever vLLM is importable: the reference's module path first, then
    the one newer releases moved the ABCs to.  Raises ImportError when there is none."""
    import importlib
    err: Optional[Exception] = None
    for name in ("vllm.attention.backends.abstract", "vllm.v1.attention.backend"):
        try:
            return bind_abcs(importlib.import_module(name), flash_layout)
        except ImportError as exc:      # try the next layout
            err = exc
    raise ImportError(f"no vLLM attention ABCs importable: {err}")

"""TEST INFRASTRUCTURE ONLY -- CPU restatement of causal chunk attention over the paged cache.

Nothing here is imported by the product.  fp32 torch on CPU tensors.

What it restates: the second kernel the reference's paged backends hand batches with more than
one query token per sequence to -- ``flash_attn_varlen_func(out, query, key_cache, value_cache,
cu_seqlens_q, cu_seqlens_kv, ..., causal=True, block_table, alibi_slopes)`` as called in
``vllm/v1/attention/backends/cpu_attn.py:584-612`` (everything read from the cache, which the
backend has written at :548-566), and the Triton ``context_attention_fwd`` behind
``PagedAttention.forward_prefix`` (``vllm/attention/ops/paged_attn.py:200-242``).  The math
follows the reference's own test oracle for exactly that call, ``ref_paged_attn``
(``tests/kernels/attention/test_flash_attn.py:27-79``): per sequence, gather the first ``kv_len``
cache rows through the block table, logits = scale * q.k, mask ``triu(diagonal = kv_len - q_len
+ 1)`` (query token i sees positions <= kv_len - q_len + i), softmax, rounding of P to the value
type, P.V -- restated here on the X-PACKED cache layout of the paged ops, with the fp8 dequant
of ``quant_utils.cuh:306-428`` (value = fp8 -> float, times k_scale / v_scale) and the ALiBi bias
of ``csrc/attention/attention_kernels.cuh:316-319`` generalised to a chunk
(slope * (key position - query position)).

Pinned by tests/golden/prefill_*.npz: outputs of ``ref_paged_attn`` itself, extracted from the
reference at fixture-generation time (tests/golden/make_prefill_golden.py) and run on the same
caches re-indexed to its token-major layout.
"""
from __future__ import annotations

from typing import Optional, Sequence

import torch


def _dequant(cache: torch.Tensor, kv_cache_dtype: str, scale: float) -> torch.Tensor:
    if kv_cache_dtype == "auto":
        return cache.float()
    if kv_cache_dtype in ("fp8", "fp8_e4m3"):
        return cache.view(torch.float8_e4m3fn).float() * scale
    if kv_cache_dtype == "fp8_e5m2":
        return cache.view(torch.float8_e5m2).float() * scale
    raise ValueError(kv_cache_dtype)


def prefill_attention(query: torch.Tensor, key_cache: torch.Tensor, value_cache: torch.Tensor,
                      block_tables: torch.Tensor, query_start_loc: Sequence[int],
                      seq_lens: Sequence[int], scale: float,
                      alibi_slopes: Optional[torch.Tensor] = None, kv_cache_dtype: str = "auto",
                      k_scale: float = 1.0, v_scale: float = 1.0, round_p: bool = False,
                      return_lse: bool = False):
    """query [T, H, D]; key_cache [NB, Hkv, D/x, BS, x]; value_cache [NB, Hkv, D, BS].
    Returns fp32 [T, H, D] (and the natural-log sum-exp [T, H] with ``return_lse``).
    ``round_p``: round the probabilities to the query's dtype before P.V as ``ref_paged_attn``
    does (``.to(v.dtype)``, test_flash_attn.py:73)."""
    T, H, D = query.shape
    NB, Hkv, Dx, BS, x = key_cache.shape
    G = H // Hkv
    k_log = _dequant(key_cache, kv_cache_dtype, k_scale).permute(0, 3, 1, 2, 4).reshape(NB, BS, Hkv, D)
    v_log = _dequant(value_cache, kv_cache_dtype, v_scale).permute(0, 3, 1, 2)
    out = torch.zeros(T, H, D, dtype=torch.float32)
    lse = torch.full((T, H), float("-inf"), dtype=torch.float32)
    for b in range(len(seq_lens)):
        q0, q1 = int(query_start_loc[b]), int(query_start_loc[b + 1])
        q_len, kv_len = q1 - q0, int(seq_lens[b])
        if q_len <= 0:
            continue
        nb = (kv_len + BS - 1) // BS
        blocks = block_tables[b, :nb].long()
        k = k_log[blocks].reshape(nb * BS, Hkv, D)[:kv_len].repeat_interleave(G, dim=1)
        v = v_log[blocks].reshape(nb * BS, Hkv, D)[:kv_len].repeat_interleave(G, dim=1)
        q = query[q0:q1].float()
        attn = scale * torch.einsum("qhd,khd->hqk", q, k)
        qpos = torch.arange(q_len).view(-1, 1) + (kv_len - q_len)
        kpos = torch.arange(kv_len).view(1, -1)
        if alibi_slopes is not None:
            attn = attn + alibi_slopes.float().view(-1, 1, 1) * (kpos - qpos).float()
        attn = attn.masked_fill(kpos > qpos, float("-inf"))
        lse[q0:q1] = torch.logsumexp(attn, dim=-1).transpose(0, 1)
        p = torch.softmax(attn, dim=-1)
        if round_p:
            p = p.to(query.dtype).float()
        out[q0:q1] = torch.einsum("hqk,khd->qhd", p, v)
    return (out, lse) if return_lse else out

"""TEST INFRASTRUCTURE ONLY -- CPU restatements of the log-sum-exp side of the decode path.

Nothing here is imported by the product.  Everything runs on CPU torch tensors in fp32.

* ``merge_attn_states``   csrc/attention/merge_attn_states.cu:14-89 of the reference (the CUDA
                          kernel; launcher :131-170) -- pinned by tests/golden/merge_*.npz, which
                          hold outputs of the reference's own ``merge_attn_states_torch``
                          (tests/kernels/attention/test_merge_attn_states.py:16-46), and on the
                          GPU box by the reference CUDA kernel itself (oracle/_ref).
* ``correct_attn_out``    vllm/attention/ops/common.py:9-68 (Triton kernel) and :86-116 (wrapper).
                          Pinned by tests/golden/correct_*.npz: the reference's kernel SOURCE,
                          extracted and executed per (batch, head) against a torch implementation
                          of the tl. primitives it uses (tests/golden/make_correct_golden.py;
                          Triton itself needs a GPU), inf / nan / all -inf rows included -- and
                          through the property it exists for (the sum over ranks of the corrected
                          partial outputs is the full attention output, :119-140).
* ``paged_attention_lse`` no reference op.  LSE of the scaled logits from the oracle's own v2
                          partials (attention_kernels.cuh:337-341 exp_sums / max_logits):
                          lse = log(sum_j exp_sums[j] * exp(max_logits[j] - m)) + m.
"""
from __future__ import annotations

from typing import Optional

import torch

from . import oracle as _o


def merge_attn_states(output: torch.Tensor, prefix_output: torch.Tensor,
                      prefix_lse: torch.Tensor, suffix_output: torch.Tensor,
                      suffix_lse: torch.Tensor,
                      output_lse: Optional[torch.Tensor] = None) -> None:
    """merge_attn_states.cu:40-87: per (token, head), fp32 arithmetic, one rounding to scalar_t.
    output [n,h,d]; lse tensors [h,n]."""
    p_lse = prefix_lse.float().clone()
    s_lse = suffix_lse.float().clone()
    p_lse[torch.isinf(p_lse)] = -torch.inf           # :45-46 (both signs of inf)
    s_lse[torch.isinf(s_lse)] = -torch.inf
    max_lse = torch.maximum(p_lse, s_lse)             # :48
    p_se = torch.exp(p_lse - max_lse)                 # :49-52
    s_se = torch.exp(s_lse - max_lse)
    out_se = p_se + s_se                              # :53
    p_scale = (p_se / out_se).t().unsqueeze(-1)       # :54-55, [n,h,1]
    s_scale = (s_se / out_se).t().unsqueeze(-1)
    # :75 "p_out_f * p_scale + (s_out_f * s_scale)" -- nvcc contracts it into
    # fma(p, p_scale, s * s_scale); the difference to two roundings is below 1 ulp of fp32 and
    # vanishes in the rounding to a 16-bit scalar_t except at ties (tests allow 1 ulp there).
    o = prefix_output.float() * p_scale + suffix_output.float() * s_scale
    output.copy_(o.to(output.dtype))
    if output_lse is not None:
        output_lse.copy_(torch.log(out_se) + max_lse)  # :85


def correct_attn_out(out: torch.Tensor, lses: torch.Tensor, cp_rank: int):
    """common.py:39-67: out [B,H,D] rescaled in place by exp(lses[rank] - lse_total);
    returns (out, lse_total [B,H])."""
    lse = lses.float().clone()
    lse[torch.isnan(lse) | (lse == torch.inf)] = -torch.inf      # :41
    lse_max = lse.max(dim=0).values                               # :42
    acc = torch.exp(lse - lse_max).sum(dim=0)                     # :43-45
    total = torch.log(acc) + lse_max                              # :46-47
    fin = lses[cp_rank].float() - total                           # :60-61
    fin[torch.isnan(fin) | (fin == torch.inf)] = -torch.inf      # :62-64
    factor = torch.exp(fin).unsqueeze(-1)                         # :65
    out.copy_((out.float() * factor).to(out.dtype))               # :66-69
    return out, total


def paged_attention_lse(out, lse, query, key_cache, value_cache, num_kv_heads, scale,
                        block_tables, seq_lens, block_size, max_seq_len, alibi_slopes,
                        kv_cache_dtype, k_scale, v_scale, *, out_f32=None) -> None:
    """Decode attention + natural-log sum-exp of the scaled logits, lse [B,H] fp32
    (-inf for an empty sequence)."""
    B, H, D = query.shape
    P = max(1, (int(max_seq_len) + 511) // 512)
    es = torch.zeros(B, H, P, dtype=torch.float32)
    ml = torch.full((B, H, P), -float("inf"), dtype=torch.float32)
    tmp = torch.zeros(B, H, P, D, dtype=query.dtype)
    _o.paged_attention_v2(out, es, ml, tmp, query, key_cache, value_cache, num_kv_heads,
                          scale, block_tables, seq_lens, block_size, max_seq_len,
                          alibi_slopes, kv_cache_dtype, k_scale, v_scale, faithful=False,
                          out_f32=out_f32)
    for b in range(B):
        n = (int(seq_lens[b]) + 511) // 512
        if n == 0:
            lse[b] = -float("inf")
            continue
        m = ml[b, :, :n].double()
        e = es[b, :, :n].double()
        mx = m.max(dim=1).values
        lse[b] = (torch.log((e * torch.exp(m - mx[:, None])).sum(dim=1)) + mx).float()

"""Decode context parallelism on top of the LSE-returning decode op.

Host-side twin of the reference's ``cp_lse_ag_out_rs`` (``vllm/attention/ops/common.py:119-140``):
every rank of the DCP group holds a slice of each sequence's KV, runs decode with all query heads
on its slice (``ops.paged_attention_lse``), then

    lses = all_gather(lse)                      # [N, B, H]
    out  = correct_attn_out(out, lses, rank)    # *= exp(lse_rank - logsumexp_n lses), in place
    out  = reduce_scatter(out, dim=1)           # sum over ranks, each rank keeps H/N heads

``group`` is a ``torch.distributed`` process group (the reference passes its ``GroupCoordinator``;
only ``all_gather`` / ``reduce_scatter`` / ``world_size`` / ``rank_in_group`` are used there, which
map 1:1 onto the calls below).  The collectives are NCCL over NVLink on the GPU box.
"""
from __future__ import annotations

from typing import Callable, Optional, Tuple

import torch
import torch.distributed as dist


def _default_correct(out: torch.Tensor, lses: torch.Tensor, rank: int):
    from . import _custom_ops as ops
    return ops.correct_attn_out(out, lses, rank)


def cp_lse_ag_out_rs(cp_attn_out: torch.Tensor, cp_attn_lse: torch.Tensor, group=None,
                     ctx=None, *,
                     correct_fn: Optional[Callable[..., Tuple[torch.Tensor, torch.Tensor]]] = None
                     ) -> torch.Tensor:
    """cp_attn_out [B, H, D], cp_attn_lse [B, H] -> this rank's [B, H/N, D] of the exact output.

    ``correct_fn`` exists for host-logic tests on CPU (gloo); the product path is the CUDA op.
    """
    world = dist.get_world_size(group)
    if world == 1:
        return cp_attn_out
    rank = dist.get_rank(group)
    B, H, D = cp_attn_out.shape
    if H % world:
        raise ValueError(f"the DCP world size {world} must divide the {H} query heads")
    cp_attn_lse = cp_attn_lse.contiguous()
    lses = torch.empty((world,) + cp_attn_lse.shape, dtype=cp_attn_lse.dtype,
                       device=cp_attn_lse.device)
    dist.all_gather_into_tensor(lses.view(world * B, H), cp_attn_lse, group=group)
    out, _ = (correct_fn or _default_correct)(cp_attn_out, lses, rank)
    # reduce_scatter along the head dimension: rank r keeps heads [r*H/N, (r+1)*H/N)
    send = out.view(B, world, H // world, D).transpose(0, 1).contiguous()
    recv = torch.empty((B, H // world, D), dtype=out.dtype, device=out.device)
    if send.device.type == "cuda":
        dist.reduce_scatter_tensor(recv, send, group=group)
    else:  # gloo has no reduce_scatter: all_reduce + slice (tests only)
        dist.all_reduce(send, group=group)
        recv.copy_(send[rank])
    return recv

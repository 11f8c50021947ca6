"""KV-head sharding of the decode path across the GPUs of one box.

The reference's tensor-parallel layout (``vllm/model_executor/models/llama.py:124-137``,
``QKVParallelLinear`` in ``vllm/model_executor/layers/linear.py:849-908``): rank ``r`` of
``tp`` owns ``H / tp`` query heads and ``max(1, Hkv / tp)`` KV heads; when ``tp > Hkv`` each
KV head is replicated on ``tp / Hkv`` consecutive ranks.  Block tables, seq_lens and
slot_mapping are replicated.  Attention for KV head ``g`` touches only query heads
``[g*G, (g+1)*G)``, so the attention path itself needs no collective.
"""
from __future__ import annotations

from dataclasses import dataclass

import torch


@dataclass(frozen=True)
class HeadShard:
    q_lo: int       # first query head owned by the rank
    q_hi: int       # one past the last
    kv_lo: int      # first KV head held by the rank
    kv_hi: int

    @property
    def num_heads(self) -> int:
        return self.q_hi - self.q_lo

    @property
    def num_kv_heads(self) -> int:
        return self.kv_hi - self.kv_lo


def head_shard(total_num_heads: int, total_num_kv_heads: int, tp_size: int, tp_rank: int) -> HeadShard:
    """The slice of heads rank ``tp_rank`` owns (``llama.py:124-137`` of the reference)."""
    if tp_size <= 0 or not 0 <= tp_rank < tp_size:
        raise ValueError(f"bad tensor-parallel coordinates {tp_rank}/{tp_size}")
    if total_num_heads % tp_size:
        raise ValueError(f"tp_size {tp_size} must divide the {total_num_heads} query heads")
    if total_num_heads % total_num_kv_heads:
        raise ValueError("query heads must be a multiple of the KV heads")
    h = total_num_heads // tp_size
    if total_num_kv_heads >= tp_size:
        if total_num_kv_heads % tp_size:
            raise ValueError(f"tp_size {tp_size} must divide the {total_num_kv_heads} KV heads")
        kv = total_num_kv_heads // tp_size
        return HeadShard(tp_rank * h, (tp_rank + 1) * h, tp_rank * kv, (tp_rank + 1) * kv)
    # more ranks than KV heads: replicate each KV head on tp_size / Hkv consecutive ranks
    if tp_size % total_num_kv_heads:
        raise ValueError(f"the {total_num_kv_heads} KV heads must divide tp_size {tp_size}")
    rep = tp_size // total_num_kv_heads
    g = tp_rank // rep
    return HeadShard(tp_rank * h, (tp_rank + 1) * h, g, g + 1)


def shard_query(query: torch.Tensor, sh: HeadShard) -> torch.Tensor:
    """[T, H, D] -> this rank's [T, H/tp, D] (a view)."""
    return query[:, sh.q_lo:sh.q_hi]


def shard_paged_cache(key_cache: torch.Tensor, value_cache: torch.Tensor, sh: HeadShard):
    """x-packed caches [NB, Hkv, ...] -> this rank's [NB, Hkv/tp, ...] (contiguous copies: each
    rank owns its own cache tensor, as each worker of the reference does)."""
    return (key_cache[:, sh.kv_lo:sh.kv_hi].contiguous(),
            value_cache[:, sh.kv_lo:sh.kv_hi].contiguous())


def gather_heads(local_out: torch.Tensor, tp_size: int, group=None) -> torch.Tensor:
    """Bench / test only: all-gather [T, H/tp, D] -> [T, H, D] in rank order (the model's
    row-parallel o_proj consumes the local slice directly; the attention path has no collective)."""
    import torch.distributed as dist
    parts = [torch.empty_like(local_out) for _ in range(tp_size)]
    dist.all_gather(parts, local_out.contiguous(), group=group)
    return torch.cat(parts, dim=1)

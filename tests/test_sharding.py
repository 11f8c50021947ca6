"""KV-head sharding on CPU: world_size-2 `gloo` run of the host-side
shard logic, with the oracle standing in for the kernels.  Each rank attends over its own
KV-head slice with replicated block tables / seq_lens; the gathered result must equal the
unsharded computation bit for bit (no collective touches the data path)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from tests import helpers as hp
from vllm_main_modify_b200.sharding import (gather_heads, head_shard, shard_paged_cache,
                                            shard_query)


def test_head_shard_layout_matches_reference_tp_rules():
    # Llama-3-70B: 64 q / 8 kv heads (llama.py:124-137 of the reference)
    for tp in (1, 2, 4, 8):
        seen_q, seen_kv = [], []
        for r in range(tp):
            sh = head_shard(64, 8, tp, r)
            assert sh.num_heads == 64 // tp and sh.num_kv_heads == 8 // tp
            # the rank's query heads are exactly the groups of its KV heads
            assert sh.q_lo == sh.kv_lo * 8 and sh.q_hi == sh.kv_hi * 8
            seen_q += list(range(sh.q_lo, sh.q_hi))
            seen_kv += list(range(sh.kv_lo, sh.kv_hi))
        assert seen_q == list(range(64)) and seen_kv == list(range(8))
    # more ranks than KV heads: each KV head replicated on tp / Hkv consecutive ranks
    shards = [head_shard(64, 8, 16, r) for r in range(16)]
    assert all(s.num_kv_heads == 1 and s.num_heads == 4 for s in shards)
    assert [s.kv_lo for s in shards] == [r // 2 for r in range(16)]
    for s in shards:   # every query head still sees its own KV head
        assert all(q // 8 == s.kv_lo for q in range(s.q_lo, s.q_hi))
    with pytest.raises(ValueError):
        head_shard(64, 8, 3, 0)
    with pytest.raises(ValueError):
        head_shard(64, 8, 8, 8)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, kv, ret):
    from oracle import oracle as ora
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        B, H, Hkv, D, BS, NB = 5, 16, 4, 128, 16, 300
        lens = [700, 1, 513, 64, 1200]
        s = D ** -0.5
        # replicated inputs: every rank draws the same full tensors from the same seed
        q, kc, vc, bt, sl = hp.make_paged_inputs(B, H, Hkv, D, BS, NB, lens, torch.bfloat16, seed=4,
                                                 kv=kv, lo=-s, hi=s, permute=True)
        ks, vs = torch.tensor(0.03), torch.tensor(0.05)
        sh = head_shard(H, Hkv, world, rank)
        q_loc = shard_query(q, sh).contiguous()
        kc_loc, vc_loc = shard_paged_cache(kc, vc, sh)
        # new-token write on the shard, then attention on the shard
        key = torch.randn(B, Hkv, D, generator=torch.Generator().manual_seed(9)).to(torch.bfloat16) * 0.05
        slots = torch.tensor([int(bt[b, (L - 1) // BS]) * BS + (L - 1) % BS for b, L in enumerate(lens)])
        ora.reshape_and_cache(key[:, sh.kv_lo:sh.kv_hi].contiguous(), key[:, sh.kv_lo:sh.kv_hi].contiguous(),
                              kc_loc, vc_loc, slots, kv, ks, vs)
        out_loc = torch.empty_like(q_loc)
        ora.paged_attention_v1(out_loc, q_loc, kc_loc, vc_loc, sh.num_kv_heads, s, bt, sl, BS,
                               max(lens), None, kv, ks, vs)
        full = gather_heads(out_loc, world)
        if rank == 0:
            ora.reshape_and_cache(key, key, kc, vc, slots, kv, ks, vs)
            want = torch.empty_like(q)
            ora.paged_attention_v1(want, q, kc, vc, Hkv, s, bt, sl, BS, max(lens), None, kv, ks, vs)
            ret["equal"] = bool(torch.equal(full.view(torch.int16), want.view(torch.int16)))
            ret["shape"] = tuple(full.shape)
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("kv", ["auto", "fp8"])
def test_two_rank_gloo_shards_reproduce_the_unsharded_result(kv):
    from oracle import oracle as ora
    ora.lib()   # build the C oracle once in the parent
    world = 2
    port = _free_port()
    with mp.Manager() as mgr:
        ret = mgr.dict()
        mp.spawn(_worker, args=(world, port, kv, ret), nprocs=world, join=True)
        assert ret["shape"] == (5, 16, 128)
        assert ret["equal"], "gathered per-rank outputs differ from the unsharded computation"


# ------------------------------------------------------------------ decode context parallelism
def _dcp_worker(rank, world, port, ret):
    """Rank r holds pages r, r+N, ... of every sequence; oracle kernels stand in for the CUDA ops,
    the host logic under test is vllm_main_modify_b200.dcp.cp_lse_ag_out_rs."""
    from oracle import lse_oracle
    from vllm_main_modify_b200.dcp import cp_lse_ag_out_rs
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        B, H, Hkv, D, BS = 3, 8, 2, 64, 16
        lens = [16, 640, 1000]           # sequence 0: rank 1 holds no page at all
        NB = sum((l + BS - 1) // BS for l in lens) + 1
        s = D ** -0.5
        q, kc, vc, bt, sl = hp.make_paged_inputs(B, H, Hkv, D, BS, NB, lens, torch.float32, seed=12,
                                                 permute=True)
        rows, lens_r = [], []
        for b, L in enumerate(lens):
            nb = (L + BS - 1) // BS
            mine = list(range(rank, nb, world))
            rows.append([int(bt[b, j]) for j in mine])
            lens_r.append(sum(min(BS, L - j * BS) for j in mine))
        mb = max(1, max(len(x) for x in rows))
        bt_r = torch.tensor([x + [0] * (mb - len(x)) for x in rows], dtype=torch.int32)
        sl_r = torch.tensor(lens_r, dtype=torch.int32)
        one = torch.tensor(1.0)
        part = torch.zeros(B, H, D)
        lse = torch.empty(B, H)
        lse_oracle.paged_attention_lse(part, lse, q, kc, vc, Hkv, s, bt_r, sl_r, BS,
                                       max(max(lens_r), 1), None, "auto", one, one)
        mine = cp_lse_ag_out_rs(part, lse, None, correct_fn=lse_oracle.correct_attn_out)
        want = hp.torch_reference_attention(q, kc, vc, bt, sl, s, Hkv).float()
        h = H // world
        ret[rank] = float((mine - want[:, rank * h:(rank + 1) * h]).abs().max())
        ret[f"shape{rank}"] = tuple(mine.shape)
        dist.barrier()
    finally:
        dist.destroy_process_group()


def test_two_rank_gloo_decode_context_parallel_reproduces_full_attention():
    from oracle import oracle as ora
    ora.lib()
    world = 2
    port = _free_port()
    with mp.Manager() as mgr:
        ret = mgr.dict()
        mp.spawn(_dcp_worker, args=(world, port, ret), nprocs=world, join=True)
        for r in range(world):
            assert ret[f"shape{r}"] == (3, 4, 64)
            assert ret[r] < 1e-5, ret[r]

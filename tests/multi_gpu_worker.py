"""Worker of tests/test_gpu_multi_rank.py: one rank per GPU under torchrun, NCCL.

Checks, on real devices and over real NCCL collectives, what tests/test_sharding.py checks on
CPU/gloo with the oracle standing in for the kernels:

  (1) KV-head sharding (llama.py:124-137 of the reference): every rank runs the CUDA
      decode step on its own KV-head shard; the all-gathered [B,H,D] must be BIT-EQUAL to the same
      kernels run unsharded on rank 0 (no collective touches the data path, and the kernel's
      arithmetic for one KV head does not depend on how many heads the launch has), and within the
      parity gate of the reference's CUDA kernels run unsharded (bench.parity_check);
  (2) decode context parallelism: `dcp.cp_lse_ag_out_rs` (all_gather LSE -> correct_attn_out ->
      NCCL reduce_scatter, the branch a gloo run cannot take) over ranks that each hold every
      world-th page of every sequence, against full attention on rank 0.

Prints one line `MULTI_GPU_OK {...}` on rank 0; any failure raises (non-zero exit).
"""
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def main():
    import bench
    from tests import helpers as hp
    from vllm_main_modify_b200 import _custom_ops as ops
    from vllm_main_modify_b200.dcp import cp_lse_ag_out_rs
    from vllm_main_modify_b200.sharding import gather_heads, head_shard, shard_paged_cache, shard_query

    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    report = {"world": world}

    # ---- (1a) the bench's own check, world ranks: per-rank + gathered vs unsharded reference CUDA
    for cfg in ("cfg2", "cfg5"):
        w = bench.make_shard(cfg, rank, world, dev, copies=1)
        out = torch.empty(w["B"], w["H"] * bench.D, dtype=torch.bfloat16, device=dev)
        res = bench.parity_check(cfg, w, out, True, rank, world, dev)
        assert res["checked"] and res["ok"], res
        assert res["gathered_vs_unsharded"]["ok"], res
        report[f"bench_parity_{cfg}"] = {k: res[k] for k in ("max_abs", "max_rel")}
        del w, out
        torch.cuda.empty_cache()

    # ---- (1b) gather of the shards == our own kernels unsharded, bit for bit
    for kv in ("auto", "fp8"):
        B, H, Hkv, D, BS = 6, 32, 8, 128, 16
        lens = [700, 1, 5000, 64, 1200, 9000]
        NB = sum((L + BS - 1) // BS for L in lens) + 3
        s = D ** -0.5
        q, kc, vc, bt, sl = hp.make_paged_inputs(B, H, Hkv, D, BS, NB, lens, torch.bfloat16, seed=4,
                                                 kv=kv, lo=-s, hi=s, permute=True)
        g = torch.Generator().manual_seed(9)
        key = (torch.randn(B, Hkv, D, generator=g) * 0.05).to(torch.bfloat16)
        val = (torch.randn(B, Hkv, D, generator=g) * 0.05).to(torch.bfloat16)
        slots = torch.tensor([int(bt[b, (L - 1) // BS]) * BS + (L - 1) % BS for b, L in enumerate(lens)])
        ks, vs = torch.tensor(0.03, device=dev), torch.tensor(0.05, device=dev)
        sh = head_shard(H, Hkv, world, rank)
        q_loc = shard_query(q, sh).contiguous().to(dev)
        kc_loc, vc_loc = (t.to(dev) for t in shard_paged_cache(kc, vc, sh))
        out_loc = torch.empty_like(q_loc)
        ops.decode_step(out_loc, None, q_loc, key[:, sh.kv_lo:sh.kv_hi].contiguous().to(dev),
                        val[:, sh.kv_lo:sh.kv_hi].contiguous().to(dev), kc_loc, vc_loc,
                        slots.to(dev), sh.num_kv_heads, s, bt.to(dev), sl.to(dev), BS, max(lens),
                        None, kv, ks, vs)
        full = gather_heads(out_loc, world)
        if rank == 0:
            qd, kcd, vcd = q.to(dev), kc.to(dev), vc.to(dev)
            want = torch.empty_like(qd)
            ops.decode_step(want, None, qd, key.to(dev), val.to(dev), kcd, vcd, slots.to(dev), Hkv,
                            s, bt.to(dev), sl.to(dev), BS, max(lens), None, kv, ks, vs)
            torch.cuda.synchronize()
            assert full.shape == want.shape
            gate = 1e-2 + 1e-2 * want.float().abs()
            assert bool(((full.float() - want.float()).abs() <= gate).all()), f"gather != unsharded ({kv})"
            report[f"gather_{kv}_bit_equal"] = bool(torch.equal(full.view(torch.int16),
                                                                want.view(torch.int16)))
            # the written cache shards, gathered, are the unsharded written cache
            assert torch.equal(kc_loc.cpu(), kcd[:, sh.kv_lo:sh.kv_hi].cpu())
            assert torch.equal(vc_loc.cpu(), vcd[:, sh.kv_lo:sh.kv_hi].cpu())
        dist.barrier()

    # ---- (2) decode context parallelism over NCCL (reduce_scatter_tensor branch)
    B, H, Hkv, D, BS = 5, 16, 4, 128, 16
    lens = [16, 640, 1000, 4097, 33]
    NB = sum((L + BS - 1) // BS for L in lens) + 1
    s = D ** -0.5
    for kv in ("auto", "fp8"):
        q, kc, vc, bt, sl = hp.make_paged_inputs(B, H, Hkv, D, BS, NB, lens, torch.bfloat16, seed=12,
                                                 kv=kv, lo=-s, hi=s, permute=True)
        rows, lens_r = [], []
        for b, L in enumerate(lens):
            nb = (L + BS - 1) // BS
            mine = list(range(rank, nb, world))
            rows.append([int(bt[b, j]) for j in mine])
            lens_r.append(sum(min(BS, L - j * BS) for j in mine))
        mb = max(1, max(len(x) for x in rows))
        bt_r = torch.tensor([x + [0] * (mb - len(x)) for x in rows], dtype=torch.int32, device=dev)
        sl_r = torch.tensor(lens_r, dtype=torch.int32, device=dev)
        ks, vs = torch.tensor(0.5, device=dev), torch.tensor(0.25, device=dev)
        qd, kcd, vcd = q.to(dev), kc.to(dev), vc.to(dev)
        part = torch.zeros_like(qd)
        lse = torch.empty(B, H, dtype=torch.float32, device=dev)
        ops.paged_attention_lse(part, lse, qd, kcd, vcd, Hkv, s, bt_r, sl_r, BS,
                                max(max(lens_r), 1), None, kv, ks, vs)
        mine = cp_lse_ag_out_rs(part, lse, None)          # NCCL all_gather + reduce_scatter
        want = torch.empty_like(qd)
        ops.paged_attention_v1(want, qd, kcd, vcd, Hkv, s, bt.to(dev), sl.to(dev), BS, max(lens),
                               None, kv, ks, vs)
        torch.cuda.synchronize()
        h = H // world
        assert mine.shape == (B, h, D), mine.shape
        ref = want[:, rank * h:(rank + 1) * h].float()
        err = float((mine.float() - ref).abs().max())
        assert bool(((mine.float() - ref).abs() <= 1e-2 + 1e-2 * ref.abs()).all()), (kv, err)
        errs = torch.tensor([err], device=dev)
        dist.all_reduce(errs, op=dist.ReduceOp.MAX)
        report[f"dcp_{kv}_max_abs"] = round(float(errs[0]), 6)
    dist.barrier()
    if rank == 0:
        print("MULTI_GPU_OK " + json.dumps(report), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()

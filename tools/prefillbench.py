#!/usr/bin/env python
"""Timing of ops.prefill_attention (development tool): chunked prefill over the paged cache.

One JSON line per case: device time (CUDA events, median of --iters), achieved TFLOP/s
(4 * D * visible (query, key) pairs per head -- the causal triangle counted exactly), the KV bytes
one pass over each sequence's context amounts to, and -- with --expand -- the per-token expansion
on the decode kernel that the backend used before (O(T x context) KV reads) on the same tensors,
plus the max abs difference between the two outputs.

  python tools/prefillbench.py [--cases chunk2k,...] [--iters 10] [--expand]
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

D, BS = 128, 16
# name: (num_seqs, q_len, ctx_len, H, Hkv, kv)
CASES = {
    "chunk2k_ctx0_bf16": (8, 2048, 0, 32, 8, "auto"),
    "chunk2k_ctx6k_bf16": (8, 2048, 6144, 32, 8, "auto"),
    "chunk2k_ctx6k_fp8": (8, 2048, 6144, 32, 8, "fp8"),
    "chunk512_ctx32k_fp8": (4, 512, 32256, 64, 8, "fp8"),
    "spec4_ctx8k_fp8": (64, 4, 8188, 32, 8, "fp8"),
    "spec8_ctx32k_fp8_gqa8": (32, 8, 32760, 64, 8, "fp8"),
}


def med(fn, iters):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    evs = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        evs.append((a, b))
    torch.cuda.synchronize()
    ts = sorted(a.elapsed_time(b) for a, b in evs)
    return ts[len(ts) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cases", default=",".join(CASES))
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--expand", action="store_true")
    ap.add_argument("--heads-per-tile", type=int, default=0)
    args = ap.parse_args()
    from vllm_main_modify_b200 import _custom_ops as ops
    from vllm_main_modify_b200 import backend as be
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    for name in args.cases.split(","):
        B, ql, ctx, H, Hkv, kv = CASES[name]
        L = ql + ctx
        nb = (L + BS - 1) // BS
        NB = B * nb
        g = torch.Generator(device=dev).manual_seed(0)
        s = D ** -0.5
        q = ((torch.rand(B * ql, H, D, generator=g, device=dev) * 2 - 1)).to(torch.bfloat16)
        if kv == "auto":
            kc = torch.empty(NB, Hkv, D // 8, BS, 8, dtype=torch.bfloat16, device=dev).uniform_(-1, 1, generator=g)
            vc = torch.empty(NB, Hkv, D, BS, dtype=torch.bfloat16, device=dev).uniform_(-1, 1, generator=g)
        else:
            kc = torch.randint(0, 0x48, (NB, Hkv, D // 16, BS, 16), generator=g, device=dev, dtype=torch.uint8)
            vc = torch.randint(0, 0x48, (NB, Hkv, D, BS), generator=g, device=dev, dtype=torch.uint8)
        bt = torch.randperm(NB, generator=torch.Generator().manual_seed(1)).to(torch.int32).view(B, nb).to(dev)
        qsl = (torch.arange(B + 1, dtype=torch.int32) * ql).to(dev)
        sl = torch.full((B,), L, dtype=torch.int32, device=dev)
        one = torch.tensor(0.05 if kv != "auto" else 1.0, device=dev)
        out = torch.empty_like(q)

        def run():
            ops.prefill_attention(out, None, q, kc, vc, s, bt, qsl, sl, ql, None, kv, one, one,
                                  args.heads_per_tile)

        ms = med(run, args.iters)
        pairs = B * H * (ql * ctx + ql * (ql + 1) // 2)
        flops = 4.0 * D * pairs
        ces = 2 if kv == "auto" else 1
        kv_once = B * Hkv * L * D * 2 * ces
        rec = {"case": name, "num_seqs": B, "q_len": ql, "ctx": ctx, "H": H, "Hkv": Hkv, "kv": kv,
               "ms": round(ms, 4), "tflops": round(flops / (ms * 1e-3) / 1e12, 1),
               "kv_bytes_one_pass": kv_once,
               "kv_one_pass_gbs": round(kv_once / (ms * 1e-3) / 1e9, 1)}
        if args.expand:
            meta = be.B200PagedAttentionMetadata(
                num_actual_tokens=B * ql, max_query_len=ql, seq_lens=sl, max_seq_len=L,
                block_table=bt, slot_mapping=torch.zeros(B * ql, dtype=torch.int64, device=dev),
                query_start_loc=qsl)
            bt_tok, sl_tok = be.expand_queries_to_sequences(meta)
            out2 = torch.empty_like(q)

            def run2():
                ops.paged_attention_v1(out2, q, kc, vc, Hkv, s, bt_tok, sl_tok, BS, L, None, kv,
                                       one, one)

            ms2 = med(run2, max(3, args.iters // 3))
            rec["expansion_on_decode_kernel_ms"] = round(ms2, 4)
            rec["speedup_vs_expansion"] = round(ms2 / ms, 2)
            rec["max_abs_diff_vs_expansion"] = float((out.float() - out2.float()).abs().max())
        print(json.dumps(rec), flush=True)
        del q, kc, vc, out
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()

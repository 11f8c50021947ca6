#!/usr/bin/env python
"""Times paged_attention_v2 on shapes outside the five BASELINE configs (development tool):
this repo's stream kernel, its generic kernel and -- with oracle/_ref/libvllm_ref_cuda.so -- the
reference's CUDA kernel on the SAME device tensors.  One JSON line per (shape, impl).

  python tools/shapebench.py [--iters 10]
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
DEV = "cuda:0"

# name, B, H, Hkv, D, ctx, kv
SHAPES = [
    ("gemma_d256_bf16", 64, 16, 8, 256, 4096, "auto"),
    ("gemma_d256_fp8", 64, 16, 8, 256, 8192, "fp8"),
    ("phi_d96_bf16", 64, 32, 32, 96, 2048, "auto"),
    ("d192_fp8", 64, 24, 8, 192, 8192, "fp8"),
    ("d64_fp8_gqa8", 128, 64, 8, 64, 8192, "fp8"),
    ("neox_d80_bf16", 64, 32, 32, 80, 2048, "auto"),
    ("d112_bf16_gqa4", 64, 32, 8, 112, 4096, "auto"),
    ("d80_fp8", 64, 32, 32, 80, 4096, "fp8"),
    ("d112_fp8_gqa4", 64, 32, 8, 112, 8192, "fp8"),
    ("llama8b_e5m2", 256, 32, 8, 128, 8192, "fp8_e5m2"),
    ("d192_bf16", 64, 24, 8, 192, 4096, "auto"),
    ("d96_fp8", 64, 32, 8, 96, 8192, "fp8"),
    ("d64_bf16_gqa8", 128, 64, 8, 64, 4096, "auto"),
    # 32-token cache blocks
    ("llama8b_bs32_bf16", 64, 32, 8, 128, 4096, "auto", 32),
    ("llama8b_bs32_fp8", 256, 32, 8, 128, 8192, "fp8", 32),
    ("gemma_d256_bs32_fp8", 64, 16, 8, 256, 8192, "fp8", 32),
]


def timeit(fn, iters):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    return ts[len(ts) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--only", default="", help="comma-separated shape names (default: all)")
    ap.add_argument("--tune", type=int, default=0, help="launch-shape candidate (test hook)")
    ap.add_argument("--sustain", type=float, default=0.0,
                    help="also run the stream kernel back to back for this many seconds and "
                         "report the rate of the last third (the power-capped regime)")
    ap.add_argument("--stream-only", action="store_true")
    args = ap.parse_args()
    from vllm_main_modify_b200 import _custom_ops as ops
    from vllm_main_modify_b200 import _native
    ref = None
    try:
        from oracle import ref_cuda
        if os.path.exists(ref_cuda.LIB_PATH):
            ref_cuda.load()
            ref = ref_cuda
    except Exception as e:  # noqa: BLE001
        print(json.dumps({"ref_cuda": f"unavailable: {e}"}))
    g = torch.Generator(device=DEV).manual_seed(0)
    only = set(filter(None, args.only.split(",")))
    for name, B, H, Hkv, D, ctx, kv, *rest in SHAPES:
        if only and name not in only:
            continue
        BS = rest[0] if rest else 16
        nblk = ctx // BS
        NB = B * nblk
        if kv == "auto":
            kc = (torch.rand(NB, Hkv, D // 8, BS, 8, device=DEV, generator=g) - 0.5).to(torch.bfloat16)
            vc = (torch.rand(NB, Hkv, D, BS, device=DEV, generator=g) - 0.5).to(torch.bfloat16)
            ces = 2
        else:
            kc = torch.randint(0, 0x48, (NB, Hkv, D // 16, BS, 16), device=DEV, generator=g,
                               dtype=torch.int32).to(torch.uint8)
            vc = torch.randint(0, 0x48, (NB, Hkv, D, BS), device=DEV, generator=g,
                               dtype=torch.int32).to(torch.uint8)
            ces = 1
        q = (torch.rand(B, H, D, device=DEV, generator=g) - 0.5).to(torch.bfloat16)
        bt = torch.randperm(NB, device=DEV, generator=g).to(torch.int32).view(B, nblk)
        sl = torch.full((B,), ctx, dtype=torch.int32, device=DEV)
        one = torch.tensor(1.0, device=DEV)
        P = (ctx + 511) // 512
        es = torch.empty(B, H, P, dtype=torch.float32, device=DEV)
        ml = torch.empty_like(es)
        tmp = torch.empty(B, H, P, D, dtype=q.dtype, device=DEV)
        out = torch.empty_like(q)
        nbytes = B * ctx * Hkv * D * 2 * ces
        s = D ** -0.5

        def ours():
            ops.paged_attention_v2(out, es, ml, tmp, q, kc, vc, Hkv, s, bt, sl, BS, ctx, None, kv,
                                   one, one)

        def theirs():
            ref.paged_attention_v2(out, es, ml, tmp, q, kc, vc, Hkv, s, bt, sl, BS, ctx, None, kv,
                                   one, one, 0, 0, 0, 64, 0)

        outs = {}
        _native.set_stream_tune(args.tune)
        for impl, label in ((_native.IMPL_STREAM, "stream"), (_native.IMPL_GENERIC, "generic")):
            if args.stream_only and label != "stream":
                continue
            _native.set_impl(impl)
            ms = timeit(ours, args.iters)
            outs[label] = out.clone()
            rec = {"shape": name, "B": B, "H": H, "Hkv": Hkv, "D": D, "ctx": ctx,
                   "block_size": BS, "kv": kv, "impl": label, "tune": args.tune,
                   "ms": round(ms, 4), "gbs": round(nbytes / ms / 1e6, 1)}
            if label == "stream" and args.sustain > 0:
                n = max(30, int(args.sustain * 1e3 / ms))
                marks = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
                marks[0].record()
                for k in range(3):
                    for _ in range(n // 3):
                        ours()
                    marks[k + 1].record()
                torch.cuda.synchronize()
                last = marks[2].elapsed_time(marks[3]) / (n // 3)
                first = marks[0].elapsed_time(marks[1]) / (n // 3)
                rec["sustained_first_third_gbs"] = round(nbytes / first / 1e6, 1)
                rec["sustained_last_third_gbs"] = round(nbytes / last / 1e6, 1)
            print(json.dumps(rec), flush=True)
        _native.set_impl(_native.IMPL_AUTO)
        if ref is not None and not args.stream_only:
            ms = timeit(theirs, args.iters)
            err = float((out.float() - outs["stream"].float()).abs().max())
            print(json.dumps({"shape": name, "impl": "reference_cuda", "ms": round(ms, 4),
                              "gbs": round(nbytes / ms / 1e6, 1),
                              "max_abs_diff_vs_stream": round(err, 5)}), flush=True)
        del kc, vc, tmp
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""Development check (GPU): paged_attention_v2 through each kernel family against the
generic SIMT kernel on the same device tensors, over shapes that stress the stream-K
scheduler (ragged, many short sequences, single long sequence, empty sequences) and the
lazy softmax rescale ("hot" logits).  Prints one line per case; exit code 1 on mismatch.

  python tools/dev_stream_check.py [--impls 2]
"""
from __future__ import annotations

import argparse
import os
import random
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from tests import helpers as hp  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--impls", default="2")
    args = ap.parse_args()
    from vllm_main_modify_b200 import _custom_ops as ops
    from vllm_main_modify_b200 import _native

    dev = "cuda:0"
    rnd = random.Random(0)
    cases = []
    # (name, B, H, Hkv, D, lens, dtype, kv, alibi, qmul)
    cases.append(("gqa8 fp8 ragged", 6, 8, 1, 128, [128, 40000, 5000, 777, 16384 + 3, 20001], torch.bfloat16, "fp8", False, 1.0))
    cases.append(("gqa8 fp8 hot", 5, 64, 8, 128, [1000, 17, 512, 640, 3333], torch.bfloat16, "fp8", False, 40.0))
    cases.append(("gqa4 bf16 hot", 5, 32, 8, 128, [1000, 17, 512, 640, 3333], torch.bfloat16, "auto", False, 40.0))
    cases.append(("mha bf16", 8, 32, 32, 128, [1024] * 8, torch.bfloat16, "auto", False, 1.0))
    cases.append(("mha f16 d64 alibi", 7, 40, 40, 64, [rnd.randint(1, 3000) for _ in range(7)], torch.float16, "auto", True, 10.0))
    cases.append(("gqa16 fp8 d64 alibi", 7, 32, 2, 64, [rnd.randint(1, 3000) for _ in range(7)], torch.float16, "fp8", True, 10.0))
    cases.append(("gqa5 bf16", 3, 10, 2, 128, [1, 513, 1025], torch.bfloat16, "auto", False, 5.0))
    cases.append(("many short fp8", 300, 16, 2, 128, [rnd.randint(1, 700) for _ in range(300)], torch.bfloat16, "fp8", False, 20.0))
    cases.append(("with empties fp8", 9, 8, 1, 128, [0, 600, 0, 0, 31, 5000, 0, 512, 0], torch.bfloat16, "fp8", False, 20.0))
    cases.append(("one long bf16", 1, 8, 1, 128, [70001], torch.bfloat16, "auto", False, 20.0))
    cases.append(("one long fp8", 1, 64, 8, 128, [131072], torch.bfloat16, "fp8", False, 20.0))

    bad = 0
    for name, B, H, Hkv, D, lens, dtype, kv, use_alibi, qmul in cases:
        BS = 16
        s = D ** -0.5
        NB = sum((L + BS - 1) // BS for L in lens) + 5
        q, kc, vc, bt, sl = hp.make_paged_inputs(B, H, Hkv, D, BS, NB, lens, dtype, seed=3, kv=kv,
                                                 lo=-1.0, hi=1.0, permute=True)
        q = (q.float() * qmul).to(dtype)
        # k_scale is a power of two: the generic kernel rounds K * k_scale to scalar_t like the
        # reference (quant_utils.cuh:362-372), which is exact then, so peaky ("hot") softmaxes
        # can be compared tightly
        ks, vs = (1.0, 1.0) if kv == "auto" else (0.5, 1.3)
        alibi = None
        if use_alibi:
            alibi = torch.tensor([2.0 ** (-(i + 1) * 8.0 / H) for i in range(H)], dtype=torch.float32, device=dev)
        qd, kcd, vcd, btd, sld = (t.to(dev) for t in (q, kc, vc, bt, sl))
        ksd, vsd = torch.tensor(ks, device=dev), torch.tensor(vs, device=dev)
        max_len = max(max(lens), 1)
        P = (max_len + 511) // 512
        outs = {}
        for impl in [1] + [int(x) for x in args.impls.split(",")]:
            _native.set_impl(impl)
            es = torch.full((B, H, P), float("nan"), dtype=torch.float32, device=dev)
            ml = torch.full_like(es, float("nan"))
            tmp = torch.full((B, H, P, D), float("nan"), dtype=dtype, device=dev)
            out = torch.full_like(qd, float("nan"))
            ops.paged_attention_v2(out, es, ml, tmp, qd, kcd, vcd, Hkv, s, btd, sld, BS, max_len,
                                   alibi, kv, ksd, vsd)
            torch.cuda.synchronize()
            outs[impl] = out.float().cpu()
            if impl == 2:
                out1 = torch.full_like(qd, float("nan"))
                ops.paged_attention_v1(out1, qd, kcd, vcd, Hkv, s, btd, sld, BS, max_len, alibi,
                                       kv, ksd, vsd)
                torch.cuda.synchronize()
                outs["2v1"] = out1.float().cpu()
        _native.set_impl(0)
        live = torch.tensor([L > 0 for L in lens])
        ref = outs[1][live]
        msg = []
        for impl, o in outs.items():
            if impl == 1:
                continue
            o = o[live]
            err = (o - ref).abs()
            tol = 1e-2 + 1e-2 * ref.abs()
            ok = bool((err <= tol).all()) and bool(torch.isfinite(o).all())
            bad += (not ok)
            msg.append(f"impl{impl}: max_abs={err.max().item():.3e} ref_max={ref.abs().max().item():.3f} {'ok' if ok else 'MISMATCH'}")
        print(f"{name:24s} " + " | ".join(msg), flush=True)
    print("FAILED" if bad else "all ok")
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())

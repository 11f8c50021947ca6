#!/usr/bin/env python
"""Times the secondary ops of the path (development tool): cache write, LSE merges, block ops.

For each case: this repo's op and -- where oracle/_ref/libvllm_ref_cuda.so has it -- the
reference's own CUDA op on the SAME device tensors, CUDA events around single launches after
3 warm-ups, median of --iters.  Working sets are sized above the 126 MB L2 where the op's natural
size allows; the small decode-step cases are latency-bound and say so.  Prints one JSON line per
(case, impl) with the algorithmic bytes (bytes that must be read + written) and GB/s.

  python tools/opbench.py [--iters 20] [--no-ref]
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
    return ts[len(ts) // 2], ts[0]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--no-ref", action="store_true")
    args = ap.parse_args()
    from vllm_main_modify_b200 import _custom_ops as ops
    ref = None
    if not args.no_ref:
        from oracle import ref_cuda
        if os.path.exists(ref_cuda.LIB_PATH):
            ref_cuda.load()
            ref = ref_cuda
    peak = 6549.8
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peak = float(json.load(f)["hbm_gbs"])
    except OSError:
        pass
    g = torch.Generator(device=DEV).manual_seed(0)

    def emit(case, impl, nbytes, med, best, note=""):
        gbs = nbytes / (med * 1e-3) / 1e9
        print(json.dumps({"case": case, "impl": impl, "us_median": round(med * 1e3, 2),
                          "us_best": round(best * 1e3, 2), "algorithmic_bytes": int(nbytes),
                          "gbs": round(gbs, 1), "frac_measured_hbm": round(gbs / peak, 4),
                          "note": note}), flush=True)

    def both(case, nbytes, ours, theirs, note=""):
        med, best = timeit(ours, args.iters)
        emit(case, "ours", nbytes, med, best, note)
        if ref is not None and theirs is not None:
            med, best = timeit(theirs, args.iters)
            emit(case, "reference_cuda", nbytes, med, best, note)

    # ---------------------------------------------------------------- cache write
    Hkv, D, BS = 8, 128, 16
    for T, tag in ((128, "decode step, 128 tokens (latency-bound)"),
                   (16384, "16384 tokens, random slots"),
                   (16384, "16384 tokens, prefill: 8 chunks of 2048 consecutive positions from "
                           "position 5, physical blocks permuted")):
        NB = max(T // BS + 64, 2048)
        key = torch.randn(T, Hkv, D, device=DEV, generator=g).to(torch.bfloat16)
        val = torch.randn(T, Hkv, D, device=DEV, generator=g).to(torch.bfloat16)
        if "prefill" in tag:
            per = T // 8
            nb = (5 + per + BS - 1) // BS
            perm = torch.randperm(NB, device=DEV, generator=g)[:8 * nb].view(8, nb)
            pos = torch.arange(5, 5 + per, device=DEV)
            slots = (perm[:, pos // BS] * BS + pos % BS).reshape(-1).to(torch.int64)
        else:
            slots = torch.randperm(NB * BS, device=DEV, generator=g)[:T].to(torch.int64)
        sc = torch.tensor(0.05, device=DEV)
        for kv, cdt, ces in (("fp8", torch.uint8, 1), ("auto", torch.bfloat16, 2)):
            x = 16 // ces
            kc = torch.zeros(NB, Hkv, D // x, BS, x, dtype=cdt, device=DEV)
            vc = torch.zeros(NB, Hkv, D, BS, dtype=cdt, device=DEV)
            nbytes = 2 * T * Hkv * D * (2 + ces)
            both(f"reshape_and_cache {kv} T={T}" + (" prefill" if "prefill" in tag else ""), nbytes,
                 lambda: ops.reshape_and_cache(key, val, kc, vc, slots, kv, sc, sc),
                 (lambda: ref.reshape_and_cache(key, val, kc, vc, slots, kv, sc, sc)) if ref else None, tag)
            kf = torch.zeros(NB, BS, Hkv, D, dtype=cdt, device=DEV)
            vf = torch.zeros_like(kf)
            both(f"reshape_and_cache_flash NHD {kv} T={T}" + (" prefill" if "prefill" in tag else ""), nbytes,
                 lambda: ops.reshape_and_cache_flash(key, val, kf, vf, slots, kv, sc, sc),
                 (lambda: ref.reshape_and_cache_flash(key, val, kf, vf, slots, kv, sc, sc)) if ref else None, tag)
            del kc, vc, kf, vf
    # ---------------------------------------------------------------- LSE merges
    n, h, d = 8192, 64, 128
    po = torch.randn(n, h, d, device=DEV, generator=g).to(torch.bfloat16)
    so = torch.randn(n, h, d, device=DEV, generator=g).to(torch.bfloat16)
    pl = torch.randn(h, n, device=DEV, generator=g)
    sl_ = torch.randn(h, n, device=DEV, generator=g)
    out, ol = torch.empty_like(po), torch.empty_like(pl)
    nbytes = 3 * n * h * d * 2 + 3 * n * h * 4
    both(f"merge_attn_states bf16 n={n} h={h} d={d}", nbytes,
         lambda: ops.merge_attn_states(out, po, pl, so, sl_, ol),
         (lambda: ref.merge_attn_states(out, ol, po, pl, so, sl_)) if ref and hasattr(torch.ops.vllm_ref_cuda, "merge_attn_states") else None)
    B_, N_ = 8192, 8
    o2 = torch.randn(B_, h, d, device=DEV, generator=g).to(torch.bfloat16)
    lses = torch.randn(N_, B_, h, device=DEV, generator=g)
    both(f"correct_attn_out bf16 B={B_} h={h} d={d} ranks={N_}", 2 * B_ * h * d * 2 + (N_ + 1) * B_ * h * 4,
         lambda: ops.correct_attn_out(o2, lses, 3), None,
         "no reference CUDA op (the reference's is a Triton kernel)")
    del po, so, out, o2
    # ---------------------------------------------------------------- gathers (MLA-style cache)
    entry, batch, per = 576, 64, 4096
    NB = batch * per // BS
    for kv, cdt in (("auto", torch.bfloat16), ("fp8", torch.uint8)):
        src = (torch.randn(NB, BS, entry, device=DEV, generator=g) * 0.5).to(torch.bfloat16)
        if cdt == torch.uint8:
            src = torch.randint(0, 0x78, (NB, BS, entry), device=DEV, generator=g, dtype=torch.uint8)
        bt = torch.randperm(NB, device=DEV, generator=g).to(torch.int32).view(batch, per // BS)
        cu = (torch.arange(batch + 1, device=DEV) * per).to(torch.int32)
        dst = torch.empty(batch * per, entry, dtype=torch.bfloat16, device=DEV)
        sc = torch.tensor(0.1, device=DEV)
        nbytes = batch * per * entry * (src.element_size() + 2)
        both(f"gather_and_maybe_dequant_cache {kv} batch={batch} x {per} tok, entry={entry}", nbytes,
             lambda: ops.gather_and_maybe_dequant_cache(src, dst, bt, cu, batch, kv, sc, None),
             (lambda: ref.gather_and_maybe_dequant_cache(src, dst, bt, cu, batch, kv, sc, None))
             if ref and hasattr(torch.ops.vllm_ref_cuda, "cp_gather_cache") else None)
        if kv == "auto":
            both(f"cp_gather_cache bf16 batch={batch} x {per} tok, entry={entry}", nbytes,
                 lambda: ops.cp_gather_cache(src, dst, bt, cu, batch, None),
                 (lambda: ref.cp_gather_cache(src, dst, bt, cu, batch, None))
                 if ref and hasattr(torch.ops.vllm_ref_cuda, "cp_gather_cache") else None)
            # ragged: one long sequence among short ones (the reference splits every sequence
            # into the same number of CTAs)
            lens = [per * batch // 2] + [per // 2 + (i % 7) for i in range(batch - 1)]
            tot = sum(lens)
            cu2 = torch.tensor([0] + list(torch.tensor(lens).cumsum(0)), dtype=torch.int32, device=DEV)
            bt2 = torch.randint(0, NB, (batch, (max(lens) + BS - 1) // BS), device=DEV, generator=g,
                                dtype=torch.int32)
            dst2 = torch.empty(tot, entry, dtype=torch.bfloat16, device=DEV)
            both(f"cp_gather_cache bf16 ragged (1 x {lens[0]} + {batch - 1} x ~{per // 2}), entry={entry}",
                 tot * entry * 4,
                 lambda: ops.cp_gather_cache(src, dst2, bt2, cu2, batch, None),
                 (lambda: ref.cp_gather_cache(src, dst2, bt2, cu2, batch, None))
                 if ref and hasattr(torch.ops.vllm_ref_cuda, "cp_gather_cache") else None)
        del src, dst
    # ---------------------------------------------------------------- copy_blocks
    L, NBc, pairs = 32, 4096, 512
    kcs = [torch.randint(0, 255, (NBc, Hkv, D // 16, BS, 16), device=DEV, generator=g, dtype=torch.uint8)
           for _ in range(L)]
    vcs = [torch.randint(0, 255, (NBc, Hkv, D, BS), device=DEV, generator=g, dtype=torch.uint8)
           for _ in range(L)]
    perm = torch.randperm(NBc, generator=torch.Generator().manual_seed(0))
    mapping = torch.stack([perm[:pairs], perm[pairs:2 * pairs]], dim=1).to(torch.int64).to(DEV)
    nbytes = L * pairs * 2 * (Hkv * D * BS) * 2
    both(f"copy_blocks fp8 layers={L} pairs={pairs} block={Hkv * D * BS} B", nbytes,
         lambda: ops.copy_blocks(kcs, vcs, mapping),
         (lambda: ref.copy_blocks(kcs, vcs, mapping)) if ref and hasattr(torch.ops.vllm_ref_cuda, "copy_blocks") else None,
         "reference: includes its blocking upload of the per-layer pointer arrays")


if __name__ == "__main__":
    main()

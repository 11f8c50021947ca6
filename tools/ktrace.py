#!/usr/bin/env python
"""Phase timeline of the decode (stream) kernel, measured INSIDE the kernel (development tool).

Arms the in-kernel trace (include/b200_paged_attn_testing.h: every CTA stamps %globaltimer at
its phase boundaries), runs the fused decode step of a BASELINE config a few times -- eager,
back to back, one stream -- and prints, for the last launch, when each phase was reached
relative to the first CTA's entry: min / median / max over the CTAs (and warps).  One JSON line
per config; `--dump` adds the raw stamps.

  python tools/ktrace.py --configs cfg1,cfg5 --tp 8
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

SLOTS = 32
NAMES = {0: "entry", 1: "dependency_wait_over", 2: "stage_count_known", 3: "span_located",
         4: "first_copies_issued", 5: "query_loaded", 24: "all_warps_over", 25: "cta_merge_over",
         26: "ticket_taken", 27: "exit"}


def stats(v):
    v = sorted(v)
    return {"min": round(v[0], 2), "med": round(v[len(v) // 2], 2), "max": round(v[-1], 2)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--configs", default="cfg1,cfg5,cfg4")
    ap.add_argument("--tp", type=int, default=1)
    ap.add_argument("--reps", type=int, default=6)
    ap.add_argument("--unfused", action="store_true")
    ap.add_argument("--dump", action="store_true")
    args = ap.parse_args()
    from vllm_main_modify_b200 import _custom_ops as ops
    from vllm_main_modify_b200 import _native
    from vllm_main_modify_b200.paged_attn import PagedAttention
    dev = torch.device("cuda", 0)
    trace = torch.zeros(1 << 16, dtype=torch.int64, device=dev)     # 512 KiB: 2048 CTAs
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    for cfg in args.configs.split(","):
        desc, B, H, Hkv, ctx, kv, ragged = bench.CONFIGS[cfg]
        w = bench.make_shard(cfg, 0, args.tp, dev)
        kc, vc = PagedAttention.split_kv_cache(w["kv_cache"], w["Hkv"], bench.D)
        out = torch.empty_like(w["q"])
        nbytes = bench.kv_bytes(w["lens"], w["Hkv"], kv)

        def step():
            if args.unfused:
                ops.paged_attention_v1(out, w["q"], kc, vc, w["Hkv"], w["scale"], w["bt"], w["sl"],
                                       bench.BS, w["max_len"], None, kv, w["ks"], w["vs"])
            else:
                ops.decode_step(out, None, w["q"], w["k_new"], w["v_new"], kc, vc, w["slots"],
                                w["Hkv"], w["scale"], w["bt"], w["sl"], bench.BS, w["max_len"],
                                None, kv, w["ks"], w["vs"])

        # the phase stamps live in the per-warp stream kernel (tune >= 100 selects it where the
        # whole-block kernel would otherwise take the call)
        _native.set_stream_tune(100)
        for _ in range(3):
            step()
        torch.cuda.synchronize()
        if nbytes < (512 << 20):
            flush.add_(1)
        trace.zero_()
        torch.cuda.synchronize()
        _native.set_trace(trace.data_ptr(), trace.numel() * 8)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(args.reps):
            step()
        b.record()
        torch.cuda.synchronize()
        _native.set_trace(0, 0)
        t = trace.cpu().view(-1, SLOTS)
        used = (t[:, 0] != 0).nonzero().flatten()
        t = t[used].double()
        n_cta = t.shape[0]
        t0 = float(t[:, 0].min())
        rel = (t - t0) / 1e3                                # microseconds
        line = {"config": cfg, "tp": args.tp, "ctas": n_cta, "kv_gib": round(nbytes / 2**30, 3),
                "us_per_launch_events": round(a.elapsed_time(b) * 1e3 / args.reps, 2),
                "ideal_us_at_measured_peak": round(nbytes / 6454.6e9 * 1e6, 2), "phases_us": {}}
        for k, name in NAMES.items():
            col = rel[:, k]
            col = col[t[:, k] != 0]
            if len(col):
                line["phases_us"][name] = stats(col.tolist())
        for base, name in ((8, "first_stage_landed_per_warp"), (16, "main_loop_over_per_warp")):
            blk = rel[:, base:base + 8]
            m = t[:, base:base + 8] != 0
            if m.any():
                line["phases_us"][name] = stats(blk[m].tolist())
        # per-CTA durations of the tail phases
        def dur(k1, k0):
            m = (t[:, k1] != 0) & (t[:, k0] != 0)
            return stats(((t[:, k1] - t[:, k0])[m] / 1e3).tolist()) if m.any() else None
        line["durations_us"] = {"prologue(entry->first stage of warp 0)": dur(8, 0),
                                "cta_merge(all warps over->merge over)": dur(25, 24),
                                "ticket": dur(26, 25), "cross_cta_merge": dur(27, 26),
                                "whole_cta": dur(27, 0)}
        line["timer_resolution_ns"] = int(torch.diff(torch.unique(t.flatten())).min().item())
        if args.dump:
            line["raw_us"] = [[round(x, 2) for x in row] for row in rel.tolist()]
        print(json.dumps(line), flush=True)
        del w, kc, vc, out
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()

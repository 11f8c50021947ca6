#!/usr/bin/env python
"""Kernel-level sweep over the five BASELINE configs (development tool).

Times ops.paged_attention_v1 / _v2 and ops.reshape_and_cache directly with CUDA
events on the current stream (KV per call >> L2 for cfg2-5; cfg1 gets an L2 flush
between calls) and prints one JSON line per (config, op).

  python tools/kbench.py [--configs cfg1,cfg2,...] [--iters 20] [--tp 1]
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
# A/B runs on one box: B200PA_PKG_ROOT=<dir holding another built copy of the package>
if os.environ.get("B200PA_PKG_ROOT"):
    sys.path.insert(0, os.environ["B200PA_PKG_ROOT"])

import bench  # noqa: E402


def time_op(fn, iters, flush=None):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    evs = []
    for _ in range(iters):
        if flush is not None:
            flush.add_(1)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        evs.append((a, b))
    torch.cuda.synchronize()
    series = [a.elapsed_time(b) for a, b in evs]
    ts = sorted(series)
    time_op.last_series = series
    return ts[len(ts) // 2], ts[0]


def graphed(fn):
    """Capture fn into a CUDA graph (side stream, after a warm-up) and return the replay callable:
    what a full-graph decode capture of the runner (gpu_model_runner.py:3468) replays per step."""
    stream = torch.cuda.Stream()
    stream.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(stream):
        fn()
        stream.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=stream):
            fn()
    torch.cuda.current_stream().wait_stream(stream)
    torch.cuda.synchronize()
    return graph.replay


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--configs", default="cfg1,cfg2,cfg3,cfg4,cfg5")
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--tp", type=int, default=1, help="emulate rank 0 of a tp-way KV-head shard")
    ap.add_argument("--impl", type=int, default=0, help="b200pa_set_impl value (0 auto)")
    ap.add_argument("--ops", default="v1,v2,reshape_and_cache")
    ap.add_argument("--series", action="store_true", help="also print the per-launch times in order")
    ap.add_argument("--graph", action="store_true",
                    help="time CUDA-graph replays of each op instead of eager launches")
    ap.add_argument("--sustain", type=float, default=0.0,
                    help="also run each op back to back for this many seconds and report the "
                         "rate of the first and the last third (the power-capped regime)")
    ap.add_argument("--ref-cuda", action="store_true",
                    help="time the reference's own CUDA kernels (oracle/_ref/libvllm_ref_cuda.so) "
                         "on the same tensors instead of this repo's")
    args = ap.parse_args()
    from vllm_main_modify_b200 import _custom_ops as ops
    from vllm_main_modify_b200 import _native
    _native.set_impl(args.impl)
    from vllm_main_modify_b200.paged_attn import PagedAttention
    if args.ref_cuda:
        from oracle import ref_cuda
        ref_cuda.load()
        ops = ref_cuda   # same positional signatures (torch.ops.vllm_ref_cuda.*)
    tail = (0, 0, 0, 64, 0) if args.ref_cuda else ()

    dev = torch.device("cuda", 0)
    peak = 6524.6
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peak = float(json.load(f)["hbm_gbs"])
    except OSError:
        pass
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    for cfg in args.configs.split(","):
        desc, B, H, Hkv, ctx, kv, ragged = bench.CONFIGS[cfg]
        w = bench.make_shard(cfg, 0, args.tp, dev)
        kc, vc = PagedAttention.split_kv_cache(w["kv_cache"], w["Hkv"], bench.D)
        nbytes = bench.kv_bytes(w["lens"], w["Hkv"], kv)
        out = torch.empty_like(w["q"])
        P = (w["max_len"] + 511) // 512
        es = torch.empty(B, w["H"], P, dtype=torch.float32, device=dev)
        ml = torch.empty_like(es)
        tmp = torch.empty(B, w["H"], P, bench.D, dtype=out.dtype, device=dev)
        fl = flush if nbytes < (512 << 20) else None

        def v1():
            ops.paged_attention_v1(out, w["q"], kc, vc, w["Hkv"], w["scale"], w["bt"], w["sl"],
                                   bench.BS, w["max_len"], None, kv, w["ks"], w["vs"], *tail)

        def v2():
            ops.paged_attention_v2(out, es, ml, tmp, w["q"], kc, vc, w["Hkv"], w["scale"], w["bt"],
                                   w["sl"], bench.BS, w["max_len"], None, kv, w["ks"], w["vs"], *tail)

        def wr():
            ops.reshape_and_cache(w["k_new"], w["v_new"], kc, vc, w["slots"], kv, w["ks"], w["vs"])

        def step():      # one decode step of one layer: write the new token, then attend
            wr()
            v2()

        def fstep():     # the same step in one launch (not available from the reference)
            ops.decode_step(out, None, w["q"], w["k_new"], w["v_new"], kc, vc, w["slots"], w["Hkv"],
                            w["scale"], w["bt"], w["sl"], bench.BS, w["max_len"], None, kv,
                            w["ks"], w["vs"])

        # token-major views of the same memory (timing only: the bytes mean other tokens there)
        NBk = kc.shape[0]
        kf_nhd = kc.reshape(NBk, -1).view(NBk, bench.BS, w["Hkv"], bench.D)
        vf_nhd = vc.reshape(NBk, -1).view(NBk, bench.BS, w["Hkv"], bench.D)
        kf_hnd = kc.reshape(NBk, -1).view(NBk, w["Hkv"], bench.BS, bench.D).permute(0, 2, 1, 3)
        vf_hnd = vc.reshape(NBk, -1).view(NBk, w["Hkv"], bench.BS, bench.D).permute(0, 2, 1, 3)

        def flash_nhd():
            ops.paged_attention_flash(out, None, w["q"], None, None, kf_nhd, vf_nhd, None,
                                      w["scale"], w["bt"], w["sl"], w["max_len"], None, kv,
                                      w["ks"], w["vs"])

        def flash_hnd():
            ops.paged_attention_flash(out, None, w["q"], None, None, kf_hnd, vf_hnd, None,
                                      w["scale"], w["bt"], w["sl"], w["max_len"], None, kv,
                                      w["ks"], w["vs"])

        def fflash_nhd():    # cache write fused
            ops.paged_attention_flash(out, None, w["q"], w["k_new"], w["v_new"], kf_nhd, vf_nhd,
                                      w["slots"], w["scale"], w["bt"], w["sl"], w["max_len"], None,
                                      kv, w["ks"], w["vs"])

        fns = {"v1": v1, "v2": v2, "reshape_and_cache": wr, "step": step, "fstep": fstep,
               "flash_nhd": flash_nhd, "flash_hnd": flash_hnd, "fflash_nhd": fflash_nhd}
        # ops run in the order given (repeats allowed: "v2,step,v2" separates op from order effects)
        for name in args.ops.split(","):
            fn = fns[name]
            if args.ref_cuda and name == "v1" and w["max_len"] * 4 > 200 * 1024:
                continue   # the reference's v1 keeps max_seq_len fp32 logits in shared memory
            if args.graph:
                fn = graphed(fn)
            med, best = time_op(fn, args.iters, fl)
            line = {"config": cfg, "tp": args.tp,
                    "impl": "reference_cuda" if args.ref_cuda else args.impl,
                    "tune": os.environ.get("B200PA_STREAM_TUNE", "0"), "op": name,
                    "launch": "graph" if args.graph else "eager",
                    "build": os.environ.get("B200PA_AB_LABEL", "tree"),
                    "ms_median": round(med, 4),
                    "ms_best": round(best, 4)}
            if name != "reshape_and_cache":
                gbs = nbytes / (med * 1e-3) / 1e9
                line.update(kv_gib=round(nbytes / 2**30, 3), gbs=round(gbs, 1),
                            frac_measured=round(gbs / peak, 4), pct_8tb=round(gbs / 80.0, 2),
                            heuristic_v1=PagedAttention.use_v1(w["max_len"], B, w["H"]))
            if args.sustain > 0 and name != "reshape_and_cache":
                n = max(30, int(args.sustain * 1e3 / med))
                marks = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
                marks[0].record()
                for k in range(3):
                    for _ in range(n // 3):
                        fn()
                    marks[k + 1].record()
                torch.cuda.synchronize()
                first = marks[0].elapsed_time(marks[1]) / (n // 3)
                last = marks[2].elapsed_time(marks[3]) / (n // 3)
                line["sustained_first_third_gbs"] = round(nbytes / (first * 1e-3) / 1e9, 1)
                line["sustained_last_third_gbs"] = round(nbytes / (last * 1e-3) / 1e9, 1)
            if args.series:
                line["series_us"] = [round(t * 1e3, 1) for t in time_op.last_series]
            print(json.dumps(line), flush=True)
        del w, kc, vc, tmp, es, ml, out
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""Sustained rate, power and clocks of the fused decode step (development tool).

Runs bench.py's sustained leg (the timed step back to back for --seconds, nvidia-smi sampled every
100 ms) on one config and prints one JSON line: GB/s of the first 300 ms, of the last third, the
power and SM clock they settled at.  With B200PA_PKG_ROOT=<dir of an ablated build>
(tools/ablate.sh) the same run shows what one ingredient of the main loop costs in time and watts.

  python tools/powerprobe.py --config cfg4 --seconds 4
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if os.environ.get("B200PA_PKG_ROOT"):
    sys.path.insert(0, os.environ["B200PA_PKG_ROOT"])
sys.path.insert(1 if os.environ.get("B200PA_PKG_ROOT") else 0, ROOT)
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="cfg4")
    ap.add_argument("--tp", type=int, default=1)
    ap.add_argument("--seconds", type=float, default=4.0)
    ap.add_argument("--tune", type=int, default=0,
                    help="test hook: 20-24 launch shapes of the whole-block kernel, 100 + i the "
                         "per-warp stream kernel with its launch shape i")
    ap.add_argument("--unfused", action="store_true")
    args = ap.parse_args()
    import vllm_main_modify_b200
    from vllm_main_modify_b200 import _native
    _native.load()
    if args.tune:
        _native.set_stream_tune(args.tune)
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    cfg = args.config
    w = bench.make_shard(cfg, 0, args.tp, dev, copies=bench.cache_copies(cfg, args.tp))
    out = torch.empty(w["B"], w["H"] * bench.D, dtype=torch.bfloat16, device=dev)
    nbytes = bench.kv_bytes(w["lens"], w["Hkv"], w["kv"])
    fused = not args.unfused
    for _ in range(5):
        at = bench.decode_op(w, out, fused, 0)
        if at is not None:
            at()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(10):
        at = bench.decode_op(w, out, fused, 0)
        if at is not None:
            at()
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / 10
    per, total_ms, steps, chunk, sens = bench.sustained_leg(w, out, fused, args.seconds, ms, nbytes,
                                                            0, torch.cuda.synchronize)
    gbs = [nbytes / (t * 1e-3) / 1e9 for t in per]
    third = max(1, len(gbs) // 3)
    d = sens.get("drift") or {}
    print(json.dumps({
        "pkg": os.path.dirname(vllm_main_modify_b200.__file__),
        "label": os.environ.get("B200PA_AB_LABEL", "tree"), "config": cfg, "tp": args.tp,
        "tune": args.tune, "fused": fused,
        "burst_gbs_10_steps": round(nbytes / (ms * 1e-3) / 1e9, 1),
        "first_300ms_gbs": round(sum(gbs[:3]) / 3, 1),
        "last_third_gbs": round(sum(gbs[-third:]) / third, 1),
        "power_w_last_third": (d.get("power_w") or {}).get("last"),
        "power_w_max": (d.get("power_w") or {}).get("max"),
        "sm_mhz_last_third": (d.get("sm_mhz") or {}).get("last"),
        "temp_hbm_last_third": (d.get("temp_hbm_c") or {}).get("last"),
        "reasons": sens.get("reasons"), "seconds": round(total_ms / 1e3, 2),
        "series_gbs": [round(g) for g in gbs]}), flush=True)


if __name__ == "__main__":
    main()

#!/usr/bin/env python
"""bench.py -- paged-decode hot path benchmark (BASELINE.json metric).

One "step" = one decode step of ONE layer over one batch: write the new token's
K/V into the paged cache (reshape_and_cache) and attend over the whole context
(paged_attention_v2), through the public Python API of this package.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--config cfg4] [--impl ours|reference]

For N > 1 it is launched under torchrun (one rank per GPU); the KV heads are
sharded across ranks exactly like the reference's tensor-parallel layout
(vllm/model_executor/models/llama.py:124-137) and there is NO collective on the
attention path.  Rank 0 prints ONE JSON line on stdout.

`--impl reference` times the reference's own CPU kernels (oracle/_ref, built from
csrc/cpu/attention.cpp + cache.cpp, unmodified) on the host cores, on a bounded
sample of the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

D, BS = 128, 16
HBM_NOMINAL_GBPS = 8000.0   # the "~8 TB/s" BASELINE.json's metric is quoted against

# name -> (description, B, H, Hkv, ctx, kv, ragged)
CONFIGS = {
    "cfg1": ("Llama-2-7B shape MHA 32q/32kv, D=128, block 16, batch 8, ctx 1024, bf16 KV",
             8, 32, 32, 1024, "auto", False),
    "cfg2": ("Llama-3-8B shape GQA 32q/8kv, D=128, block 16, batch 64, ctx 4096, bf16 KV",
             64, 32, 8, 4096, "auto", False),
    "cfg3": ("Llama-3-8B shape GQA 32q/8kv, D=128, block 16, batch 256, ctx 8192, FP8-e4m3 KV",
             256, 32, 8, 8192, "fp8", False),
    "cfg4": ("Llama-3-70B shape GQA 64q/8kv, D=128, block 16, batch 128, ctx 32768, FP8-e4m3 KV, "
             "sharded by KV head", 128, 64, 8, 32768, "fp8", False),
    "cfg5": ("Qwen2-72B shape GQA 64q/8kv, D=128, block 16, batch 64 ragged ctx 128..131072 "
             "(log-uniform), permuted block tables, FP8-e4m3 KV, sharded by KV head",
             64, 64, 8, 131072, "fp8", True),
}


def seq_lens_of(cfg):
    _, B, _, _, ctx, _, ragged = CONFIGS[cfg]
    if not ragged:
        return [ctx] * B
    import math
    import random
    rnd = random.Random(0)
    lens = [int(round(math.exp(rnd.uniform(math.log(128), math.log(131072))))) for _ in range(B)]
    lens[0], lens[1] = 128, 131072
    return lens


def kv_bytes(lens, hkv, kv):
    es = 1 if kv != "auto" else 2
    return sum(lens) * hkv * D * 2 * es


# ----------------------------------------------------------------------------
# clocks sampler (nvidia-smi during the timed region)
# ----------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, smax, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            p = [x.strip() for x in ln.split(",")]
            if len(p) < 6:
                continue
            try:
                sm.append(float(p[0]))
                smax = float(p[1])
            except ValueError:
                continue
            for n, v in zip(names, p[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": smax,
                "reasons": sorted(reasons), "samples": len(sm)}


# ----------------------------------------------------------------------------
# workload
# ----------------------------------------------------------------------------
def make_shard(cfg, rank, world, device):
    """Seeded synthetic tensors of this rank's KV-head shard, resident in HBM."""
    _, B, H, Hkv, ctx, kv, _ = CONFIGS[cfg]
    assert Hkv % world == 0, "tp must divide the KV heads"
    h_loc, hkv_loc = H // world, Hkv // world
    lens = seq_lens_of(cfg)
    nblk = [(L + BS - 1) // BS for L in lens]
    NB = sum(nblk)
    maxblk = max(nblk)
    s = D ** -0.5
    g = torch.Generator(device=device).manual_seed(1234 + rank)
    gc = torch.Generator().manual_seed(0)            # replicated metadata: same on all ranks
    q = ((torch.rand(B, h_loc, D, generator=g, device=device) * 2 - 1) * s).to(torch.bfloat16)
    k_new = ((torch.rand(B, hkv_loc, D, generator=g, device=device) * 2 - 1) * s).to(torch.bfloat16)
    v_new = ((torch.rand(B, hkv_loc, D, generator=g, device=device) * 2 - 1) * s).to(torch.bfloat16)
    if kv == "auto":
        kv_cache = torch.empty(2, NB, BS * hkv_loc * D, dtype=torch.bfloat16, device=device)
        kv_cache.uniform_(-s, s, generator=g)
    else:
        # finite e4m3 codes (no 0x7f / 0xff NaN), both signs
        kv_cache = torch.empty(2, NB, BS * hkv_loc * D, dtype=torch.uint8, device=device)
        for half in range(2):
            for c in range(0, NB, 1 << 15):
                sl_ = kv_cache[half, c:c + (1 << 15)]
                sl_.copy_(torch.randint(0, 0x78, sl_.shape, generator=g, device=device,
                                        dtype=torch.uint8))
                sl_ |= (torch.randint(0, 2, sl_.shape, generator=g, device=device,
                                      dtype=torch.uint8) << 7)
    perm = torch.randperm(NB, generator=gc).to(torch.int32)
    bt = torch.zeros(B, maxblk, dtype=torch.int32)
    pos = 0
    for b, n in enumerate(nblk):
        bt[b, :n] = perm[pos:pos + n]
        pos += n
    sl = torch.tensor(lens, dtype=torch.int32)
    last = torch.tensor([(L - 1) for L in lens])
    slots = (bt[torch.arange(B), (last // BS)].long() * BS + (last % BS)).to(torch.int64)
    ks = torch.tensor(s / 1.7 if kv != "auto" else 1.0, dtype=torch.float32, device=device)
    vs = torch.tensor(s / 2.9 if kv != "auto" else 1.0, dtype=torch.float32, device=device)
    return dict(B=B, H=h_loc, Hkv=hkv_loc, kv=kv, lens=lens, max_len=max(lens), scale=s,
                q=q, k_new=k_new, v_new=v_new, kv_cache=kv_cache, bt=bt.to(device),
                sl=sl.to(device), slots=slots.to(device), ks=ks, vs=vs,
                bt_host=bt, sl_host=sl, slots_host=slots)


class Layer:
    """AttentionLayer protocol (vllm/attention/backends/abstract.py:92-110)."""

    def __init__(self, ks, vs):
        self._k_scale, self._v_scale = ks, vs


def run_ours(args):
    import torch.distributed as dist
    from vllm_main_modify_b200 import _native
    from vllm_main_modify_b200.backend import B200PagedAttentionBackend, B200PagedAttentionMetadata
    from vllm_main_modify_b200.paged_attn import PagedAttention

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (this package has no CPU path)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    assert world == args.gpus, f"--gpus {args.gpus} but WORLD_SIZE={world}"

    cfg = args.config
    desc, B, H, Hkv, ctx, kv, _ = CONFIGS[cfg]
    w = make_shard(cfg, rank, world, dev)
    impl = B200PagedAttentionBackend.get_impl_cls()(w["H"], D, w["scale"], w["Hkv"], None, None, kv)
    layer = Layer(w["ks"], w["vs"])
    meta = B200PagedAttentionMetadata(num_actual_tokens=B, max_query_len=1, seq_lens=w["sl"],
                                      max_seq_len=w["max_len"], block_table=w["bt"],
                                      slot_mapping=w["slots"])
    q2 = w["q"].view(B, -1)
    k2, v2 = w["k_new"].view(B, -1), w["v_new"].view(B, -1)
    out = torch.empty_like(q2)
    key_cache, value_cache = PagedAttention.split_kv_cache(w["kv_cache"], w["Hkv"], D)

    def step():
        impl.forward(layer, q2, k2, v2, w["kv_cache"], meta, output=out)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step()
    barrier()

    # ---- timed region 1: device-resident inputs ("value") -------------------
    sampler = ClockSampler(local_rank)
    sampler.start()
    time.sleep(0.3)
    K = args.steps
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a_ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
            for _ in range(K)]
    launches0 = _native.launch_count()
    barrier()
    ev0.record()
    fused = not args.unfused
    impl.fused_decode_step = fused
    for i in range(K):
        # the step, with events around the attention op (the dominant kernel)
        if fused:
            # one launch: the decode kernel also quantises + stores the new token (decode_step)
            a_ev[i][0].record()
            PagedAttention.write_and_decode(w["q"], w["k_new"], w["v_new"], key_cache, value_cache,
                                            w["slots"], w["bt"], w["sl"], w["max_len"], kv,
                                            w["Hkv"], w["scale"], None, w["ks"], w["vs"],
                                            output=out.view(B, w["H"], D))
            a_ev[i][1].record()
            continue
        PagedAttention.write_to_paged_cache(w["k_new"], w["v_new"], key_cache, value_cache,
                                            w["slots"], kv, w["ks"], w["vs"])
        a_ev[i][0].record()
        PagedAttention.forward_decode(w["q"], key_cache, value_cache, w["bt"], w["sl"],
                                      w["max_len"], kv, w["Hkv"], w["scale"], None, w["ks"],
                                      w["vs"], output=out.view(B, w["H"], D))
        a_ev[i][1].record()
    ev1.record()
    barrier()
    launches = _native.launch_count() - launches0
    ms_total = ev0.elapsed_time(ev1)
    attn_ms = sum(a.elapsed_time(b) for a, b in a_ev) / K
    clocks = sampler.stop()

    # ---- timed region 2: end to end through AttentionImpl.forward with HOST buffers
    # One pinned staging buffer on the host and its twin on the device hold every per-step input
    # (q | k | v | block_table | seq_lens | slot_mapping, 256-byte aligned segments): ONE
    # host->device copy per step instead of six (what a runner's input-batch staging does).
    parts = [("q", q2), ("k", k2), ("v", v2), ("bt", w["bt"]), ("sl", w["sl"]), ("slots", w["slots"])]
    offs, total = {}, 0
    for name, t in parts:
        offs[name] = total
        total += (t.numel() * t.element_size() + 255) // 256 * 256
    hbuf = torch.empty(total, dtype=torch.uint8).pin_memory()
    dbuf = torch.empty(total, dtype=torch.uint8, device=dev)

    def seg(buf, name, like):
        n = like.numel() * like.element_size()
        return buf[offs[name]:offs[name] + n].view(like.dtype).view(like.shape)

    host_src = {"q": q2, "k": k2, "v": v2, "bt": w["bt_host"], "sl": w["sl_host"],
                "slots": w["slots_host"]}
    for name, t in parts:
        seg(hbuf, name, t).copy_(host_src[name].detach().cpu())
    dq, dk, dv = seg(dbuf, "q", q2), seg(dbuf, "k", k2), seg(dbuf, "v", v2)
    dbt, dsl, dslots = seg(dbuf, "bt", w["bt"]), seg(dbuf, "sl", w["sl"]), seg(dbuf, "slots", w["slots"])
    hout = torch.empty(out.shape, dtype=out.dtype).pin_memory()
    meta2 = B200PagedAttentionMetadata(num_actual_tokens=B, max_query_len=1, seq_lens=dsl,
                                       max_seq_len=w["max_len"], block_table=dbt,
                                       slot_mapping=dslots)
    h2d = total   # bytes actually copied (segments padded to 256 B)
    d2h = hout.numel() * hout.element_size()

    def e2e_step():
        dbuf.copy_(hbuf, non_blocking=True)
        impl.forward(layer, dq, dk, dv, w["kv_cache"], meta2, output=out)
        hout.copy_(out, non_blocking=True)
        torch.cuda.current_stream().synchronize()     # the caller reads the result

    for _ in range(3):
        e2e_step()
    if os.environ.get("B200PA_E2E_BREAKDOWN") == "1" and rank == 0:
        # development aid: host time of each part of the eager step (median of 30), to stderr
        parts_us = {"h2d_enqueue": [], "forward": [], "d2h_enqueue": [], "sync": []}
        for _ in range(30):
            t0 = time.perf_counter()
            dbuf.copy_(hbuf, non_blocking=True)
            t1 = time.perf_counter()
            impl.forward(layer, dq, dk, dv, w["kv_cache"], meta2, output=out)
            t2 = time.perf_counter()
            hout.copy_(out, non_blocking=True)
            t3 = time.perf_counter()
            torch.cuda.current_stream().synchronize()
            t4 = time.perf_counter()
            for name, dt in zip(parts_us, (t1 - t0, t2 - t1, t3 - t2, t4 - t3)):
                parts_us[name].append(dt * 1e6)
        sys.stderr.write("[bench] e2e host time per part, us (median): " + json.dumps(
            {k: round(sorted(v)[len(v) // 2], 1) for k, v in parts_us.items()}) + "\n")
        # ... and the device-side segments of the same step (CUDA events between the parts)
        seg_us = {"h2d": [], "forward": [], "d2h": []}
        for _ in range(30):
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
            ev[0].record()
            dbuf.copy_(hbuf, non_blocking=True)
            ev[1].record()
            impl.forward(layer, dq, dk, dv, w["kv_cache"], meta2, output=out)
            ev[2].record()
            hout.copy_(out, non_blocking=True)
            ev[3].record()
            torch.cuda.current_stream().synchronize()
            for name, i in zip(seg_us, range(3)):
                seg_us[name].append(ev[i].elapsed_time(ev[i + 1]) * 1e3)
        sys.stderr.write("[bench] e2e device time per segment, us (median): " + json.dumps(
            {k: round(sorted(v)[len(v) // 2], 1) for k, v in seg_us.items()}) + "\n")
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(K):
        e2e_step()
    e1.record()
    barrier()
    ms_e2e = e0.elapsed_time(e1)

    # The same step replayed from ONE CUDA graph (H2D memcpy node -> AttentionImpl.forward ->
    # D2H memcpy node), the way the reference's runner replays a full-graph decode capture
    # (gpu_model_runner.py:3468): what is left of the host's share when no Python runs per step.
    # Reported beside the eager figure, which stays e2e.value.
    ms_e2e_graph = float("nan")
    try:
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=side):
                dbuf.copy_(hbuf, non_blocking=True)
                impl.forward(layer, dq, dk, dv, w["kv_cache"], meta2, output=out)
                hout.copy_(out, non_blocking=True)
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()

        def e2e_graph_step():
            graph.replay()
            torch.cuda.current_stream().synchronize()

        for _ in range(3):
            e2e_graph_step()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(K):
            e2e_graph_step()
        e1.record()
        barrier()
        ms_e2e_graph = e0.elapsed_time(e1)
        # the replay must order H2D -> forward -> D2H like the eager step: change the host-side
        # query, replay, and compare the host-side output with an eager step on the same input
        hq = seg(hbuf, "q", q2)
        saved = hq.clone()
        hq.copy_(-saved)
        e2e_graph_step()
        got = hout.clone()
        e2e_step()
        if not torch.equal(got.view(torch.uint8), hout.view(torch.uint8)):
            raise RuntimeError("graph replay and eager step disagree on a changed input")
        hq.copy_(saved)
        e2e_step()
    except Exception as exc:   # the eager figure stands; say why there is no replay figure
        sys.stderr.write(f"[bench] e2e graph replay dropped: {exc}\n")
        if ms_e2e_graph != ms_e2e_graph:   # failed before the timed loop: keep the ranks in step
            barrier()
            barrier()
        ms_e2e_graph = float("nan")

    # ---- reduce over ranks ---------------------------------------------------
    my_bytes = kv_bytes(w["lens"], w["Hkv"], kv)
    stats = torch.tensor([ms_total, ms_e2e, attn_ms, ms_e2e_graph], dtype=torch.float64, device=dev)
    tot = torch.tensor([float(my_bytes), float(launches)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(stats, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
        # bench-only output all-gather: [B, H/n, D] -> [B, H, D]
        gathered = torch.empty(world, B, w["H"], D, dtype=out.dtype, device=dev)
        g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        dist.all_gather_into_tensor(gathered, out.view(B, w["H"], D))
        g0.record()
        dist.all_gather_into_tensor(gathered, out.view(B, w["H"], D))
        g1.record()
        torch.cuda.synchronize()
        allgather_ms = g0.elapsed_time(g1)
    else:
        allgather_ms = None
    ms_total, ms_e2e, attn_ms, ms_e2e_graph = [float(x) for x in stats.tolist()]
    total_bytes, total_launches = tot.tolist()

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu = cpu_reference_sample(cfg, budget_s=args.cpu_budget)

    if rank == 0:
        peaks = {}
        try:
            with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
                peaks = json.load(f)
        except OSError:
            pass
        # DRAM traffic of the dominant kernel per launch, from the committed ncu --set full
        # capture of this exact workload (profiles/): only quoted for the config it was taken on
        traffic = None
        try:
            with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
                tr = json.load(f)
            if tr.get("config") == cfg and tr.get("tp") == world:
                traffic = int(tr["dram_bytes_read"]) + int(tr["dram_bytes_write"])
        except (OSError, ValueError, KeyError):
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"
        ms_step = ms_total / K
        value = total_bytes / (ms_step * 1e-3) / 1e9
        e2e_val = total_bytes / (ms_e2e / K * 1e-3) / 1e9
        # roofline of the dominant kernel, per GPU: this rank's KV bytes / attention time
        achieved = (total_bytes / world) / (attn_ms * 1e-3) / 1e9
        line = {
            "metric": "paged_decode_attn_kv_bandwidth", "value": round(value, 1), "unit": "GB/s",
            "n_gpus": world, "steps": K, "warmup": max(args.warmup, 3),
            "ms_per_step": round(ms_step, 4), "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "fp8_e4m3 KV, bf16 q/out, f32 accumulate" if kv != "auto"
            else "bf16 KV/q/out, f32 accumulate", "data": "synthetic",
            "config": {"workload": f"{cfg}: {desc}", "tp": world, "batch": B,
                       "kv_bytes_total": int(total_bytes),
                       "step": ("decode_step: new token K/V quantised + stored and the whole "
                                "context attended in one launch (= reshape_and_cache + "
                                "paged_attention)") if fused else
                               "reshape_and_cache (new token K/V) + paged_attention_v2",
                       "l2": "inputs larger than L2: KV read per GPU per step = "
                             f"{int(total_bytes / world) >> 20} MiB vs 126 MB L2; no flush"},
            "tokens_per_s": round(B / (ms_step * 1e-3), 1),
            "pct_of_8TBps": round(100.0 * value / world / HBM_NOMINAL_GBPS, 2),
            "pct_of_measured_hbm": round(100.0 * value / world / peak, 2),
            "e2e": {"value": round(e2e_val, 1), "unit": "GB/s", "h2d_bytes_per_step": int(h2d),
                    "d2h_bytes_per_step": int(d2h), "ms_per_step": round(ms_e2e / K, 4),
                    "tokens_per_s": round(B / (ms_e2e / K * 1e-3), 1),
                    "graph_replay": (None if ms_e2e_graph != ms_e2e_graph else {
                        "value": round(total_bytes / (ms_e2e_graph / K * 1e-3) / 1e9, 1),
                        "ms_per_step": round(ms_e2e_graph / K, 4),
                        "what": "the same step (H2D copy, AttentionImpl.forward, D2H copy) "
                                "captured once and replayed from one CUDA graph per step"}),
                    "note": "AttentionImpl.forward with pinned HOST q/k/v/block_table/seq_lens/"
                            "slot_mapping copied in (one staging buffer, one copy) and the output "
                            "copied back every step; the "
                            "paged KV cache is device-resident state, as in the reference"},
            "gpu_launches": int(total_launches),
            "roofline": {"bound": "hbm", "achieved": round(achieved, 1), "peak": peak,
                         "unit": "GB/s", "frac": round(achieved / peak, 4), "traffic": traffic,
                         "kernel": "paged_attention_stream_kernel (+ its merge launch; CUDA "
                                   "events around the attention op, in-loop"
                                   + ("; includes the fused cache write)" if fused else ")"),
                         "kernel_ms": round(attn_ms, 4), "peak_source": peak_src,
                         "bytes": "sum(seq_len) * Hkv_local * D * 2 (K,V) * sizeof(cache_t)"},
            "clocks": clocks,
        }
        if allgather_ms is not None:
            line["bench_allgather_ms"] = round(allgather_ms, 4)
        if cpu is not None:
            line["cpu_baseline"] = cpu
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


# ----------------------------------------------------------------------------
# CPU reference arm (oracle/_ref = the reference's own csrc/cpu kernels)
# ----------------------------------------------------------------------------
def _cpu_sample_inputs(cfg, b_sample):
    _, B, H, Hkv, ctx, kv, _ = CONFIGS[cfg]
    lens = seq_lens_of(cfg)
    # the sample: the first b_sample sequences (for ragged configs: a spread)
    if CONFIGS[cfg][6]:
        order = sorted(range(B), key=lambda i: lens[i])
        pick = [order[int(j * (B - 1) / max(b_sample - 1, 1))] for j in range(b_sample)]
        lens = [min(lens[i], 32768) for i in pick]
    else:
        lens = lens[:b_sample]
    nblk = [(L + BS - 1) // BS for L in lens]
    NB = sum(nblk)
    s = D ** -0.5
    g = torch.Generator().manual_seed(0)
    q = ((torch.rand(len(lens), H, D, generator=g) * 2 - 1) * s).to(torch.bfloat16)
    # the CPU backend ignores kv_cache_dtype (csrc/cpu/attention.cpp:458-476): it runs on the
    # bf16 (dequantised) cache, as the reference's own test does (test_attention.py:303-319)
    kc = torch.empty(NB, Hkv, D // 8, BS, 8, dtype=torch.bfloat16).uniform_(-s, s, generator=g)
    vc = torch.empty(NB, Hkv, D, BS, dtype=torch.bfloat16).uniform_(-s, s, generator=g)
    perm = torch.randperm(NB, generator=g).to(torch.int32)
    bt = torch.zeros(len(lens), max(nblk), dtype=torch.int32)
    pos = 0
    for b, n in enumerate(nblk):
        bt[b, :n] = perm[pos:pos + n]
        pos += n
    sl = torch.tensor(lens, dtype=torch.int32)
    k_new = ((torch.rand(len(lens), Hkv, D, generator=g) * 2 - 1) * s).to(torch.bfloat16)
    last = torch.tensor([L - 1 for L in lens])
    slots = (bt[torch.arange(len(lens)), last // BS].long() * BS + last % BS).to(torch.int64)
    return q, kc, vc, bt, sl, k_new, slots, lens, s


def _cpu_step_fn(cfg, b_sample):
    from oracle import ref_cpu
    _, B, H, Hkv, ctx, kv, _ = CONFIGS[cfg]
    q, kc, vc, bt, sl, k_new, slots, lens, s = _cpu_sample_inputs(cfg, b_sample)
    out = torch.empty_like(q)
    one = torch.tensor(1.0)
    max_len = max(lens)
    kind = "reference"
    if ref_cpu.available():
        ref_cpu.load()

        def step():
            # the CPU backend always uses v1 (cpu_attn.py:823 of the reference)
            ref_cpu.reshape_and_cache(k_new, k_new, kc, vc, slots, "auto", one, one)
            ref_cpu.paged_attention_v1(out, q, kc, vc, Hkv, s, bt, sl, BS, max_len, None, "auto",
                                       one, one, 0, 0, 0, 64, 0)
        cores = torch.get_num_threads()
    else:
        from oracle import oracle as ora
        kind = "port"

        def step():
            ora.reshape_and_cache(k_new, k_new, kc, vc, slots, "auto", one, one)
            ora.paged_attention_v1(out, q, kc, vc, Hkv, s, bt, sl, BS, max_len, None, "auto",
                                   one, one, faithful=False)
        cores = ora.num_threads()
    nbytes = kv_bytes(lens, Hkv, "auto")
    return step, nbytes, len(lens), kind, cores, lens


def cpu_reference_sample(cfg, budget_s=15.0):
    """Time the reference CPU kernels on a bounded sample; returns the cpu_baseline object."""
    torch.set_num_threads(os.cpu_count() or 1)
    b_sample = 4 if CONFIGS[cfg][4] >= 8192 else 8
    b_sample = min(b_sample, CONFIGS[cfg][1])
    step, nbytes, nseq, kind, cores, lens = _cpu_step_fn(cfg, b_sample)
    t0 = time.perf_counter()
    step()
    warm = time.perf_counter() - t0
    n = int(max(3, min(20, budget_s / max(warm, 1e-3))))
    t0 = time.perf_counter()
    for _ in range(n):
        step()
    dt = (time.perf_counter() - t0) / n
    return {"value": round(nbytes / dt / 1e9, 3), "unit": "GB/s", "cores": int(cores), "kind": kind,
            "sample": f"{nseq} of {CONFIGS[cfg][1]} sequences of {cfg} (same per-sequence shape, "
                      f"all KV heads, ctx {max(lens)}), bf16 KV (the CPU backend has no fp8 path), "
                      f"reshape_and_cache + paged_attention_v1, {n} timed calls after 1 warm-up",
            "ms_per_step": round(dt * 1e3, 3), "tokens_per_s": round(nseq / dt, 2),
            "kv_bytes_per_step": int(nbytes)}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cfg = args.config
    desc, B, H, Hkv, ctx, kv, _ = CONFIGS[cfg]
    torch.set_num_threads(os.cpu_count() or 1)
    b_sample = 4 if ctx >= 8192 else 8
    b_sample = min(b_sample, B)
    step, nbytes, nseq, kind, cores, lens = _cpu_step_fn(cfg, b_sample)
    W, K = max(args.warmup, 1), args.steps
    # keep the whole run within a few minutes: cap K by a wall-clock budget
    t0 = time.perf_counter()
    step()
    first = time.perf_counter() - t0
    for _ in range(W - 1):
        step()
    K = int(max(1, min(K, 120.0 / max(first, 1e-3))))
    t0 = time.perf_counter()
    for _ in range(K):
        step()
    dt = (time.perf_counter() - t0) / K
    value = nbytes / dt / 1e9
    sample = (f"{nseq} of {B} sequences of {cfg} per step (same per-sequence shape, all KV heads), "
              f"bf16 KV (the CPU backend ignores kv_cache_dtype), reshape_and_cache + "
              f"paged_attention_v1 (the CPU backend always uses v1)")
    line = {
        "impl": "reference", "metric": "paged_decode_attn_kv_bandwidth",
        "value": round(value, 3), "unit": "GB/s", "n_gpus": args.gpus, "steps": K, "warmup": W,
        "ms_per_step": round(dt * 1e3, 3), "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "bf16 KV/q/out, f32 accumulate (CPU)", "data": "synthetic",
        "config": {"workload": f"{cfg}: {desc}", "tp": args.gpus, "batch": B, "sample": sample},
        "tokens_per_s": round(nseq / dt, 2),
        "cpu_baseline": {"value": round(value, 3), "unit": "GB/s", "cores": int(cores),
                         "kind": kind, "sample": sample},
        "e2e": {"value": round(value, 3), "unit": "GB/s", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--config", default="cfg4", choices=sorted(CONFIGS))
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--unfused", action="store_true",
                    help="step = reshape_and_cache + paged_attention_v2 as two ops (A/B)")
    ap.add_argument("--cpu-budget", type=float, default=15.0)
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()

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

Besides the timed legs the line carries
  parity     -- BEFORE any timing, one step of exactly the tensors that are timed is compared
                with the reference's own CUDA kernels (oracle/_ref/libvllm_ref_cuda.so =
                csrc/attention/paged_attention_v2.cu + csrc/cache_kernels.cu compiled unmodified
                for sm_100a; used ONLY as the checker, never timed here): the whole output of
                every rank within 1e-2 + 1e-2|ref|, the cache bytes torch.equal; at N > 1 the
                all-gathered [B,H,D] is also compared with an UNSHARDED reference run on rank 0
  sustained  -- the same step back to back for >= 3 s with a 100 ms series and the sensors
                (SM / memory clock, power, GPU / HBM temperature) sampled meanwhile
  configs    -- the other four BASELINE.json configs, 10 steps each, with their own parity check
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
# sensors sampler (nvidia-smi during a timed region)
# ----------------------------------------------------------------------------
class ClockSampler:
    FIELDS = ("clocks.sm", "clocks.max.sm", "clocks.mem", "clocks.max.mem", "power.draw",
              "temperature.gpu", "temperature.memory",
              "clocks_event_reasons.hw_slowdown", "clocks_event_reasons.hw_thermal_slowdown",
              "clocks_event_reasons.sw_thermal_slowdown", "clocks_event_reasons.sw_power_cap")
    REASONS = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")

    def __init__(self, index: int):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + ",".join(self.FIELDS),
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    @staticmethod
    def _num(s):
        try:
            return float(s)
        except ValueError:      # "[N/A]" on parts without that sensor
            return None

    def stop(self, series: bool = False):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        cols = {k: [] for k in ("sm", "mem", "power", "tgpu", "tmem")}
        smax = mmax = None
        reasons = set()
        for ln in self.lines:
            p = [x.strip() for x in ln.split(",")]
            if len(p) < len(self.FIELDS):
                continue
            sm = self._num(p[0])
            if sm is None:
                continue
            cols["sm"].append(sm)
            smax = self._num(p[1]) or smax
            mmax = self._num(p[3]) or mmax
            for key, idx in (("mem", 2), ("power", 4), ("tgpu", 5), ("tmem", 6)):
                v = self._num(p[idx])
                if v is not None:
                    cols[key].append(v)
            for n, v in zip(self.REASONS, p[7:11]):
                if v.lower().startswith("active"):
                    reasons.add(n)

        def med(v):
            return sorted(v)[len(v) // 2] if v else None

        out = {"sm_mhz": med(cols["sm"]), "sm_max_mhz": smax, "reasons": sorted(reasons),
               "samples": len(cols["sm"]), "mem_mhz": med(cols["mem"]), "mem_max_mhz": mmax,
               "power_w": med(cols["power"]), "temp_gpu_c": med(cols["tgpu"]),
               "temp_hbm_c": med(cols["tmem"])}
        if series:
            # first / last thirds: the sensor that moves over a multi-second run shows here
            def thirds(v):
                if len(v) < 3:
                    return None
                n = max(1, len(v) // 3)
                return {"first": round(sum(v[:n]) / n, 1), "last": round(sum(v[-n:]) / n, 1),
                        "min": min(v), "max": max(v)}
            out["drift"] = {"sm_mhz": thirds(cols["sm"]), "mem_mhz": thirds(cols["mem"]),
                            "power_w": thirds(cols["power"]), "temp_gpu_c": thirds(cols["tgpu"]),
                            "temp_hbm_c": thirds(cols["tmem"])}
        return out


# ----------------------------------------------------------------------------
# workload
# ----------------------------------------------------------------------------
L2_BYTES = 126 << 20


def cache_copies(cfg, world):
    """Per-GPU KV of a step far above the 126 MB L2: one cache.  Otherwise (cfg1; small shards)
    the step rotates over enough distinct copies of the cache that a copy is long evicted when
    its turn comes again (>= 1 GiB in rotation)."""
    _, B, H, Hkv, ctx, kv, _ = CONFIGS[cfg]
    per_gpu = kv_bytes(seq_lens_of(cfg), Hkv // world, kv)
    if per_gpu >= 4 * L2_BYTES:
        return 1
    return int(min(32, max(2, -(-(1 << 30) // per_gpu))))


def make_shard(cfg, rank, world, device, copies=1):
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
    caches = [kv_cache] + [kv_cache.clone() for _ in range(copies - 1)]
    from vllm_main_modify_b200.paged_attn import PagedAttention
    views = [PagedAttention.split_kv_cache(c, hkv_loc, D) for c in caches]
    return dict(B=B, H=h_loc, Hkv=hkv_loc, kv=kv, lens=lens, max_len=max(lens), scale=s,
                q=q, k_new=k_new, v_new=v_new, kv_cache=kv_cache, kv_caches=caches, views=views,
                bt=bt.to(device), sl=sl.to(device), slots=slots.to(device), ks=ks, vs=vs,
                bt_host=bt, sl_host=sl, slots_host=slots)


class Layer:
    """AttentionLayer protocol (vllm/attention/backends/abstract.py:92-110)."""

    def __init__(self, ks, vs):
        self._k_scale, self._v_scale = ks, vs


def decode_op(w, out, fused, ci=0):
    """The device-resident step on copy `ci` of the cache, through the package's public helper
    (= the reference's PagedAttention.write_to_paged_cache + forward_decode, paged_attn.py:73-198)."""
    from vllm_main_modify_b200.paged_attn import PagedAttention
    kc, vc = w["views"][ci]
    o = out.view(w["B"], w["H"], D)
    if fused:
        PagedAttention.write_and_decode(w["q"], w["k_new"], w["v_new"], kc, vc, w["slots"],
                                        w["bt"], w["sl"], w["max_len"], w["kv"], w["Hkv"],
                                        w["scale"], None, w["ks"], w["vs"], output=o)
        return None
    PagedAttention.write_to_paged_cache(w["k_new"], w["v_new"], kc, vc, w["slots"], w["kv"],
                                        w["ks"], w["vs"])
    return lambda: PagedAttention.forward_decode(w["q"], kc, vc, w["bt"], w["sl"], w["max_len"],
                                                 w["kv"], w["Hkv"], w["scale"], None, w["ks"],
                                                 w["vs"], output=o)


# ----------------------------------------------------------------------------
# parity of exactly what is timed, against the reference's own CUDA kernels (CHECKER ONLY)
# ----------------------------------------------------------------------------
PARITY_ATOL, PARITY_RTOL = 1e-2, 1e-2
PARITY_AGAINST = ("reference CUDA paged_attention_v2 + reshape_and_cache (csrc/attention/"
                  "paged_attention_v2.cu, csrc/cache_kernels.cu compiled unmodified for sm_100a: "
                  "oracle/_ref/libvllm_ref_cuda.so)")


def _ref_cuda_step(ref, q, k_new, v_new, kv_cache, hkv, slots, bt, sl, max_len, kv, scale, ks, vs):
    """reshape_and_cache + paged_attention_v2 of the reference on (a copy of) the cache."""
    from vllm_main_modify_b200.paged_attn import PagedAttention
    kc, vc = PagedAttention.split_kv_cache(kv_cache, hkv, D)
    ref.reshape_and_cache(k_new, v_new, kc, vc, slots, kv, ks, vs)
    B, H, _ = q.shape
    P = (max_len + 511) // 512
    es = torch.empty(B, H, P, dtype=torch.float32, device=q.device)
    ml = torch.empty_like(es)
    tmp = torch.empty(B, H, P, D, dtype=q.dtype, device=q.device)
    want = torch.full_like(q, float("nan"))
    ref.paged_attention_v2(want, es, ml, tmp, q, kc, vc, hkv, scale, bt, sl, BS, max_len, None, kv,
                           ks, vs, 0, 0, 0, 64, 0)
    torch.cuda.synchronize()
    return want


def _diff(got, want):
    g, r = got.float(), want.float()
    d = (g - r).abs()
    viol = d - (PARITY_ATOL + PARITY_RTOL * r.abs())
    big = r.abs() > PARITY_ATOL
    max_rel = float((d[big] / r.abs()[big]).max()) if bool(big.any()) else 0.0
    ok = bool(torch.isfinite(g).all()) and float(viol.max()) <= 0.0
    return ok, float(d.max()), max_rel


def time_ref_cuda(ref, w, steps, graph):
    """The kernel to beat, on the same box and the same tensors: the reference's own
    reshape_and_cache + paged_attention_v2 (compiled unmodified for sm_100a) on a clone of the
    cache, `steps` steps after one warm-up, CUDA events around them (one graph replay when our own
    leg of this config is timed that way).  Scratch is allocated once, outside the timed steps --
    kinder to the reference than its PagedAttention.forward_decode, which allocates per call
    (paged_attn.py:163-173).  Harness shape: benchmarks/kernels/benchmark_paged_attention.py:
    183-193 of the reference.  CHECKER-side only: never inside our timed region."""
    from vllm_main_modify_b200.paged_attn import PagedAttention
    cache = w["kv_caches"][0].clone()
    kc, vc = PagedAttention.split_kv_cache(cache, w["Hkv"], D)
    q = w["q"]
    B, H, _ = q.shape
    P = (w["max_len"] + 511) // 512
    es = torch.empty(B, H, P, dtype=torch.float32, device=q.device)
    ml = torch.empty_like(es)
    tmp = torch.empty(B, H, P, D, dtype=q.dtype, device=q.device)
    o = torch.empty_like(q)

    def one():
        ref.reshape_and_cache(w["k_new"], w["v_new"], kc, vc, w["slots"], w["kv"], w["ks"], w["vs"])
        ref.paged_attention_v2(o, es, ml, tmp, q, kc, vc, w["Hkv"], w["scale"], w["bt"], w["sl"],
                               BS, w["max_len"], None, w["kv"], w["ks"], w["vs"], 0, 0, 0, 64, 0)

    one()
    torch.cuda.synchronize()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if graph:
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=side):
                for _ in range(steps):
                    one()
        torch.cuda.current_stream().wait_stream(side)
        g.replay()
        torch.cuda.synchronize()
        ev0.record()
        g.replay()
        ev1.record()
    else:
        ev0.record()
        for _ in range(steps):
            one()
        ev1.record()
    torch.cuda.synchronize()
    ms = ev0.elapsed_time(ev1) / steps
    del cache, kc, vc, es, ml, tmp, o
    torch.cuda.empty_cache()
    return ms


def ref_cuda_entry(w, world, dev, our_ms_step, total_bytes, graph, steps=3):
    """{"ms_per_step", "gbs", "speedup"} of the reference's CUDA kernels on this workload (max over
    ranks), or why there is none."""
    import torch.distributed as dist
    try:
        from oracle import ref_cuda as ref
        if not os.path.exists(ref.LIB_PATH):
            raise RuntimeError("oracle/_ref/libvllm_ref_cuda.so is not built")
        ref.load()
        ms = time_ref_cuda(ref, w, steps, graph)
    except Exception as exc:   # every rank fails alike (same files): no collective to keep in step
        return {"unavailable": str(exc)}
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t[0])
    return {"ms_per_step": round(ms, 4), "gbs": round(total_bytes / (ms * 1e-3) / 1e9, 1),
            "speedup": round(ms / our_ms_step, 2), "steps": steps,
            "what": "reference CUDA reshape_and_cache + paged_attention_v2 (oracle/_ref/"
                    "libvllm_ref_cuda.so, sm_100a build of the reference's own sources) on a clone "
                    "of the timed tensors; speedup = its ms_per_step / ours (device-timed)"}


def parity_status(parity):
    """ "ok" | "failed" | "unchecked" -- an absent checker must not read as a pass."""
    if not parity.get("checked"):
        return "unchecked"
    return "ok" if parity.get("ok") else "failed"


def parity_check(cfg, w, out, fused, rank, world, dev):
    """Runs ONE step of the op that is about to be timed, on the tensors that are about to be
    timed, and compares output and cache bytes with the reference's CUDA kernels run on a clone
    of the same cache.  Must be called before any other step touched the cache."""
    import torch.distributed as dist
    try:
        from oracle import ref_cuda as ref     # the checker; the product never loads it
        if not os.path.exists(ref.LIB_PATH):
            raise RuntimeError("oracle/_ref/libvllm_ref_cuda.so is not built")
        ref.load()
    except Exception as exc:
        return {"checked": False, "why": str(exc)}
    res = {"checked": True, "against": PARITY_AGAINST,
           "gate": f"|out - ref| <= {PARITY_ATOL} + {PARITY_RTOL}|ref| over the whole "
                   "[B, H/n, D] output of every rank; written cache torch.equal (whole tensor)"}
    ok_all, max_abs, max_rel, cache_ok = True, 0.0, 0.0, True
    for ci, kvc in enumerate(w["kv_caches"][:2]):   # the rotation's copies are clones: two suffice
        ref_cache = kvc.clone()
        want = _ref_cuda_step(ref, w["q"], w["k_new"], w["v_new"], ref_cache, w["Hkv"], w["slots"],
                              w["bt"], w["sl"], w["max_len"], w["kv"], w["scale"], w["ks"], w["vs"])
        out.fill_(float("nan"))
        attend = decode_op(w, out, fused, ci)
        if attend is not None:
            attend()
        torch.cuda.synchronize()
        ok, a, r = _diff(out.view_as(want), want)
        ok_all &= ok
        max_abs, max_rel = max(max_abs, a), max(max_rel, r)
        cache_ok &= bool(torch.equal(kvc, ref_cache))
        del ref_cache
    flags = torch.tensor([float(ok_all), float(cache_ok), -max_abs, -max_rel], dtype=torch.float64,
                         device=dev)
    if world > 1:
        dist.all_reduce(flags, op=dist.ReduceOp.MIN)
    ok_all, cache_ok = bool(flags[0] > 0), bool(flags[1] > 0)
    res.update(ok=ok_all and cache_ok, max_abs=round(-float(flags[2]), 6),
               max_rel=round(-float(flags[3]), 6), cache_bytes_equal=cache_ok, ranks=world)
    if world > 1:
        # the all-gathered [B, H, D] against ONE unsharded reference run on rank 0: rank 0
        # regenerates every rank's shard (same seeds, same device model) into the full tensors
        _, B, H, Hkv, _, kv, _ = CONFIGS[cfg]
        gathered = torch.empty(world, B, w["H"], D, dtype=out.dtype, device=dev)
        dist.all_gather_into_tensor(gathered, out.view(B, w["H"], D).contiguous())
        gres = torch.zeros(3, dtype=torch.float64, device=dev)
        if rank == 0:
            hkv_loc = Hkv // world
            full_q = torch.empty(B, H, D, dtype=out.dtype, device=dev)
            full_k = torch.empty(B, Hkv, D, dtype=out.dtype, device=dev)
            full_v = torch.empty_like(full_k)
            NB = w["kv_cache"].shape[1]
            full_cache = torch.empty(2, NB, Hkv, BS * D, dtype=w["kv_cache"].dtype, device=dev)
            for r_ in range(world):
                sh = w if r_ == 0 else make_shard(cfg, r_, world, dev)
                full_q[:, r_ * w["H"]:(r_ + 1) * w["H"]] = sh["q"]
                full_k[:, r_ * hkv_loc:(r_ + 1) * hkv_loc] = sh["k_new"]
                full_v[:, r_ * hkv_loc:(r_ + 1) * hkv_loc] = sh["v_new"]
                # rank 0's own cache already holds the new token; the reference rewrites it
                full_cache[:, :, r_ * hkv_loc:(r_ + 1) * hkv_loc] = \
                    sh["kv_cache"].view(2, NB, hkv_loc, BS * D)
                del sh
            want = _ref_cuda_step(ref, full_q, full_k, full_v, full_cache.view(2, NB, -1), Hkv,
                                  w["slots"], w["bt"], w["sl"], w["max_len"], kv, w["scale"],
                                  w["ks"], w["vs"])
            got = gathered.permute(1, 0, 2, 3).reshape(B, H, D)
            ok, a, r = _diff(got, want)
            gres = torch.tensor([float(ok), a, r], dtype=torch.float64, device=dev)
            del full_cache, want
        dist.broadcast(gres, src=0)
        res["gathered_vs_unsharded"] = {"ok": bool(gres[0] > 0), "max_abs": round(float(gres[1]), 6),
                                        "max_rel": round(float(gres[2]), 6),
                                        "what": f"all-gather of the {world} ranks' [B,H/{world},D] vs "
                                                "one unsharded reference run (all KV heads) on rank 0"}
        res["ok"] = res["ok"] and res["gathered_vs_unsharded"]["ok"]
    torch.cuda.empty_cache()
    return res


def timed_steps(w, out, fused, K, barrier, native, graph=False):
    """K back-to-back steps, CUDA events around the whole loop and around every attention op.
    graph=True (steps of tens of microseconds, where the Python caller -- not the GPU -- would
    set the pace): the K steps are captured into ONE CUDA graph and its replay is timed."""
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = len(w["kv_caches"])
    if graph:
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        launches0 = native.launch_count()
        with torch.cuda.stream(side):
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=side):
                for i in range(K):
                    attend = decode_op(w, out, fused, i % n)
                    if attend is not None:
                        attend()
        launches = native.launch_count() - launches0
        torch.cuda.current_stream().wait_stream(side)
        g.replay()                       # warm-up replay
        barrier()
        ev0.record()
        g.replay()
        ev1.record()
        barrier()
        ms = ev0.elapsed_time(ev1)
        return ms, ms / K, launches
    a_ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
            for _ in range(K)]
    launches0 = native.launch_count()
    barrier()
    ev0.record()
    for i in range(K):
        if fused:
            # one launch: the decode kernel also quantises + stores the new token (decode_step)
            a_ev[i][0].record()
            decode_op(w, out, True, i % n)
            a_ev[i][1].record()
            continue
        attend = decode_op(w, out, False, i % n)
        a_ev[i][0].record()
        attend()
        a_ev[i][1].record()
    ev1.record()
    barrier()
    launches = native.launch_count() - launches0
    return ev0.elapsed_time(ev1), sum(a.elapsed_time(b) for a, b in a_ev) / K, launches


def sustained_plan(seconds, ms_step_hint):
    """(steps per chunk, chunks) of the sustained leg: ~100 ms of work per chunk.  A pure function
    of its arguments -- every rank must be given the SAME hint (the per-chunk times are
    all-reduced as one tensor; see tests/test_bench_host.py)."""
    chunk = max(1, int(round(100.0 / max(ms_step_hint, 1e-3))))
    n_chunks = max(3, int(seconds * 1e3 / (chunk * ms_step_hint)) + 1)
    return chunk, n_chunks


def sustained_leg(w, out, fused, seconds, ms_step_hint, my_bytes, sampler_index, barrier):
    """The same step back to back for >= `seconds`, an event every ~100 ms of work."""
    chunk, n_chunks = sustained_plan(seconds, ms_step_hint)
    caches = w["kv_caches"]
    sampler = ClockSampler(sampler_index)
    sampler.start()
    time.sleep(0.25)
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(n_chunks + 1)]
    barrier()
    evs[0].record()
    i = 0
    for c in range(n_chunks):
        for _ in range(chunk):
            attend = decode_op(w, out, fused, i % len(caches))
            if attend is not None:
                attend()
            i += 1
        evs[c + 1].record()
        if c >= 2:
            evs[c - 1].synchronize()      # keep the launch queue bounded, never drained
    barrier()
    sens = sampler.stop(series=True)
    per = [evs[c].elapsed_time(evs[c + 1]) / chunk for c in range(n_chunks)]
    total_ms = evs[0].elapsed_time(evs[-1])
    return per, total_ms, n_chunks * chunk, chunk, sens


def run_side_config(cfg, rank, world, dev, fused, check, barrier, native, peak, steps=10):
    """One of the other BASELINE configs: parity check, 3 warm-ups, `steps` timed steps."""
    import torch.distributed as dist
    _, B, H, Hkv, ctx, kv, _ = CONFIGS[cfg]
    copies = cache_copies(cfg, world)
    w = make_shard(cfg, rank, world, dev, copies)
    out = torch.empty(B, w["H"] * D, dtype=torch.bfloat16, device=dev)
    par = parity_check(cfg, w, out, fused, rank, world, dev) if check else {"checked": False,
                                                                            "why": "--no-check"}
    for i in range(3):
        attend = decode_op(w, out, fused, i % copies)
        if attend is not None:
            attend()
    # steps shorter than ~0.1 ms: replayed from one CUDA graph (the Python caller would pace them)
    per_gpu = kv_bytes(w["lens"], w["Hkv"], kv)
    graph = per_gpu < 512 << 20
    ms_total, attn_ms, launches = timed_steps(w, out, fused, steps, barrier, native, graph)
    my_bytes = kv_bytes(w["lens"], w["Hkv"], kv)
    st = torch.tensor([ms_total, attn_ms], dtype=torch.float64, device=dev)
    tot = torch.tensor([float(my_bytes)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(st, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    ms_step = float(st[0]) / steps
    gbs = float(tot[0]) / (ms_step * 1e-3) / 1e9
    op_gbs = float(tot[0]) / world / (float(st[1]) * 1e-3) / 1e9
    ref_cuda = (ref_cuda_entry(w, world, dev, ms_step, float(tot[0]), graph) if check
                else {"unavailable": "--no-check"})
    del w, out
    torch.cuda.empty_cache()
    return {"gbs": round(gbs, 1), "ms_per_step": round(ms_step, 4), "steps": steps,
            "ref_cuda": ref_cuda, "parity_status": parity_status(par),
            "kv_bytes_total": int(tot[0]), "gbs_per_gpu": round(gbs / world, 1),
            "frac_of_measured_hbm": round(gbs / world / peak, 4),
            "pct_of_8TBps": round(100.0 * gbs / world / HBM_NOMINAL_GBPS, 2),
            "attn_op_gbs_per_gpu": round(op_gbs, 1), "launches_per_step": launches / steps,
            "timing": ("one CUDA graph of the %d steps, events around its replay" % steps) if graph
                      else "eager launches, events around the loop",
            "l2": ("one cache, %d MiB per GPU per step" % (int(tot[0]) / world / 2 ** 20)
                   if copies == 1 else
                   "%d MiB per GPU per step: rotates over %d copies of the cache (L2 is 126 MB)"
                   % (int(tot[0]) / world / 2 ** 20, copies)),
            "tokens_per_s": round(B / (ms_step * 1e-3), 1),
            "parity": {k: par[k] for k in ("checked", "ok", "max_abs", "max_rel",
                                           "cache_bytes_equal", "why", "gathered_vs_unsharded")
                       if k in par}}


def run_ours(args):
    import torch.distributed as dist
    from vllm_main_modify_b200 import _native
    from vllm_main_modify_b200.backend import B200PagedAttentionBackend, B200PagedAttentionMetadata

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

    peaks = {}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peaks = json.load(f)
    except OSError:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"

    cfg = args.config
    desc, B, H, Hkv, ctx, kv, _ = CONFIGS[cfg]
    copies = cache_copies(cfg, world)
    w = make_shard(cfg, rank, world, dev, copies)
    impl = B200PagedAttentionBackend.get_impl_cls()(w["H"], D, w["scale"], w["Hkv"], None, None, kv)
    layer = Layer(w["ks"], w["vs"])
    meta = B200PagedAttentionMetadata(num_actual_tokens=B, max_query_len=1, seq_lens=w["sl"],
                                      max_seq_len=w["max_len"], block_table=w["bt"],
                                      slot_mapping=w["slots"])
    q2 = w["q"].view(B, -1)
    k2, v2 = w["k_new"].view(B, -1), w["v_new"].view(B, -1)
    out = torch.empty_like(q2)
    fused = not args.unfused
    impl.fused_decode_step = fused

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- parity of what is about to be timed (outside every timed region) ----
    parity = (parity_check(cfg, w, out, fused, rank, world, dev) if args.check
              else {"checked": False, "why": "--no-check"})

    def step():
        impl.forward(layer, q2, k2, v2, w["kv_cache"], meta, output=out)

    for _ in range(max(args.warmup, 3)):
        step()
    barrier()

    # ---- timed region 1: device-resident inputs ("value") -------------------
    sampler = ClockSampler(local_rank)
    sampler.start()
    time.sleep(0.3)
    K = args.steps
    ms_total, attn_ms, launches = timed_steps(w, out, fused, K, barrier, _native)
    clocks = sampler.stop()

    # ---- sustained leg: the same step for >= args.sustain seconds -----------
    my_bytes = kv_bytes(w["lens"], w["Hkv"], kv)
    sustained = None
    if args.sustain > 0:
        # every rank must cut the leg into the SAME number of chunks (the per-chunk times are
        # all-reduced below): the step-time hint is the maximum over ranks, not this rank's own
        hint = torch.tensor([ms_total / K], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(hint, op=dist.ReduceOp.MAX)
        per, sus_ms, sus_steps, chunk, sens = sustained_leg(w, out, fused, args.sustain,
                                                            float(hint[0]), my_bytes, local_rank,
                                                            barrier)
        sus = torch.tensor([sus_ms] + per, dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(sus, op=dist.ReduceOp.MAX)
        sustained = (sus.tolist(), sus_steps, chunk, sens)

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
    e2e_caches = w["kv_caches"]
    e2e_i = [0]

    def e2e_step():
        dbuf.copy_(hbuf, non_blocking=True)
        impl.forward(layer, dq, dk, dv, e2e_caches[e2e_i[0] % len(e2e_caches)], meta2, output=out)
        e2e_i[0] += 1
        hout.copy_(out, non_blocking=True)
        torch.cuda.current_stream().synchronize()     # the caller reads the result

    for _ in range(3):
        e2e_step()
    if os.environ.get("B200PA_E2E_BREAKDOWN") == "1" and rank == 0:
        # development aid: device-side segments of the eager step (CUDA events between the parts)
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
        e2e_i[0] = 0
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
        del gathered
    else:
        allgather_ms = None
    ms_total, ms_e2e, attn_ms, ms_e2e_graph = [float(x) for x in stats.tolist()]
    total_bytes, total_launches = tot.tolist()

    # ---- the kernel to beat on this box: the reference's own CUDA kernels, same tensors -------
    ref_cuda_main = (ref_cuda_entry(w, world, dev, ms_total / K, total_bytes, False)
                     if args.check else {"unavailable": "--no-check"})

    # ---- the other BASELINE configs (driver-visible, 10 steps each) ---------
    del w, impl, meta, meta2, q2, k2, v2, dbuf, dq, dk, dv, dbt, dsl, dslots, e2e_caches
    torch.cuda.empty_cache()
    side_configs = {}
    if args.side_configs:
        for c in sorted(CONFIGS):
            if c == cfg or CONFIGS[c][3] % world:
                continue
            try:
                side_configs[c] = run_side_config(c, rank, world, dev, fused, args.check, barrier,
                                                  _native, peak)
            except Exception as exc:       # a side config must not take the main line down
                side_configs[c] = {"error": f"{type(exc).__name__}: {exc}"}
                barrier()

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu = cpu_reference_sample(cfg, budget_s=args.cpu_budget)

    if rank == 0:
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
        ms_step = ms_total / K
        value = total_bytes / (ms_step * 1e-3) / 1e9
        e2e_val = total_bytes / (ms_e2e / K * 1e-3) / 1e9
        # roofline of the dominant kernel, per GPU: this rank's KV bytes / attention time
        achieved = (total_bytes / world) / (attn_ms * 1e-3) / 1e9
        per_gpu_mib = int(total_bytes / world) >> 20
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
                       "l2": (f"inputs larger than L2: KV read per GPU per step = {per_gpu_mib} MiB "
                              "vs 126 MB L2; no flush") if copies == 1 else
                             (f"KV read per GPU per step = {per_gpu_mib} MiB: the step rotates over "
                              f"{copies} distinct copies of the cache (126 MB L2)")},
            "tokens_per_s": round(B / (ms_step * 1e-3), 1),
            "pct_of_8TBps": round(100.0 * value / world / HBM_NOMINAL_GBPS, 2),
            "pct_of_measured_hbm": round(100.0 * value / world / peak, 2),
            "parity": parity,
            "parity_status": parity_status(parity),
            "ref_cuda": ref_cuda_main,
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
                         "traffic_source": "profiles/ncu_traffic.json (ncu --set full capture of "
                                           "this workload, committed; not measured in this run)",
                         "kernel": ("paged_attention_nhd_kernel (whole-block decode kernel"
                                    if world <= 4 else
                                    "paged_attention_stream_kernel (per-warp decode kernel") +
                                   "; CUDA events around the attention op, in-loop"
                                   + ("; includes the fused cache write)" if fused else ")"),
                         "kernel_ms": round(attn_ms, 4), "peak_source": peak_src,
                         "bytes": "sum(seq_len) * Hkv_local * D * 2 (K,V) * sizeof(cache_t)"},
            "clocks": clocks,
        }
        if sustained is not None:
            vals, sus_steps, chunk, sens = sustained
            sus_ms, per = vals[0], sorted(vals[1:])
            series = [round(total_bytes / (m * 1e-3) / 1e9, 1) for m in vals[1:]]
            line["sustained"] = {
                "seconds": round(sus_ms / 1e3, 3), "steps": sus_steps,
                "gbs": round(total_bytes * sus_steps / (sus_ms * 1e-3) / 1e9, 1),
                "pct_of_8TBps": round(100.0 * total_bytes * sus_steps / (sus_ms * 1e-3) / 1e9
                                      / world / HBM_NOMINAL_GBPS, 2),
                "ms_per_step_p50": round(per[len(per) // 2], 4),
                "ms_per_step_p95": round(per[min(len(per) - 1, int(0.95 * len(per)))], 4),
                "series_100ms": series, "steps_per_point": chunk,
                "first_vs_last_third": [round(sum(series[:max(1, len(series) // 3)])
                                              / max(1, len(series) // 3), 1),
                                        round(sum(series[-max(1, len(series) // 3):])
                                              / max(1, len(series) // 3), 1)],
                "sensors": sens,
                "what": "the timed step back to back (whole job, max over ranks per point)"}
        if side_configs:
            line["configs"] = side_configs
        if allgather_ms is not None:
            line["bench_allgather_ms"] = round(allgather_ms, 4)
        if cpu is not None:
            line["cpu_baseline"] = cpu
        print(json.dumps(line), flush=True)
    bad = parity.get("checked") and not parity.get("ok")
    unchecked = args.check and not parity.get("checked")
    if world > 1:
        dist.destroy_process_group()
    if bad:
        raise SystemExit("bench.py: PARITY FAILED against the reference's CUDA kernels: "
                         + json.dumps(parity))
    if unchecked:
        # the line above carries "parity_status": "unchecked"; a run without its checker is not a
        # pass (build oracle/_ref with __graft_entry__.build(), or say --no-check)
        raise SystemExit("bench.py: PARITY UNCHECKED (" + str(parity.get("why")) + "); the number "
                         "above was not compared with the reference's CUDA kernels")


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
    ap.add_argument("--no-check", dest="check", action="store_false",
                    help="skip the parity check against the reference's CUDA kernels")
    ap.add_argument("--sustain", type=float, default=3.0,
                    help="seconds of the back-to-back sustained leg (0 = off)")
    ap.add_argument("--no-side-configs", dest="side_configs", action="store_false",
                    help="do not run the other four BASELINE configs after the main one")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()

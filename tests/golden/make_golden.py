#!/usr/bin/env python
"""Generate the committed golden fixtures under tests/golden/.

Runs ONLY in the build container (needs /root/reference, nvcc and oracle/_ref);
the fixtures it writes are what travels.  Nothing here is imported by the
product or by the tests -- tests only read the .npz files.

Sources of truth used (all the reference's own code, none of ours):
  fp8_golden.npz    CUDA toolkit 12.9 host implementations of
                    __nv_cvt_float_to_fp8 / __nv_cvt_fp8_to_halfraw /
                    __float2bfloat16 / __float2half_rn (gen_fp8_golden.cu) --
                    the primitives quant_utils.cuh:300-512 is written against.
  attn_*.npz        (a) `ref_masked_attention` +
                    `ref_single_query_cached_kv_attention`, extracted verbatim
                    at run time (ast) from the reference's
                    tests/kernels/attention/test_attention.py:50-118 and
                    executed here; (b) the reference's CPU kernels
                    csrc/cpu/attention.cpp (v1 and v2) from oracle/_ref.
  cache_*.npz       the reference's CPU reshape_and_cache (csrc/cpu/cache.cpp)
                    from oracle/_ref.
  merge_*.npz       `merge_attn_states_torch`, extracted verbatim at run time (ast)
                    from the reference's
                    tests/kernels/attention/test_merge_attn_states.py:16-46 and
                    executed here (inputs include +inf / -inf LSEs like that test).
"""
from __future__ import annotations

import ast
import os
import random
import subprocess
import sys
import tempfile
from typing import Optional  # noqa: F401  (used by the exec'd reference code)

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = os.environ.get("REFERENCE_ROOT", "/root/reference")

from oracle import ref_cpu  # noqa: E402


def make_fp8():
    with tempfile.TemporaryDirectory() as td:
        exe = os.path.join(td, "gen_fp8")
        subprocess.check_call(["nvcc", "-O1", "-w", "-o", exe,
                               os.path.join(HERE, "gen_fp8_golden.cu")])
        raw = subprocess.check_output([exe])
    off = 0

    def take(n, dt):
        nonlocal off
        a = np.frombuffer(raw, dtype=dt, count=n, offset=off).copy()
        off += a.nbytes
        return a

    out = dict(
        e4m3_from_bf16=take(65536, np.uint8),
        e5m2_from_bf16=take(65536, np.uint8),
        e4m3_from_f16=take(65536, np.uint8),
        e5m2_from_f16=take(65536, np.uint8),
        half_from_e4m3=take(256, np.uint16),
        half_from_e5m2=take(256, np.uint16),
        bf16_from_f32=take(65536 * 4, np.uint16),
        f16_from_f32=take(65536 * 4, np.uint16),
        scales=np.array([0.5, 0.0123, 3.7], dtype=np.float32),
        e4m3_scaled=take(3 * 65536, np.uint8).reshape(3, 65536),
    )
    assert off == len(raw)
    np.savez_compressed(os.path.join(HERE, "fp8_golden.npz"), **out)
    print("fp8_golden.npz", os.path.getsize(os.path.join(HERE, "fp8_golden.npz")))


def load_reference_test_oracle():
    """exec the two reference functions straight from the reference's test."""
    path = os.path.join(REF, "tests/kernels/attention/test_attention.py")
    src = open(path).read()
    tree = ast.parse(src)
    wanted = {"ref_masked_attention", "ref_single_query_cached_kv_attention"}
    chunks = [ast.get_source_segment(src, n) for n in tree.body
              if isinstance(n, ast.FunctionDef) and n.name in wanted]
    assert len(chunks) == 2
    ns = {"torch": torch, "Optional": Optional}
    exec("\n\n".join(chunks), ns)
    return ns["ref_single_query_cached_kv_attention"]


def bits(t: torch.Tensor) -> np.ndarray:
    if t.dtype == torch.bfloat16 or t.dtype == torch.float16:
        return t.contiguous().view(torch.int16).numpy().view(np.uint16)
    return t.contiguous().numpy()


def make_attn_case(name, B, H, Hkv, D, BS, NB, lens, dtype, alibi, seed):
    torch.manual_seed(seed)
    random.seed(seed)
    x = 16 // torch.tensor([], dtype=dtype).element_size()
    scale = float(D ** -0.5)
    # wider than the reference generator (U(-s,s) makes softmax ~uniform):
    q = torch.empty(B, H, D, dtype=dtype).uniform_(-1.0, 1.0)
    kc = torch.empty(NB, Hkv, D // x, BS, x, dtype=dtype).uniform_(-1.0, 1.0)
    vc = torch.empty(NB, Hkv, D, BS, dtype=dtype).uniform_(-1.0, 1.0)
    max_len = max(lens)
    mb = (max_len + BS - 1) // BS
    bt = torch.tensor([[random.randint(0, NB - 1) for _ in range(mb)]
                       for _ in range(B)], dtype=torch.int32)
    sl = torch.tensor(lens, dtype=torch.int32)
    slopes = torch.randn(H, dtype=torch.float32) * 0.1 if alibi else None
    one = torch.tensor(1.0)

    ref_fn = load_reference_test_oracle()
    out_test = torch.empty_like(q)
    ref_fn(out_test, q, H // Hkv, kc, vc, bt, sl, scale, slopes)

    res = dict(q=bits(q), key_cache=bits(kc), value_cache=bits(vc),
               block_tables=bt.numpy(), seq_lens=sl.numpy(),
               scale=np.float32(scale), dtype=str(dtype).split(".")[-1],
               num_kv_heads=np.int32(Hkv), block_size=np.int32(BS),
               out_test_ref=bits(out_test))
    if slopes is not None:
        res["alibi_slopes"] = slopes.numpy()
    if BS == 16:  # the reference CPU kernels support block_size 16 only
        o1 = torch.empty_like(q)
        ref_cpu.paged_attention_v1(o1, q, kc, vc, Hkv, scale, bt, sl, BS,
                                   max_len, slopes, "auto", one, one, 0, 0, 0,
                                   64, 0)
        P = (max_len + 511) // 512
        es = torch.zeros(B, H, P)
        ml = torch.zeros(B, H, P)
        tmp = torch.zeros(B, H, P, D, dtype=dtype)
        o2 = torch.empty_like(q)
        ref_cpu.paged_attention_v2(o2, es, ml, tmp, q, kc, vc, Hkv, scale, bt,
                                   sl, BS, max_len, slopes, "auto", one, one,
                                   0, 0, 0, 64, 0)
        res["out_cpu_v1"] = bits(o1)
        res["out_cpu_v2"] = bits(o2)
    path = os.path.join(HERE, f"attn_{name}.npz")
    np.savez_compressed(path, **res)
    print(os.path.basename(path), os.path.getsize(path))


def make_cache_case(name, T, Hkv, D, BS, NB, dtype, seed, strided):
    torch.manual_seed(seed)
    random.seed(seed)
    x = 16 // torch.tensor([], dtype=dtype).element_size()
    if strided:  # key/value are slices of a fused qkv tensor (row stride > Hkv*D)
        qkv = torch.randn(T, 3, Hkv, D, dtype=dtype)
        key, value = qkv[:, 1], qkv[:, 2]
    else:
        key = torch.randn(T, Hkv, D, dtype=dtype)
        value = torch.randn(T, Hkv, D, dtype=dtype)
    slots = random.sample(range(NB * BS), T)
    slot_mapping = torch.tensor(slots, dtype=torch.int64)
    kc = torch.zeros(NB, Hkv, D // x, BS, x, dtype=dtype)
    vc = torch.zeros(NB, Hkv, D, BS, dtype=dtype)
    one = torch.tensor(1.0)
    ref_cpu.reshape_and_cache(key, value, kc, vc, slot_mapping, "auto", one, one)
    path = os.path.join(HERE, f"cache_{name}.npz")
    np.savez_compressed(path, key=bits(key.contiguous()),
                        value=bits(value.contiguous()),
                        slot_mapping=slot_mapping.numpy(),
                        key_cache=bits(kc), value_cache=bits(vc),
                        dtype=str(dtype).split(".")[-1],
                        block_size=np.int32(BS), strided=np.int32(strided))
    print(os.path.basename(path), os.path.getsize(path))


def load_reference_merge_oracle():
    path = os.path.join(REF, "tests/kernels/attention/test_merge_attn_states.py")
    src = open(path).read()
    tree = ast.parse(src)
    chunks = [ast.get_source_segment(src, n) for n in tree.body
              if isinstance(n, ast.FunctionDef) and n.name == "merge_attn_states_torch"]
    assert len(chunks) == 1
    ns = {"torch": torch, "Optional": Optional}
    exec(chunks[0], ns)
    return ns["merge_attn_states_torch"]


def make_merge_case(name, n, h, d, dtype, seed):
    torch.manual_seed(seed)
    p_lse = torch.randn(h, n, dtype=torch.float32)
    s_lse = torch.randn(h, n, dtype=torch.float32)
    # like the reference test (:113-131): some +inf / -inf entries, never both sides at once
    mask_p = torch.rand(h, n) < 0.1
    mask_s = (torch.rand(h, n) < 0.1) & ~mask_p
    p_lse[mask_p] = float("inf")
    s_lse[mask_s] = float("inf")
    s_lse[0, 0] = -float("inf")
    p_out = torch.randn(n, h, d, dtype=dtype)
    s_out = torch.randn(n, h, d, dtype=dtype)
    fn = load_reference_merge_oracle()
    out, out_lse = fn(torch.empty_like(p_out), p_out.float(), p_lse.clone(), s_out.float(),
                      s_lse.clone(), torch.empty(h, n))
    path = os.path.join(HERE, f"merge_{name}.npz")
    np.savez_compressed(path, prefix_output=bits(p_out), suffix_output=bits(s_out),
                        prefix_lse=p_lse.numpy(), suffix_lse=s_lse.numpy(),
                        output_f32=out.float().numpy(), output_lse=out_lse.numpy(),
                        dtype=str(dtype).split(".")[-1])
    print(os.path.basename(path), os.path.getsize(path))


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "merge":
        make_merge_case("bf16_d128", 37, 8, 128, torch.bfloat16, 0)
        make_merge_case("f16_d96", 19, 4, 96, torch.float16, 1)
        make_merge_case("f32_d32", 23, 3, 32, torch.float32, 2)
        sys.exit(0)
    make_fp8()
    bf16 = torch.bfloat16
    make_attn_case("gqa_bf16_d128", 4, 8, 2, 128, 16, 24, [1, 16, 333, 700], bf16, False, 0)
    make_attn_case("mha_bf16_d64_alibi", 3, 4, 4, 64, 16, 16, [5, 129, 512], bf16, True, 1)
    make_attn_case("gqa_f16_d80", 3, 6, 3, 80, 16, 16, [31, 32, 600], torch.float16, False, 2)
    make_attn_case("gqa_bf16_d128_bs32", 3, 8, 1, 128, 32, 12, [7, 64, 555], bf16, False, 3)
    make_attn_case("mha_f32_d32", 2, 2, 2, 32, 16, 8, [9, 100], torch.float32, False, 4)
    make_cache_case("bf16", 13, 4, 128, 16, 8, bf16, 0, False)
    make_cache_case("bf16_strided", 9, 2, 64, 16, 8, bf16, 1, True)
    make_cache_case("f16", 7, 3, 80, 16, 4, torch.float16, 2, False)
    make_merge_case("bf16_d128", 37, 8, 128, torch.bfloat16, 0)
    make_merge_case("f16_d96", 19, 4, 96, torch.float16, 1)
    make_merge_case("f32_d32", 23, 3, 32, torch.float32, 2)

#!/usr/bin/env python
"""Generate tests/golden/prefill_*.npz (build container only: needs /root/reference).

Source of truth: ``ref_paged_attn``, extracted verbatim at run time (ast) from the reference's
tests/kernels/attention/test_flash_attn.py:27-79 and executed here -- the reference's own oracle
for ``flash_attn_varlen_func(..., block_table)``, the call its paged backends make for batches
with more than one query token per sequence (vllm/v1/attention/backends/cpu_attn.py:584-612).
It works on the token-major cache ``[NB, BS, Hkv, D]``; the fixtures store the SAME cache in the
x-packed layout of the paged ops (pure re-indexing, oracle/cache_ops_oracle.py::flash_to_xpacked).
Nothing here is imported by the product or by the tests -- tests only read the .npz files.
"""
from __future__ import annotations

import ast
import os
import random
import sys
from typing import Optional  # noqa: F401  (used by the exec'd reference code)

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = os.environ.get("REFERENCE_ROOT", "/root/reference")

from oracle.cache_ops_oracle import flash_to_xpacked  # noqa: E402


def load_ref_paged_attn():
    path = os.path.join(REF, "tests/kernels/attention/test_flash_attn.py")
    src = open(path).read()
    tree = ast.parse(src)
    chunks = [ast.get_source_segment(src, n) for n in tree.body
              if isinstance(n, ast.FunctionDef) and n.name == "ref_paged_attn"]
    assert len(chunks) == 1
    ns = {"torch": torch, "Optional": Optional}
    exec(chunks[0], ns)
    return ns["ref_paged_attn"]


def bits(t: torch.Tensor) -> np.ndarray:
    if t.dtype in (torch.bfloat16, torch.float16):
        return t.contiguous().view(torch.int16).numpy().view(np.uint16)
    return t.contiguous().numpy()


def make_case(name, query_lens, kv_lens, H, Hkv, D, BS, NB, dtype, seed):
    torch.manual_seed(seed)
    random.seed(seed)
    T = sum(query_lens)
    scale = float(D ** -0.5)
    q = torch.randn(T, H, D, dtype=dtype)
    kc = torch.randn(NB, BS, Hkv, D, dtype=dtype)
    vc = torch.randn(NB, BS, Hkv, D, dtype=dtype)
    mb = (max(kv_lens) + BS - 1) // BS
    bt = torch.tensor([[random.randint(0, NB - 1) for _ in range(mb)]
                       for _ in range(len(kv_lens))], dtype=torch.int32)
    ref = load_ref_paged_attn()
    # (the reference oracle scales its query argument in place: hand it a copy)
    out = ref(q.clone(), kc, vc, query_lens, kv_lens, bt, scale)
    kx, vx = flash_to_xpacked(kc, vc)
    qsl = np.concatenate([[0], np.cumsum(query_lens)]).astype(np.int32)
    path = os.path.join(HERE, f"prefill_{name}.npz")
    np.savez_compressed(path, q=bits(q), key_cache=bits(kx.contiguous()),
                        value_cache=bits(vx.contiguous()), block_tables=bt.numpy(),
                        query_start_loc=qsl, seq_lens=np.asarray(kv_lens, dtype=np.int32),
                        scale=np.float32(scale), dtype=str(dtype).split(".")[-1],
                        num_kv_heads=np.int32(Hkv), block_size=np.int32(BS),
                        out_ref=bits(out.to(dtype)))
    print(os.path.basename(path), os.path.getsize(path))


if __name__ == "__main__":
    bf16, f16 = torch.bfloat16, torch.float16
    # (query_len, kv_len) pairs in the style of the reference's test (:83, :179-181): pure prefill,
    # chunked prefill with context, single-token decode and multi-token decode in one batch
    make_case("gqa_bf16_d128", [1, 5, 129, 70], [1328 // 4, 18, 463, 70], 8, 2, 128, 16, 32, bf16, 0)
    make_case("mha_f16_d64_bs32", [7, 1, 200], [7, 54, 293], 4, 4, 64, 32, 12, f16, 1)
    make_case("gqa8_bf16_d128_spec", [4, 4, 2, 8], [260, 33, 2, 100], 8, 1, 128, 16, 32, bf16, 2)
    make_case("gqa_f16_d256", [65, 3], [130, 67], 4, 2, 256, 16, 16, f16, 3)

"""Shared helpers for the tests (input generators, golden loaders)."""
from __future__ import annotations

import os
import random

import numpy as np
import torch

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

_DT = {"bfloat16": torch.bfloat16, "float16": torch.float16, "float32": torch.float32}


def from_bits(a: np.ndarray, dtype: torch.dtype) -> torch.Tensor:
    if dtype in (torch.bfloat16, torch.float16):
        return torch.from_numpy(a.view(np.int16).copy()).view(dtype)
    return torch.from_numpy(a.copy())


def load_golden(name: str) -> dict:
    z = np.load(os.path.join(GOLDEN, name), allow_pickle=False)
    return {k: z[k] for k in z.files}


def golden_dtype(g: dict) -> torch.dtype:
    return _DT[str(g["dtype"])]


def x_of(dtype: torch.dtype) -> int:
    return 16 // torch.tensor([], dtype=dtype).element_size()


def make_paged_inputs(B, H, Hkv, D, BS, NB, lens, dtype, seed, kv="auto",
                      lo=-1.0, hi=1.0, permute=False, k_scale=1.0, v_scale=1.0):
    """Seeded synthetic (query, key_cache, value_cache, block_tables, seq_lens).

    kv == "fp8": the caches are uint8 e4m3 codes drawn uniformly over the
    finite codes whose magnitude is <= hi / scale (no NaN codes).
    """
    g = torch.Generator().manual_seed(seed)
    rnd = random.Random(seed)
    q = (torch.rand(B, H, D, generator=g) * (hi - lo) + lo).to(dtype)
    if kv == "auto":
        x = x_of(dtype)
        kc = (torch.rand(NB, Hkv, D // x, BS, x, generator=g) * (hi - lo) + lo).to(dtype)
        vc = (torch.rand(NB, Hkv, D, BS, generator=g) * (hi - lo) + lo).to(dtype)
    else:
        x = 16
        # finite e4m3 codes: exponent/mantissa 0x00..0x7E, both signs; keep |v|<=2
        mag = torch.randint(0, 0x48, (NB, Hkv, D // x, BS, x), generator=g, dtype=torch.int32)
        sgn = torch.randint(0, 2, mag.shape, generator=g, dtype=torch.int32) << 7
        kc = (mag | sgn).to(torch.uint8)
        mag = torch.randint(0, 0x48, (NB, Hkv, D, BS), generator=g, dtype=torch.int32)
        sgn = torch.randint(0, 2, mag.shape, generator=g, dtype=torch.int32) << 7
        vc = (mag | sgn).to(torch.uint8)
    max_len = max(lens)
    mb = max((max_len + BS - 1) // BS, 1)
    if permute:
        perm = list(range(NB))
        rnd.shuffle(perm)
        rows, pos = [], 0
        for L in lens:
            nb = (L + BS - 1) // BS
            row = perm[pos:pos + nb]
            pos += nb
            rows.append(row + [0] * (mb - nb))
        assert pos <= NB, "not enough blocks for a non-aliased table"
    else:
        rows = [[rnd.randint(0, NB - 1) for _ in range(mb)] for _ in range(B)]
    bt = torch.tensor(rows, dtype=torch.int32)
    sl = torch.tensor(lens, dtype=torch.int32)
    return q, kc, vc, bt, sl


def torch_reference_attention(q, kc, vc, bt, sl, scale, num_kv_heads, alibi=None,
                              k_deq=None, v_deq=None):
    """fp64 torch restatement of softmax(q.K^T*scale+alibi).V on a paged cache.

    kc/vc: x-packed layouts; k_deq/v_deq: optional dequantised float caches
    (same logical shapes) when kc/vc hold fp8 codes.
    """
    B, H, D = q.shape
    kc_f = (k_deq if k_deq is not None else kc).double()
    vc_f = (v_deq if v_deq is not None else vc).double()
    NB, Hkv, Dx, BS, x = kc_f.shape
    k_log = kc_f.permute(0, 3, 1, 2, 4).reshape(NB, BS, Hkv, D)   # [NB,BS,Hkv,D]
    v_log = vc_f.permute(0, 3, 1, 2)                               # [NB,BS,Hkv,D]
    G = H // num_kv_heads
    out = torch.zeros(B, H, D, dtype=torch.float64)
    for b in range(B):
        L = int(sl[b])
        if L == 0:
            continue
        nb = (L + BS - 1) // BS
        blocks = bt[b, :nb].long()
        k = k_log[blocks].reshape(nb * BS, Hkv, D)[:L]
        v = v_log[blocks].reshape(nb * BS, Hkv, D)[:L]
        k = k.repeat_interleave(G, dim=1)
        v = v.repeat_interleave(G, dim=1)
        logits = scale * torch.einsum("hd,lhd->hl", q[b].double(), k)
        if alibi is not None:
            pos = torch.arange(L, dtype=torch.float64) - L + 1
            logits = logits + alibi.double().view(-1, 1) * pos.view(1, -1)
        p = torch.softmax(logits, dim=-1)
        out[b] = torch.einsum("hl,lhd->hd", p, v)
    return out


def e4m3_to_float(codes: torch.Tensor) -> torch.Tensor:
    """Exact value of e4m3 codes (table lookup via torch.float8_e4m3fn)."""
    return codes.view(torch.float8_e4m3fn).float()

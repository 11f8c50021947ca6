"""GPU parity of the KV-cache write / quantise ops against the CPU oracle.

Mirrors the reference's tests/kernels/attention/test_cache.py (:123-215
reshape_and_cache, :218-365 reshape_and_cache_flash, :449-480 fp8 round trip):
same shapes, slot sampling and scale choice, but the check is BIT-EXACT against
the oracle (and the committed golden fixtures) instead of a tolerance.
All calls go through the torch shim -> C ABI -> sm_100a kernels.
"""
import random

import numpy as np
import pytest
import torch

from tests import helpers as hp

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


@pytest.fixture(scope="module")
def ops():
    from vllm_main_modify_b200 import _custom_ops
    return _custom_ops


@pytest.fixture(scope="module")
def ora():
    from oracle import oracle
    return oracle


def _bytes(t: torch.Tensor) -> torch.Tensor:
    return t.contiguous().view(torch.uint8).cpu()


def _scales(key, value, kv):
    if kv == "auto":
        return torch.tensor(1.0), torch.tensor(1.0)
    # test_cache.py:165-166 of the reference: amax / 64
    return ((key.float().abs().max() / 64.0).reshape(()).float().cpu(),
            (value.float().abs().max() / 64.0).reshape(()).float().cpu())


@pytest.mark.parametrize("kv", ["auto", "fp8", "fp8_e5m2"])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("head_size", [64, 80, 128, 256])
@pytest.mark.parametrize("block_size", [8, 16, 32])
def test_reshape_and_cache_bit_exact(ops, ora, kv, dtype, head_size, block_size):
    T, Hkv, NB = 42, 8, 257
    x = 16 // (1 if kv != "auto" else torch.tensor([], dtype=dtype).element_size())
    if head_size % x:
        pytest.skip("head_size not divisible by x")
    g = torch.Generator().manual_seed(0)
    # strided like a fused qkv projection output (comment paged_attention_v1.cu:172)
    qkv = torch.randn(T, 3, Hkv, head_size, generator=g).to(dtype)
    key, value = qkv[:, 1], qkv[:, 2]
    slots = random.Random(0).sample(range(NB * block_size), T)
    slots[5] = -1
    slots[17] = -1
    slots = torch.tensor(slots, dtype=torch.int64)
    cdt = dtype if kv == "auto" else torch.uint8
    g2 = torch.Generator().manual_seed(1)
    kc0 = torch.randint(0, 255, (NB, Hkv, head_size // x, block_size, x), generator=g2,
                        dtype=torch.uint8)
    vc0 = torch.randint(0, 255, (NB, Hkv, head_size, block_size), generator=g2,
                        dtype=torch.uint8)
    if kv == "auto":
        es = torch.tensor([], dtype=dtype).element_size()
        kc0 = torch.zeros(NB, Hkv, head_size // x, block_size, x, dtype=dtype)
        vc0 = torch.zeros(NB, Hkv, head_size, block_size, dtype=dtype)
    ks, vs = _scales(key, value, kv)

    kc_o, vc_o = kc0.clone(), vc0.clone()
    ora.reshape_and_cache(key, value, kc_o, vc_o, slots, kv, ks, vs)

    qkv_d = qkv.to(DEV)
    kc_d, vc_d = kc0.to(DEV), vc0.to(DEV)
    ops.reshape_and_cache(qkv_d[:, 1], qkv_d[:, 2], kc_d, vc_d, slots.to(DEV), kv,
                          ks.to(DEV), vs.to(DEV))
    torch.cuda.synchronize()
    assert torch.equal(_bytes(kc_d), _bytes(kc_o))
    assert torch.equal(_bytes(vc_d), _bytes(vc_o))


@pytest.mark.parametrize("name", ["cache_bf16.npz", "cache_bf16_strided.npz", "cache_f16.npz"])
def test_reshape_and_cache_vs_reference_golden(ops, name):
    """Fixtures produced by the reference's own CPU reshape_and_cache."""
    g = hp.load_golden(name)
    dt = hp.golden_dtype(g)
    key, value = hp.from_bits(g["key"], dt), hp.from_bits(g["value"], dt)
    want_k, want_v = hp.from_bits(g["key_cache"], dt), hp.from_bits(g["value_cache"], dt)
    slots = torch.from_numpy(g["slot_mapping"].copy())
    if int(g["strided"]):
        T, Hkv, D = key.shape
        qkv = torch.zeros(T, 3, Hkv, D, dtype=dt)
        qkv[:, 1], qkv[:, 2] = key, value
        qkv = qkv.to(DEV)
        key_d, value_d = qkv[:, 1], qkv[:, 2]
    else:
        key_d, value_d = key.to(DEV), value.to(DEV)
    kc, vc = torch.zeros_like(want_k).to(DEV), torch.zeros_like(want_v).to(DEV)
    one = torch.tensor(1.0, device=DEV)
    ops.reshape_and_cache(key_d, value_d, kc, vc, slots.to(DEV), "auto", one, one)
    assert torch.equal(_bytes(kc), _bytes(want_k))
    assert torch.equal(_bytes(vc), _bytes(want_v))


@pytest.mark.parametrize("layout", ["NHD", "HND"])
@pytest.mark.parametrize("kv", ["auto", "fp8", "fp8_e5m2"])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("head_size", [64, 80, 256])
@pytest.mark.parametrize("block_size", [8, 16, 32])
def test_reshape_and_cache_flash_bit_exact(ops, ora, layout, kv, dtype, head_size, block_size):
    T, Hkv, NB = 42, 8, 129
    g = torch.Generator().manual_seed(2)
    # key has MORE rows than slot_mapping (CUDA-graph padding, cache_kernels.cu:468-478)
    key = torch.randn(T + 3, Hkv, head_size, generator=g).to(dtype)
    value = torch.randn(T + 3, Hkv, head_size, generator=g).to(dtype)
    slots = random.Random(1).sample(range(NB * block_size), T)
    slots[0] = -1
    slots = torch.tensor(slots, dtype=torch.int64)
    cdt = dtype if kv == "auto" else torch.uint8
    if layout == "NHD":
        base = torch.zeros(NB, 2, block_size, Hkv, head_size, dtype=cdt)
        view = lambda t: (t[:, 0], t[:, 1])
    else:
        base = torch.zeros(NB, 2, Hkv, block_size, head_size, dtype=cdt)
        view = lambda t: (t.permute(0, 1, 3, 2, 4)[:, 0], t.permute(0, 1, 3, 2, 4)[:, 1])
    ks, vs = _scales(key, value, kv)
    base_o = base.clone()
    kc_o, vc_o = view(base_o)
    ora.reshape_and_cache_flash(key, value, kc_o, vc_o, slots, kv, ks, vs)
    base_d = base.to(DEV)
    kc_d, vc_d = view(base_d)
    ops.reshape_and_cache_flash(key.to(DEV), value.to(DEV), kc_d, vc_d, slots.to(DEV), kv,
                                ks.to(DEV), vs.to(DEV))
    torch.cuda.synchronize()
    assert torch.equal(_bytes(base_d), _bytes(base_o))


def test_reshape_and_cache_flash_python_loop(ops):
    """The reference test's own expectation (test_cache.py:312-335)."""
    T, Hkv, D, BS, NB = 42, 8, 128, 16, 64
    dt = torch.bfloat16
    torch.manual_seed(0)
    key = torch.randn(T, Hkv, D, dtype=dt, device=DEV)
    value = torch.randn(T, Hkv, D, dtype=dt, device=DEV)
    slots = torch.tensor(random.Random(0).sample(range(NB * BS), T), dtype=torch.int64, device=DEV)
    kc = torch.zeros(NB, BS, Hkv, D, dtype=dt, device=DEV)
    vc = torch.zeros_like(kc)
    one = torch.tensor(1.0, device=DEV)
    ops.reshape_and_cache_flash(key, value, kc, vc, slots, "auto", one, one)
    wk, wv = torch.zeros_like(kc), torch.zeros_like(vc)
    for i, s in enumerate(slots.tolist()):
        wk[s // BS, s % BS] = key[i]
        wv[s // BS, s % BS] = value[i]
    assert torch.equal(kc, wk) and torch.equal(vc, wv)


def test_convert_fp8_every_bf16_and_f16_matches_toolkit_golden(ops):
    g = hp.load_golden("fp8_golden.npz")
    bf = (torch.arange(65536, dtype=torch.int32) << 16).view(torch.float32).to(torch.bfloat16)
    fh = torch.arange(65536, dtype=torch.int32).to(torch.int16).view(torch.float16)
    for src, key_e4, key_e5 in ((bf, "e4m3_from_bf16", "e5m2_from_bf16"),
                                (fh, "e4m3_from_f16", "e5m2_from_f16")):
        finite = torch.isfinite(src.float())
        for kvd, key in (("fp8", key_e4), ("fp8_e5m2", key_e5)):
            dst = torch.empty(65536, dtype=torch.uint8, device=DEV)
            ops.convert_fp8(dst, src.to(DEV), 1.0, kvd)
            want = torch.from_numpy(g[key].copy())
            got = dst.cpu()
            assert torch.equal(got[finite], want[finite]), key
    # non-unit scales (fp32 IEEE divide before the conversion)
    finite = torch.isfinite(bf.float())
    for j, s in enumerate(g["scales"]):
        dst = torch.empty(65536, dtype=torch.uint8, device=DEV)
        ops.convert_fp8(dst, bf.to(DEV), float(s), "fp8")
        want = torch.from_numpy(g["e4m3_scaled"][j].copy())
        assert torch.equal(dst.cpu()[finite], want[finite]), float(s)


def test_convert_fp8_decode_all_codes(ops, ora):
    codes = torch.arange(256, dtype=torch.int32).to(torch.uint8)
    for kvd in ("fp8", "fp8_e5m2"):
        for dt in (torch.bfloat16, torch.float16, torch.float32):
            for scale in (1.0, 0.37):
                out_d = torch.empty(256, dtype=dt, device=DEV)
                ops.convert_fp8(out_d, codes.to(DEV), scale, kvd)
                out_o = torch.empty(256, dtype=dt)
                ora.convert_fp8(out_o, codes, scale, kvd)
                a, b = out_d.cpu().float(), out_o.float()
                nan = torch.isnan(b)
                assert torch.equal(torch.isnan(a), nan)
                assert torch.equal(a[~nan], b[~nan]), (kvd, dt, scale)


def test_fp8_round_trip_like_reference(ops):
    """test_cache.py:449-480: U(-224, 224) -> fp8 -> back, atol 1e-3 rtol 0.1."""
    torch.manual_seed(0)
    for dt in (torch.bfloat16, torch.float16, torch.float32):
        x = torch.empty(4, 8, 128, 16, dtype=dt, device=DEV).uniform_(-224.0, 224.0)
        c = torch.empty_like(x, dtype=torch.uint8)
        ops.convert_fp8(c, x, 1.0, "fp8")
        y = torch.empty_like(x)
        ops.convert_fp8(y, c, 1.0, "fp8")
        torch.testing.assert_close(y, x, atol=1e-3, rtol=0.1)


def test_cache_write_is_idempotent_and_touches_only_its_slots(ops):
    T, Hkv, D, BS, NB = 64, 8, 128, 16, 96
    torch.manual_seed(3)
    key = torch.randn(T, Hkv, D, dtype=torch.bfloat16, device=DEV)
    value = torch.randn_like(key)
    slots = torch.tensor(random.Random(5).sample(range(NB * BS), T), dtype=torch.int64, device=DEV)
    kc = torch.full((NB, Hkv, D // 16, BS, 16), 0x5A, dtype=torch.uint8, device=DEV)
    vc = torch.full((NB, Hkv, D, BS), 0x5A, dtype=torch.uint8, device=DEV)
    ks = torch.tensor(0.05, device=DEV)
    vs = torch.tensor(0.07, device=DEV)
    ops.reshape_and_cache(key, value, kc, vc, slots, "fp8", ks, vs)
    k1, v1 = kc.clone(), vc.clone()
    ops.reshape_and_cache(key, value, kc, vc, slots, "fp8", ks, vs)
    assert torch.equal(kc, k1) and torch.equal(vc, v1)
    # untouched slots keep the fill pattern
    mask = torch.ones(NB * BS, dtype=torch.bool, device=DEV)
    mask[slots] = False
    v_by_slot = vc.permute(0, 3, 1, 2).reshape(NB * BS, -1)
    k_by_slot = kc.permute(0, 3, 1, 2, 4).reshape(NB * BS, -1)
    assert bool((v_by_slot[mask] == 0x5A).all()) and bool((k_by_slot[mask] == 0x5A).all())


def test_empty_and_all_skipped(ops):
    Hkv, D, BS, NB = 2, 64, 16, 4
    one = torch.tensor(1.0, device=DEV)
    kc = torch.zeros(NB, Hkv, D // 8, BS, 8, dtype=torch.bfloat16, device=DEV)
    vc = torch.zeros(NB, Hkv, D, BS, dtype=torch.bfloat16, device=DEV)
    key = torch.randn(0, Hkv, D, dtype=torch.bfloat16, device=DEV)
    ops.reshape_and_cache(key, key, kc, vc, torch.empty(0, dtype=torch.int64, device=DEV),
                          "auto", one, one)
    key = torch.randn(3, Hkv, D, dtype=torch.bfloat16, device=DEV)
    ops.reshape_and_cache(key, key, kc, vc, torch.full((3,), -1, dtype=torch.int64, device=DEV),
                          "auto", one, one)
    torch.cuda.synchronize()
    assert float(kc.float().abs().sum()) == 0.0 and float(vc.float().abs().sum()) == 0.0


def test_error_convention(ops):
    one = torch.tensor(1.0, device=DEV)
    key = torch.randn(2, 2, 64, dtype=torch.bfloat16, device=DEV)
    kc = torch.zeros(4, 2, 8, 16, 8, dtype=torch.bfloat16, device=DEV)
    vc = torch.zeros(4, 2, 64, 16, dtype=torch.bfloat16, device=DEV)
    slots = torch.tensor([0, 1], dtype=torch.int64, device=DEV)
    with pytest.raises(RuntimeError, match="Unsupported data type of kv cache"):
        ops.reshape_and_cache(key, key, kc, vc, slots, "int8", one, one)
    with pytest.raises(RuntimeError):
        ops.convert_fp8(torch.empty(4, dtype=torch.uint8, device=DEV),
                        torch.empty(4, dtype=torch.bfloat16, device=DEV), 1.0, "auto")


@pytest.mark.parametrize("kv", ["auto", "fp8", "fp8_e5m2"])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("head_size,block_size", [(128, 16), (64, 32), (256, 16), (128, 8), (80, 16),
                                                  (256, 32)])
def test_reshape_and_cache_prefill_sized_batches_bit_exact(ops, ora, kv, dtype, head_size,
                                                           block_size):
    """>= 1024 tokens take the tile kernel (64 tokens x one kv head per CTA, V rows staged in
    shared memory and stored by 16-byte groups when the tile is a run of consecutive positions):
    prefill chunks at every slot offset, decode tokens, padding slots and sequence boundaries
    inside a tile -- the same bytes as the reference's per-token kernel
    (csrc/cache_kernels.cu:212-264)."""
    x = 16 // (1 if kv != "auto" else 2)
    if head_size % x:
        pytest.skip("head_size not divisible by x")
    Hkv = 3
    rnd = random.Random(head_size + block_size)
    # (kind, length, first position): chunks start at arbitrary positions of their sequence
    parts = [("chunk", 300, 0), ("chunk", 77, 5), ("decode", 10, 0), ("chunk", 500, block_size - 1),
             ("pad", 7, 0), ("chunk", 64, 64), ("chunk", 129, 9), ("decode", 3, 0), ("chunk", 200, 3)]
    slots, nblocks = [], 0
    for kind, n, first in parts:
        if kind == "pad":
            slots += [-1] * n
        elif kind == "decode":
            for _ in range(n):
                slots.append(nblocks * block_size + rnd.randrange(block_size))
                nblocks += 1
        else:
            nb = (first + n + block_size - 1) // block_size
            blocks = list(range(nblocks, nblocks + nb))
            rnd.shuffle(blocks)                      # physical blocks in any order
            nblocks += nb
            slots += [blocks[p // block_size] * block_size + p % block_size
                      for p in range(first, first + n)]
    T, NB = len(slots), nblocks + 1
    assert T >= 1024
    g = torch.Generator().manual_seed(3)
    qkv = torch.randn(T, 3, Hkv, head_size, generator=g).to(dtype)
    key, value = qkv[:, 1], qkv[:, 2]
    slots = torch.tensor(slots, dtype=torch.int64)
    if kv == "auto":
        kc0 = torch.zeros(NB, Hkv, head_size // x, block_size, x, dtype=dtype)
        vc0 = torch.zeros(NB, Hkv, head_size, block_size, dtype=dtype)
    else:
        g2 = torch.Generator().manual_seed(1)
        kc0 = torch.randint(0, 255, (NB, Hkv, head_size // x, block_size, x), generator=g2,
                            dtype=torch.uint8)
        vc0 = torch.randint(0, 255, (NB, Hkv, head_size, block_size), generator=g2,
                            dtype=torch.uint8)
    ks, vs = _scales(key, value, kv)
    kc_o, vc_o = kc0.clone(), vc0.clone()
    ora.reshape_and_cache(key, value, kc_o, vc_o, slots, kv, ks, vs)
    qkv_d = qkv.to(DEV)
    kc_d, vc_d = kc0.to(DEV), vc0.to(DEV)
    ops.reshape_and_cache(qkv_d[:, 1], qkv_d[:, 2], kc_d, vc_d, slots.to(DEV), kv,
                          ks.to(DEV), vs.to(DEV))
    torch.cuda.synchronize()
    assert torch.equal(_bytes(kc_d), _bytes(kc_o))
    assert torch.equal(_bytes(vc_d), _bytes(vc_o))

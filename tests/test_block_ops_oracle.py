"""CPU checks of oracle/cache_ops_oracle.py through round trips: what reshape_and_cache_flash
(oracle, pinned) scatters into an NHD cache, the gather oracles must bring back in order."""
import random

import torch

from oracle import cache_ops_oracle as co
from oracle import oracle


def _scattered_cache(dtype, kv, BS=16, Hkv=2, D=32, lens=(0, 5, 16, 37), NB=12, seed=0):
    g = torch.Generator().manual_seed(seed)
    rnd = random.Random(seed)
    T = sum(lens)
    key = torch.randn(T, Hkv, D, generator=g).to(dtype)
    cache_dt = dtype if kv == "auto" else torch.uint8
    kc = torch.zeros(NB, BS, Hkv, D, dtype=cache_dt)
    vc = torch.zeros_like(kc)
    nblk = [(l + BS - 1) // BS for l in lens]
    perm = list(range(NB))
    rnd.shuffle(perm)
    bt = torch.zeros(len(lens), max(max(nblk), 1), dtype=torch.int32)
    slots, pos = [], 0
    for b, l in enumerate(lens):
        blocks = perm[pos:pos + nblk[b]]
        pos += nblk[b]
        bt[b, :nblk[b]] = torch.tensor(blocks, dtype=torch.int32)
        slots += [blocks[p // BS] * BS + p % BS for p in range(l)]
    scale = torch.tensor(0.05)
    oracle.reshape_and_cache_flash(key, key, kc, vc, torch.tensor(slots, dtype=torch.int64), kv,
                                   scale, scale)
    cu = torch.tensor([0] + list(torch.tensor(lens).cumsum(0)), dtype=torch.int32)
    return key, kc, bt, cu, scale


def test_gather_auto_round_trip():
    key, kc, bt, cu, scale = _scattered_cache(torch.bfloat16, "auto")
    dst = torch.zeros(key.shape[0], key.shape[1] * key.shape[2], dtype=torch.bfloat16)
    co.gather_and_maybe_dequant_cache(kc, dst, bt, cu, 4, "auto", scale)
    assert torch.equal(dst, key.flatten(1))
    dst2 = torch.zeros_like(dst)
    co.cp_gather_cache(kc, dst2, bt, cu, 4)
    assert torch.equal(dst2, dst)


def test_gather_fp8_matches_convert_fp8_of_the_gathered_bytes():
    key, kc, bt, cu, scale = _scattered_cache(torch.float16, "fp8")
    raw = torch.zeros(key.shape[0], key.shape[1] * key.shape[2], dtype=torch.uint8)
    co.cp_gather_cache(kc, raw, bt, cu, 4)
    want = torch.empty(raw.shape, dtype=torch.float16)
    oracle.convert_fp8(want, raw, float(scale), "fp8")
    dst = torch.zeros_like(want)
    co.gather_and_maybe_dequant_cache(kc, dst, bt, cu, 4, "fp8", scale)
    assert torch.equal(dst.view(torch.int16), want.view(torch.int16))
    # and dequantised values are the quantisation of the originals: close to them
    torch.testing.assert_close(dst.float(), key.flatten(1).float(), atol=0.05 * 0.07 * 16, rtol=0.07)


def test_seq_starts_semantics_differ_between_the_two_gathers():
    key, kc, bt, cu, scale = _scattered_cache(torch.float32, "auto", lens=(40, 33), NB=8)
    E = key.shape[1] * key.shape[2]
    # take 10 tokens of each sequence starting at token 16 (one whole block in)
    cu2 = torch.tensor([0, 10, 20], dtype=torch.int32)
    starts = torch.tensor([16, 16], dtype=torch.int32)
    a = torch.zeros(20, E)
    co.gather_and_maybe_dequant_cache(kc, a, bt, cu2, 2, "auto", scale, starts)
    b = torch.zeros(20, E)
    co.cp_gather_cache(kc, b, bt, cu2, 2, starts)
    want = torch.cat([key[16:26], key[40 + 16:40 + 26]]).flatten(1)
    assert torch.equal(a, want) and torch.equal(b, want)
    # a start that is not a multiple of the block size: only cp_gather_cache honours the remainder
    starts = torch.tensor([19, 3], dtype=torch.int32)
    co.cp_gather_cache(kc, b, bt, cu2, 2, starts)
    assert torch.equal(b, torch.cat([key[19:29], key[40 + 3:40 + 13]]).flatten(1))
    co.gather_and_maybe_dequant_cache(kc, a, bt, cu2, 2, "auto", scale, starts)
    assert torch.equal(a, torch.cat([key[16:26], key[40:50]]).flatten(1))   # 19 // 16, 3 // 16 blocks


def test_copy_and_swap_blocks():
    g = torch.Generator().manual_seed(1)
    kcs = [torch.randn(10, 4, 8, generator=g) for _ in range(3)]
    vcs = [torch.randn(10, 4, 8, generator=g) for _ in range(3)]
    k0 = [k.clone() for k in kcs]
    mapping = torch.tensor([[0, 5], [0, 6], [3, 9]], dtype=torch.int64)
    co.copy_blocks(kcs, vcs, mapping)
    for k, o in zip(kcs, k0):
        assert torch.equal(k[5], o[0]) and torch.equal(k[6], o[0]) and torch.equal(k[9], o[3])
        assert torch.equal(k[1], o[1])
    src, dst = torch.randn(6, 20, generator=g), torch.zeros(8, 20)
    co.swap_blocks(src, dst, torch.tensor([[1, 7], [4, 0]], dtype=torch.int64))
    assert torch.equal(dst[7], src[1]) and torch.equal(dst[0], src[4]) and (dst[1:7] == 0).all()


def test_layout_converters_pinned_on_the_two_cache_write_oracles():
    """reshape_and_cache (x-packed) and reshape_and_cache_flash (token-major) are both pinned
    against the reference (tests/test_oracle.py): writing the same tokens through either and
    converting must give the same bytes -- which pins xpacked_to_flash / flash_to_xpacked, the
    re-indexing the flash-layout attention tests rely on."""
    import pytest
    for dtype, kv in ((torch.bfloat16, "auto"), (torch.float16, "auto"), (torch.bfloat16, "fp8")):
        g = torch.Generator().manual_seed(3)
        NB, BS, Hkv, D, T = 7, 16, 3, 64, 60
        cache_dt = dtype if kv == "auto" else torch.uint8
        x = 16 // torch.tensor([], dtype=cache_dt).element_size()
        key = torch.randn(T, Hkv, D, generator=g).to(dtype)
        value = torch.randn(T, Hkv, D, generator=g).to(dtype)
        slots = torch.randperm(NB * BS, generator=g)[:T].to(torch.int64)
        ks, vs = torch.tensor(0.03), torch.tensor(0.04)
        kx = torch.zeros(NB, Hkv, D // x, BS, x, dtype=cache_dt)
        vx = torch.zeros(NB, Hkv, D, BS, dtype=cache_dt)
        oracle.reshape_and_cache(key, value, kx, vx, slots, kv, ks, vs)
        kf = torch.zeros(NB, BS, Hkv, D, dtype=cache_dt)
        vf = torch.zeros_like(kf)
        oracle.reshape_and_cache_flash(key, value, kf, vf, slots, kv, ks, vs)
        kf2, vf2 = co.xpacked_to_flash(kx, vx)
        assert torch.equal(kf2.view(torch.uint8), kf.view(torch.uint8))
        assert torch.equal(vf2.view(torch.uint8), vf.view(torch.uint8))
        kx2, vx2 = co.flash_to_xpacked(kf, vf)
        assert torch.equal(kx2.view(torch.uint8), kx.view(torch.uint8))
        assert torch.equal(vx2.view(torch.uint8), vx.view(torch.uint8))
        # HND-strided views convert the same
        kh = kf.permute(0, 2, 1, 3).contiguous().permute(0, 2, 1, 3)
        vh = vf.permute(0, 2, 1, 3).contiguous().permute(0, 2, 1, 3)
        kx3, vx3 = co.flash_to_xpacked(kh, vh)
        assert torch.equal(kx3.view(torch.uint8), kx.view(torch.uint8))
        assert torch.equal(vx3.view(torch.uint8), vx.view(torch.uint8))

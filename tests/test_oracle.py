"""Pin the CPU oracle (oracle/paged_attn_oracle.c) before anything trusts it.

Checked against (i) golden vectors made by the CUDA toolkit's own fp8/bf16/f16
conversions, (ii) golden outputs of the reference's own test oracle and CPU
kernels (tests/golden/make_golden.py), (iii) oracle/_ref live when present,
(iv) an independent fp64 torch restatement.  CPU only.
"""
import numpy as np
import pytest
import torch

from oracle import oracle as ora
from oracle import ref_cpu
from tests import helpers as hp

ONE = torch.tensor(1.0)


# ------------------------------------------------------------------ fp8 bytes

@pytest.fixture(scope="module")
def fp8_golden():
    return hp.load_golden("fp8_golden.npz")


def _all_bf16_as_f32():
    return (torch.arange(65536, dtype=torch.int32) << 16).view(torch.float32)


def test_oracle_e4m3_e5m2_from_every_bf16(fp8_golden):
    lib = ora.lib()
    xs = _all_bf16_as_f32().numpy()
    for kind, key in ((ora.KV_E4M3, "e4m3_from_bf16"), (ora.KV_E5M2, "e5m2_from_bf16")):
        got = np.array([lib.ora_f32_to_fp8(float(x), kind) for x in xs], dtype=np.uint8)
        want = fp8_golden[key]
        bad = np.nonzero(got != want)[0]
        assert bad.size == 0, (key, bad[:8], got[bad[:8]], want[bad[:8]])


def test_oracle_e4m3_from_every_f16(fp8_golden):
    lib = ora.lib()
    xs = torch.arange(65536, dtype=torch.int32).to(torch.int16).view(torch.float16).float().numpy()
    for kind, key in ((ora.KV_E4M3, "e4m3_from_f16"), (ora.KV_E5M2, "e5m2_from_f16")):
        got = np.array([lib.ora_f32_to_fp8(float(x), kind) for x in xs], dtype=np.uint8)
        assert np.array_equal(got, fp8_golden[key]), key


def test_oracle_fp8_decode_all_codes(fp8_golden):
    lib = ora.lib()
    for kind, key in ((ora.KV_E4M3, "half_from_e4m3"), (ora.KV_E5M2, "half_from_e5m2")):
        half_bits = fp8_golden[key]
        want = torch.from_numpy(half_bits.view(np.int16).copy()).view(torch.float16).float()
        got = torch.tensor([lib.ora_fp8_to_f32(i, kind) for i in range(256)])
        nan = torch.isnan(want)
        assert torch.equal(torch.isnan(got), nan)
        assert torch.equal(got[~nan], want[~nan])
    # and torch's own table for e4m3
    t = hp.e4m3_to_float(torch.arange(256, dtype=torch.int32).to(torch.uint8))
    got = torch.tensor([lib.ora_fp8_to_f32(i, ora.KV_E4M3) for i in range(256)])
    nan = torch.isnan(t)
    assert torch.equal(got[~nan], t[~nan])


def test_oracle_bf16_f16_rounding(fp8_golden):
    lib = ora.lib()
    n = 65536 * 4
    idx = np.arange(0, n, 37)  # a 1/37 sample keeps ctypes overhead small
    hi = (idx // 4).astype(np.uint32) << 16
    lo = np.array([0x0000, 0x7FFF, 0x8000, 0x8001], dtype=np.uint32)[idx % 4]
    xs = (hi | lo).view(np.float32)
    for fn, key in ((lib.ora_f32_to_bf16, "bf16_from_f32"), (lib.ora_f32_to_f16, "f16_from_f32")):
        want = fp8_golden[key][idx]
        got = np.array([fn(float(x)) for x in xs], dtype=np.uint16)
        isnan = np.isnan(xs)
        assert np.array_equal(got[~isnan], want[~isnan]), key


def test_oracle_scaled_quantise_matches_toolkit(fp8_golden):
    """convert_fp8 with a non-unit scale: fp8(float(x) / scale), fp32 divide."""
    src = _all_bf16_as_f32().to(torch.bfloat16)
    finite = torch.isfinite(src.float())
    for j, s in enumerate(fp8_golden["scales"]):
        dst = torch.empty(65536, dtype=torch.uint8)
        ora.convert_fp8(dst, src, float(s), "fp8")
        want = torch.from_numpy(fp8_golden["e4m3_scaled"][j].copy())
        assert torch.equal(dst[finite], want[finite]), float(s)


def test_oracle_quantise_agrees_with_torch_float8_in_range():
    """torch's e4m3fn cast is RNE without saturation: equal where |x| <= 448."""
    g = torch.Generator().manual_seed(0)
    x = (torch.rand(20000, generator=g) - 0.5) * 800
    x = x[(x.abs() <= 448)]
    dst = torch.empty(x.numel(), dtype=torch.uint8)
    ora.convert_fp8(dst, x.contiguous(), 1.0, "fp8")
    want = x.to(torch.float8_e4m3fn).view(torch.uint8)
    assert torch.equal(dst, want)


def test_oracle_convert_roundtrip_like_reference_test():
    """tests/kernels/attention/test_cache.py:449-480 (fp8 round trip)."""
    g = torch.Generator().manual_seed(0)
    x = ((torch.rand(4096, generator=g) - 0.5) * 448).to(torch.bfloat16)
    c = torch.empty(4096, dtype=torch.uint8)
    ora.convert_fp8(c, x, 1.0, "fp8")
    y = torch.empty_like(x)
    ora.convert_fp8(y, c, 1.0, "fp8")
    torch.testing.assert_close(y, x, atol=1e-3, rtol=0.1)


# ------------------------------------------------------------ attention gold

ATTN_GOLDEN = ["attn_gqa_bf16_d128.npz", "attn_mha_bf16_d64_alibi.npz",
               "attn_gqa_f16_d80.npz", "attn_gqa_bf16_d128_bs32.npz",
               "attn_mha_f32_d32.npz"]


def _load_attn(name):
    g = hp.load_golden(name)
    dt = hp.golden_dtype(g)
    q = hp.from_bits(g["q"], dt)
    kc = hp.from_bits(g["key_cache"], dt)
    vc = hp.from_bits(g["value_cache"], dt)
    bt = torch.from_numpy(g["block_tables"].copy())
    sl = torch.from_numpy(g["seq_lens"].copy())
    alibi = torch.from_numpy(g["alibi_slopes"].copy()) if "alibi_slopes" in g else None
    return g, dt, q, kc, vc, bt, sl, alibi


def _tol(dt):
    return (1e-4, 1e-4) if dt == torch.float32 else (1e-2, 1e-2)


@pytest.mark.parametrize("name", ATTN_GOLDEN)
@pytest.mark.parametrize("version", ["v1", "v2"])
@pytest.mark.parametrize("faithful", [True, False])
def test_oracle_attention_vs_golden(name, version, faithful):
    g, dt, q, kc, vc, bt, sl, alibi = _load_attn(name)
    Hkv, BS, scale = int(g["num_kv_heads"]), int(g["block_size"]), float(g["scale"])
    B, H, D = q.shape
    max_len = int(sl.max())
    out = torch.empty_like(q)
    if version == "v1":
        ora.paged_attention_v1(out, q, kc, vc, Hkv, scale, bt, sl, BS, max_len,
                               alibi, "auto", ONE, ONE, faithful=faithful)
    else:
        P = (max_len + 511) // 512
        es, ml = torch.zeros(B, H, P), torch.zeros(B, H, P)
        tmp = torch.zeros(B, H, P, D, dtype=dt)
        ora.paged_attention_v2(out, es, ml, tmp, q, kc, vc, Hkv, scale, bt, sl,
                               BS, max_len, alibi, "auto", ONE, ONE, faithful=faithful)
    atol, rtol = _tol(dt)
    want = hp.from_bits(g["out_test_ref"], dt)
    torch.testing.assert_close(out.float(), want.float(), atol=atol, rtol=rtol)
    for key in ("out_cpu_v1", "out_cpu_v2"):
        if key in g:
            torch.testing.assert_close(out.float(), hp.from_bits(g[key], dt).float(),
                                       atol=atol, rtol=rtol)
    # independent fp64 restatement
    ref64 = hp.torch_reference_attention(q, kc, vc, bt, sl, scale, Hkv, alibi)
    torch.testing.assert_close(out.double(), ref64, atol=atol, rtol=rtol)


@pytest.mark.skipif(not ref_cpu.available(), reason="oracle/_ref not built")
@pytest.mark.parametrize("version", ["v1", "v2"])
def test_oracle_attention_vs_live_reference_cpu(version):
    """BASELINE config 1 shape at reduced batch, against the reference's CPU op."""
    B, H, Hkv, D, BS, NB = 3, 32, 32, 128, 16, 3 * 64 + 7
    lens = [1024, 1000, 513]
    q, kc, vc, bt, sl = hp.make_paged_inputs(B, H, Hkv, D, BS, NB, lens,
                                             torch.bfloat16, seed=0, permute=True)
    scale = D ** -0.5
    o_ref, o_ora = torch.empty_like(q), torch.empty_like(q)
    if version == "v1":
        ref_cpu.paged_attention_v1(o_ref, q, kc, vc, Hkv, scale, bt, sl, BS, 1024,
                                   None, "auto", ONE, ONE, 0, 0, 0, 64, 0)
        ora.paged_attention_v1(o_ora, q, kc, vc, Hkv, scale, bt, sl, BS, 1024,
                               None, "auto", ONE, ONE)
    else:
        P = 2
        bufs = [(torch.zeros(B, H, P), torch.zeros(B, H, P),
                 torch.zeros(B, H, P, D, dtype=torch.bfloat16)) for _ in range(2)]
        ref_cpu.paged_attention_v2(o_ref, *bufs[0], q, kc, vc, Hkv, scale, bt, sl,
                                   BS, 1024, None, "auto", ONE, ONE, 0, 0, 0, 64, 0)
        ora.paged_attention_v2(o_ora, *bufs[1], q, kc, vc, Hkv, scale, bt, sl,
                               BS, 1024, None, "auto", ONE, ONE)
    torch.testing.assert_close(o_ora.float(), o_ref.float(), atol=1e-2, rtol=1e-2)


def test_oracle_fp8_attention_matches_dequantised_fp64():
    """fp8 cache + non-unit scales against fp64 math on the dequantised cache
    (what the reference test does with convert_fp8, test_attention.py:303-319)."""
    B, H, Hkv, D, BS, NB = 3, 8, 2, 128, 16, 40
    lens = [3, 200, 530]
    q, kc, vc, bt, sl = hp.make_paged_inputs(B, H, Hkv, D, BS, NB, lens,
                                             torch.bfloat16, seed=5, kv="fp8")
    ks, vs = 0.37, 1.9
    kd = hp.e4m3_to_float(kc) * ks
    vd = hp.e4m3_to_float(vc) * vs
    scale = D ** -0.5
    want = hp.torch_reference_attention(q, kc, vc, bt, sl, scale, Hkv, None, kd, vd)
    for faithful in (True, False):
        out = torch.empty_like(q)
        ora.paged_attention_v1(out, q, kc, vc, Hkv, scale, bt, sl, BS, 530, None,
                               "fp8", torch.tensor(ks), torch.tensor(vs),
                               faithful=faithful)
        torch.testing.assert_close(out.double(), want, atol=1e-2, rtol=1e-2)


def test_oracle_blocksparse_equals_masked_dense():
    """Block-sparse semantics (attention_kernels.cuh:204-251, 382-387): a skipped
    block behaves like -inf logits."""
    B, H, Hkv, D, BS, NB = 2, 4, 2, 64, 16, 60
    lens = [300, 513]
    q, kc, vc, bt, sl = hp.make_paged_inputs(B, H, Hkv, D, BS, NB, lens,
                                             torch.float32, seed=7, permute=True)
    scale = D ** -0.5
    local, vert, bsz, step = 2, 3, 32, 1
    out = torch.empty_like(q)
    ora.paged_attention_v1(out, q, kc, vc, Hkv, scale, bt, sl, BS, 513, None, "auto",
                           ONE, ONE, 0, local, vert, bsz, step, faithful=False)
    # dense restatement with a mask
    x = 4
    k_log = kc.permute(0, 3, 1, 2, 4).reshape(NB, BS, Hkv, D).double()
    v_log = vc.permute(0, 3, 1, 2).double()
    for b in range(B):
        L = lens[b]
        nb = (L + BS - 1) // BS
        blocks = bt[b, :nb].long()
        k = k_log[blocks].reshape(nb * BS, Hkv, D)[:L]
        v = v_log[blocks].reshape(nb * BS, Hkv, D)[:L]
        for h in range(H):
            kvh = h // (H // Hkv)
            logits = scale * (k[:, kvh] @ q[b, h].double())
            qb = (L - 1) // bsz
            off = (0 * H + h) * step + 1
            tok = torch.arange(L)
            kb = (tok // BS) * BS // bsz
            keep = (((kb + off) % vert) == 0) | (kb > qb - local)
            logits = torch.where(keep, logits, torch.tensor(-float("inf"), dtype=torch.float64))
            p = torch.softmax(logits, dim=0)
            want = p @ v[:, kvh]
            torch.testing.assert_close(out[b, h].double(), want, atol=1e-4, rtol=1e-4)


# ---------------------------------------------------------------- cache gold

@pytest.mark.parametrize("name", ["cache_bf16.npz", "cache_bf16_strided.npz", "cache_f16.npz"])
def test_oracle_reshape_and_cache_vs_golden(name):
    g = hp.load_golden(name)
    dt = hp.golden_dtype(g)
    key, value = hp.from_bits(g["key"], dt), hp.from_bits(g["value"], dt)
    want_k, want_v = hp.from_bits(g["key_cache"], dt), hp.from_bits(g["value_cache"], dt)
    slots = torch.from_numpy(g["slot_mapping"].copy())
    if int(g["strided"]):
        T, Hkv, D = key.shape
        qkv = torch.zeros(T, 3, Hkv, D, dtype=dt)
        qkv[:, 1], qkv[:, 2] = key, value
        key, value = qkv[:, 1], qkv[:, 2]
    kc, vc = torch.zeros_like(want_k), torch.zeros_like(want_v)
    ora.reshape_and_cache(key, value, kc, vc, slots, "auto", ONE, ONE)
    assert torch.equal(kc.view(torch.int16), want_k.view(torch.int16))
    assert torch.equal(vc.view(torch.int16), want_v.view(torch.int16))


def test_oracle_reshape_and_cache_python_loop_and_skip():
    """The reference test's own check (test_cache.py:181-196): token i lands in
    key_cache[block, :, :, off, :] / value_cache[block, :, :, off]; slot<0 skipped."""
    T, Hkv, D, BS, NB = 10, 3, 64, 16, 5
    dt = torch.bfloat16
    x = hp.x_of(dt)
    g = torch.Generator().manual_seed(3)
    key = torch.randn(T, Hkv, D, generator=g).to(dt)
    value = torch.randn(T, Hkv, D, generator=g).to(dt)
    slots = torch.tensor([5, 17, -1, 79, 0, 33, -1, 64, 2, 48], dtype=torch.int64)
    kc = torch.zeros(NB, Hkv, D // x, BS, x, dtype=dt)
    vc = torch.zeros(NB, Hkv, D, BS, dtype=dt)
    ora.reshape_and_cache(key, value, kc, vc, slots, "auto", ONE, ONE)
    wk, wv = torch.zeros_like(kc), torch.zeros_like(vc)
    rk = key.reshape(T, Hkv, D // x, x)
    for i in range(T):
        s = int(slots[i])
        if s < 0:
            continue
        wk[s // BS, :, :, s % BS, :] = rk[i]
        wv[s // BS, :, :, s % BS] = value[i]
    assert torch.equal(kc, wk) and torch.equal(vc, wv)


@pytest.mark.parametrize("layout", ["NHD", "HND"])
@pytest.mark.parametrize("kv", ["auto", "fp8"])
def test_oracle_reshape_and_cache_flash(layout, kv):
    """test_cache.py:218-365: NHD / HND by strides; fp8 = quantise(x / scale)."""
    T, Hkv, D, BS, NB = 9, 2, 64, 16, 4
    dt = torch.bfloat16
    g = torch.Generator().manual_seed(4)
    key = torch.randn(T, Hkv, D, generator=g).to(dt)
    value = torch.randn(T, Hkv, D, generator=g).to(dt)
    slots = torch.tensor([3, 20, 63, -1, 8, 41, 17, 1, 30], dtype=torch.int64)
    cdt = dt if kv == "auto" else torch.uint8
    if layout == "NHD":
        kvc = torch.zeros(NB, 2, BS, Hkv, D, dtype=cdt)
    else:
        kvc = torch.zeros(NB, 2, Hkv, BS, D, dtype=cdt).permute(0, 1, 3, 2, 4)
    kc, vc = kvc[:, 0], kvc[:, 1]
    ks = torch.tensor(float(key.float().abs().max()) / 64.0)
    vs = torch.tensor(float(value.float().abs().max()) / 64.0)
    ora.reshape_and_cache_flash(key, value, kc, vc, slots, kv, ks, vs)
    for i in range(T):
        s = int(slots[i])
        if s < 0:
            continue
        if kv == "auto":
            assert torch.equal(kc[s // BS, s % BS], key[i])
            assert torch.equal(vc[s // BS, s % BS], value[i])
        else:
            wk = (key[i].float() / ks).to(torch.float8_e4m3fn).view(torch.uint8)
            wv = (value[i].float() / vs).to(torch.float8_e4m3fn).view(torch.uint8)
            assert torch.equal(kc[s // BS, s % BS], wk)
            assert torch.equal(vc[s // BS, s % BS], wv)
    # untouched slots stay zero
    used = {int(s) for s in slots if s >= 0}
    for s in range(NB * BS):
        if s not in used:
            assert int(kc[s // BS, s % BS].view(torch.uint8 if kv != "auto" else torch.int16).abs().sum()) == 0


@pytest.mark.skipif(not ref_cpu.available(), reason="oracle/_ref not built")
def test_oracle_reshape_and_cache_vs_live_reference_cpu():
    T, Hkv, D, BS, NB = 42, 8, 128, 16, 32
    dt = torch.bfloat16
    x = hp.x_of(dt)
    g = torch.Generator().manual_seed(9)
    qkv = torch.randn(T, 3, Hkv, D, generator=g).to(dt)
    key, value = qkv[:, 1], qkv[:, 2]
    import random
    slots = torch.tensor(random.Random(0).sample(range(NB * BS), T), dtype=torch.int64)
    a = [torch.zeros(NB, Hkv, D // x, BS, x, dtype=dt), torch.zeros(NB, Hkv, D, BS, dtype=dt)]
    b = [torch.zeros_like(a[0]), torch.zeros_like(a[1])]
    ref_cpu.reshape_and_cache(key, value, a[0], a[1], slots, "auto", ONE, ONE)
    ora.reshape_and_cache(key, value, b[0], b[1], slots, "auto", ONE, ONE)
    assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])

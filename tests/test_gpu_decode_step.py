"""GPU parity of ``decode_step`` (cache write + attention in one launch).

The contract is "the effect of reshape_and_cache followed by paged_attention_v1": the cache
bytes must be IDENTICAL to reshape_and_cache's, and the output must be bitwise the output of
paged_attention_v1 run on the updated cache by the same kernel family (same schedule, same data),
plus the oracle gate (1e-2 abs + 1e-2 rel) against the CPU oracle doing the same two steps.
The new token's slot in the cache is poisoned before the call (NaN codes), so an implementation
that attends to what the TMA copy found in the cache instead of this step's key/value fails.
"""
import random

import pytest
import torch

from tests import helpers as hp

pytestmark = pytest.mark.gpu

DEV = "cuda:0"
ATOL, RTOL = 1e-2, 1e-2
GENERIC, STREAM = 1, 2
BS = 16


@pytest.fixture(scope="module")
def ops():
    from vllm_main_modify_b200 import _custom_ops
    return _custom_ops


@pytest.fixture(scope="module")
def native():
    from vllm_main_modify_b200 import _native
    yield _native
    _native.set_impl(_native.IMPL_AUTO)


@pytest.fixture(scope="module")
def ora():
    from oracle import oracle
    return oracle


_RND = random.Random(5)
CASES = [
    # name, B, H, Hkv, D, lens (INCLUDING the new token), dtype, kv, alibi
    ("gqa8_fp8_ragged", 6, 8, 1, 128, [128, 40000, 5000, 777, 16384 + 3, 20001], torch.bfloat16, "fp8", False),
    ("gqa8_fp8_page_edges", 8, 64, 8, 128, [1, 16, 17, 32, 33, 511, 512, 513], torch.bfloat16, "fp8", False),
    ("gqa4_bf16_page_edges", 8, 32, 8, 128, [1, 16, 17, 32, 33, 511, 512, 513], torch.bfloat16, "auto", False),
    ("gqa16_two_tiles_fp8", 5, 32, 2, 128, [1000, 17, 512, 640, 3333], torch.bfloat16, "fp8", False),
    ("mha_f16_d64_alibi", 7, 40, 40, 64, [_RND.randint(1, 3000) for _ in range(7)], torch.float16, "auto", True),
    ("gqa16_fp8_d64_alibi", 7, 32, 2, 64, [_RND.randint(1, 3000) for _ in range(7)], torch.float16, "fp8", True),
    ("many_short_fp8", 300, 16, 2, 128, [_RND.randint(1, 700) for _ in range(300)], torch.bfloat16, "fp8", False),
    ("one_long_fp8", 1, 64, 8, 128, [131072], torch.bfloat16, "fp8", False),
    ("llama70b_shard_fp8", 16, 8, 1, 128, [32768] * 16, torch.bfloat16, "fp8", False),
    ("gemma_d256_fp8", 5, 16, 8, 256, [1000, 17, 512, 4100, 3333], torch.bfloat16, "fp8", False),
    ("gemma_d256_bf16", 5, 16, 8, 256, [1000, 17, 512, 4100, 3333], torch.bfloat16, "auto", False),
    ("d192_f16_fp8", 4, 24, 3, 192, [700, 33, 2049, 16], torch.float16, "fp8", False),
    ("phi_d96_bf16", 6, 32, 32, 96, [1, 16, 17, 999, 2048, 31], torch.bfloat16, "auto", False),
    ("d96_fp8_gqa12", 4, 24, 2, 96, [300, 5000, 12, 640], torch.bfloat16, "fp8", False),
    ("d32_fp8", 4, 8, 8, 32, [300, 1500, 12, 640], torch.float16, "fp8", False),
    ("gqa8_e5m2_page_edges", 8, 64, 8, 128, [1, 16, 17, 32, 33, 511, 512, 513], torch.bfloat16, "fp8_e5m2", False),
    ("gqa4_e5m2_d64_f16_alibi", 7, 32, 8, 64, [_RND.randint(1, 3000) for _ in range(7)], torch.float16, "fp8_e5m2", True),
    ("e5m2_ragged_long", 4, 16, 2, 128, [40000, 129, 5000, 16384 + 3], torch.bfloat16, "fp8_e5m2", False),
    ("d80_fp8_f16", 6, 32, 32, 80, [1, 16, 17, 999, 2048, 31], torch.float16, "fp8", False),
    ("d112_fp8_gqa4", 5, 16, 4, 112, [1000, 17, 512, 4100, 3333], torch.bfloat16, "fp8", True),
    ("d80_e5m2_f16", 6, 32, 32, 80, [1, 16, 17, 999, 2048, 31], torch.float16, "fp8_e5m2", False),
    ("d256_e5m2_bf16", 5, 16, 8, 256, [1000, 17, 512, 4100, 3333], torch.bfloat16, "fp8_e5m2", False),
    ("neox_d80_f16", 6, 32, 32, 80, [1, 16, 17, 999, 2048, 31], torch.float16, "auto", False),
    ("d112_bf16_gqa4", 5, 16, 4, 112, [1000, 17, 512, 4100, 3333], torch.bfloat16, "auto", True),
]


def _inputs(B, H, Hkv, D, lens, dtype, kv, seed=7, BS=BS):
    s = D ** -0.5
    NB = sum((L + BS - 1) // BS for L in lens) + 5
    q, kc, vc, bt, sl = hp.make_paged_inputs(B, H, Hkv, D, BS, NB, lens, dtype, seed=seed, kv=kv,
                                             lo=-1.0, hi=1.0, permute=True)
    g = torch.Generator().manual_seed(seed + 1)
    key = (torch.rand(B, Hkv, D, generator=g) * 2 - 1).to(dtype)
    value = (torch.rand(B, Hkv, D, generator=g) * 2 - 1).to(dtype)
    slots = torch.tensor([int(bt[b, (L - 1) // BS]) * BS + (L - 1) % BS
                          for b, L in enumerate(lens)], dtype=torch.int64)
    return s, q, kc, vc, bt, sl, key, value, slots


def _poison(kc, vc, slots, kv, BS=BS):
    """NaN in the slots the step is about to fill (fp8: code 0x7F; 16-bit: a NaN pattern)."""
    blk, off = slots // BS, slots % BS
    live = slots >= 0
    blk, off = blk[live], off[live]
    if kv == "auto":
        nan = torch.tensor(float("nan"), dtype=kc.dtype)
        kc[blk, :, :, off, :] = nan
        vc[blk, :, :, off] = nan
    else:
        kc[blk, :, :, off, :] = 0x7F
        vc[blk, :, :, off] = 0x7F


def _scales(kv):
    return (1.0, 1.0) if kv == "auto" else (0.5, 1.3)


def _two_ops(ops, q, key, value, kc, vc, slots, Hkv, s, bt, sl, max_len, alibi, kv, ks, vs,
             lse=None, BS=BS):
    ops.reshape_and_cache(key, value, kc, vc, slots, kv, ks, vs)
    out = torch.full_like(q, float("nan"))
    if lse is None:
        ops.paged_attention_v1(out, q, kc, vc, Hkv, s, bt, sl, BS, max_len, alibi, kv, ks, vs)
    else:
        ops.paged_attention_lse(out, lse, q, kc, vc, Hkv, s, bt, sl, BS, max_len, alibi, kv, ks, vs)
    return out


@pytest.mark.parametrize("impl", [STREAM, GENERIC])
@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_decode_step_equals_write_then_attend(ops, native, ora, case, impl):
    name, B, H, Hkv, D, lens, dtype, kv, use_alibi = case
    s, q, kc, vc, bt, sl, key, value, slots = _inputs(B, H, Hkv, D, lens, dtype, kv)
    _poison(kc, vc, slots, kv)
    ks, vs = _scales(kv)
    alibi = None
    if use_alibi:
        alibi = torch.tensor([2.0 ** (-(i + 1) * 8.0 / H) for i in range(H)], dtype=torch.float32)
    max_len = max(lens)
    d = lambda t: None if t is None else t.to(DEV)
    qd, btd, sld, keyd, vald, slotsd, ald = d(q), d(bt), d(sl), d(key), d(value), d(slots), d(alibi)
    ksd, vsd = torch.tensor(ks, device=DEV), torch.tensor(vs, device=DEV)
    native.set_impl(impl)
    try:
        kc1, vc1 = d(kc), d(vc)
        want = _two_ops(ops, qd, keyd, vald, kc1, vc1, slotsd, Hkv, s, btd, sld, max_len, ald, kv,
                        ksd, vsd)
        kc2, vc2 = d(kc), d(vc)
        got = torch.full_like(qd, float("nan"))
        n0 = native.launch_count()
        ops.decode_step(got, None, qd, keyd, vald, kc2, vc2, slotsd, Hkv, s, btd, sld, BS, max_len,
                        ald, kv, ksd, vsd)
        launches = native.launch_count() - n0
        torch.cuda.synchronize()
    finally:
        native.set_impl(native.IMPL_AUTO)
    assert torch.equal(kc2.view(torch.uint8), kc1.view(torch.uint8)), "key cache bytes"
    assert torch.equal(vc2.view(torch.uint8), vc1.view(torch.uint8)), "value cache bytes"
    assert not torch.isnan(got.float()).any()
    assert torch.equal(got, want), "decode_step != reshape_and_cache + paged_attention_v1"
    if impl == STREAM:
        assert launches <= 2, launches      # stream kernel (+ boundary merge): no cache-write launch
    if sum(lens) * H <= 64 * 140000:
        kc_o, vc_o = kc.clone(), vc.clone()
        ora.reshape_and_cache(key, value, kc_o, vc_o, slots, kv, torch.tensor(ks), torch.tensor(vs))
        assert torch.equal(kc2.cpu().view(torch.uint8), kc_o.view(torch.uint8))
        assert torch.equal(vc2.cpu().view(torch.uint8), vc_o.view(torch.uint8))
        ref = torch.empty_like(q)
        ora.paged_attention_v1(ref, q, kc_o, vc_o, Hkv, s, bt, sl, BS, max_len, alibi, kv,
                               torch.tensor(ks), torch.tensor(vs))
        atol = ATOL * max(1.0, float(ref.float().abs().max()))
        torch.testing.assert_close(got.cpu().float(), ref.float(), atol=atol, rtol=RTOL)


BS32_CASES = [c for c in CASES if c[0] in ("gqa8_fp8_page_edges", "gqa4_bf16_page_edges",
                                            "gqa16_two_tiles_fp8", "mha_f16_d64_alibi",
                                            "gemma_d256_fp8", "gemma_d256_bf16", "d192_f16_fp8",
                                            "phi_d96_bf16", "many_short_fp8", "neox_d80_f16",
                                            "d112_bf16_gqa4", "d80_fp8_f16", "d112_fp8_gqa4")]


@pytest.mark.parametrize("case", BS32_CASES, ids=[c[0] for c in BS32_CASES])
def test_decode_step_block_size_32(ops, native, ora, case):
    """The same contract on 32-token cache blocks (two 16-token tiles per block in the kernel)."""
    name, B, H, Hkv, D, lens, dtype, kv, use_alibi = case
    bs = 32
    s, q, kc, vc, bt, sl, key, value, slots = _inputs(B, H, Hkv, D, lens, dtype, kv, BS=bs)
    _poison(kc, vc, slots, kv, BS=bs)
    ks, vs = _scales(kv)
    alibi = None
    if use_alibi:
        alibi = torch.tensor([2.0 ** (-(i + 1) * 8.0 / H) for i in range(H)], dtype=torch.float32)
    max_len = max(lens)
    d = lambda t: None if t is None else t.to(DEV)
    qd, btd, sld, keyd, vald, slotsd, ald = d(q), d(bt), d(sl), d(key), d(value), d(slots), d(alibi)
    ksd, vsd = torch.tensor(ks, device=DEV), torch.tensor(vs, device=DEV)
    native.set_impl(STREAM)
    try:
        kc1, vc1 = d(kc), d(vc)
        want = _two_ops(ops, qd, keyd, vald, kc1, vc1, slotsd, Hkv, s, btd, sld, max_len, ald, kv,
                        ksd, vsd, BS=bs)
        kc2, vc2 = d(kc), d(vc)
        got = torch.full_like(qd, float("nan"))
        n0 = native.launch_count()
        ops.decode_step(got, None, qd, keyd, vald, kc2, vc2, slotsd, Hkv, s, btd, sld, bs, max_len,
                        ald, kv, ksd, vsd)
        assert native.launch_count() - n0 <= 2
        torch.cuda.synchronize()
    finally:
        native.set_impl(native.IMPL_AUTO)
    assert torch.equal(kc2.view(torch.uint8), kc1.view(torch.uint8)), "key cache bytes"
    assert torch.equal(vc2.view(torch.uint8), vc1.view(torch.uint8)), "value cache bytes"
    assert not torch.isnan(got.float()).any()
    assert torch.equal(got, want)
    kc_o, vc_o = kc.clone(), vc.clone()
    ora.reshape_and_cache(key, value, kc_o, vc_o, slots, kv, torch.tensor(ks), torch.tensor(vs))
    assert torch.equal(kc2.cpu().view(torch.uint8), kc_o.view(torch.uint8))
    assert torch.equal(vc2.cpu().view(torch.uint8), vc_o.view(torch.uint8))
    ref = torch.empty_like(q)
    ora.paged_attention_v1(ref, q, kc_o, vc_o, Hkv, s, bt, sl, bs, max_len, alibi, kv,
                           torch.tensor(ks), torch.tensor(vs))
    atol = ATOL * max(1.0, float(ref.float().abs().max()))
    torch.testing.assert_close(got.cpu().float(), ref.float(), atol=atol, rtol=RTOL)


@pytest.mark.parametrize("kv", ["auto", "fp8"])
def test_decode_step_skips_negative_slots(ops, native, kv):
    """slot < 0: the token is neither stored nor attended to (what the two ops do)."""
    B, H, Hkv, D = 5, 32, 8, 128
    lens = [100, 17, 512, 640, 33]
    s, q, kc, vc, bt, sl, key, value, slots = _inputs(B, H, Hkv, D, lens, torch.bfloat16, kv)
    slots[1] = -1
    slots[3] = -1
    ks, vs = _scales(kv)
    d = lambda t: t.to(DEV)
    ksd, vsd = torch.tensor(ks, device=DEV), torch.tensor(vs, device=DEV)
    native.set_impl(STREAM)
    try:
        kc1, vc1 = d(kc), d(vc)
        want = _two_ops(ops, d(q), d(key), d(value), kc1, vc1, d(slots), Hkv, s, d(bt), d(sl),
                        max(lens), None, kv, ksd, vsd)
        kc2, vc2 = d(kc), d(vc)
        got = torch.full_like(want, float("nan"))
        ops.decode_step(got, None, d(q), d(key), d(value), kc2, vc2, d(slots), Hkv, s, d(bt), d(sl),
                        BS, max(lens), None, kv, ksd, vsd)
        torch.cuda.synchronize()
    finally:
        native.set_impl(native.IMPL_AUTO)
    assert torch.equal(kc2.view(torch.uint8), kc1.view(torch.uint8))
    assert torch.equal(vc2.view(torch.uint8), vc1.view(torch.uint8))
    assert torch.equal(got, want)


@pytest.mark.parametrize("kv", ["auto", "fp8"])
def test_decode_step_lse_and_strided_key_value(ops, native, kv):
    """key / value as slices of one fused qkv row (the model's layout), lse requested; and a
    2-byte-misaligned key, which the fused kernel does not take (16-byte loads): the entry point
    then launches the two kernels, same result."""
    B, H, Hkv, D = 6, 32, 8, 128
    lens = [1, 16, 17, 900, 512, 513]
    dt = torch.bfloat16
    s, q, kc, vc, bt, sl, key, value, slots = _inputs(B, H, Hkv, D, lens, dt, kv, seed=9)
    _poison(kc, vc, slots, kv)
    ks, vs = _scales(kv)
    d = lambda t: t.to(DEV)
    ksd, vsd = torch.tensor(ks, device=DEV), torch.tensor(vs, device=DEV)
    qkv = torch.zeros(B, (H + 2 * Hkv) * D, dtype=dt, device=DEV)
    qkv[:, :H * D] = d(q).view(B, -1)
    qkv[:, H * D:(H + Hkv) * D] = d(key).view(B, -1)
    qkv[:, (H + Hkv) * D:] = d(value).view(B, -1)
    qv = qkv[:, :H * D].view(B, H, D)
    kview = qkv[:, H * D:(H + Hkv) * D].view(B, Hkv, D)
    vview = qkv[:, (H + Hkv) * D:].view(B, Hkv, D)
    flat = torch.zeros(B * Hkv * D + 1, dtype=dt, device=DEV)
    flat[1:] = d(key).flatten()
    kmis = flat[1:].view(B, Hkv, D)
    assert kmis.data_ptr() % 16 != 0
    native.set_impl(STREAM)
    try:
        kc1, vc1 = d(kc), d(vc)
        lse1 = torch.empty(B, H, dtype=torch.float32, device=DEV)
        want = _two_ops(ops, qv, kview, vview, kc1, vc1, d(slots), Hkv, s, d(bt), d(sl), max(lens),
                        None, kv, ksd, vsd, lse=lse1)
        for kk in (kview, kmis):
            kc2, vc2 = d(kc), d(vc)
            lse2 = torch.empty(B, H, dtype=torch.float32, device=DEV)
            got = torch.full_like(want, float("nan"))
            n0 = native.launch_count()
            ops.decode_step(got, lse2, qv, kk, vview, kc2, vc2, d(slots), Hkv, s, d(bt), d(sl), BS,
                            max(lens), None, kv, ksd, vsd)
            launches = native.launch_count() - n0
            torch.cuda.synchronize()
            assert launches == (2 if kk is kmis else 1), launches   # (write +) ONE decode kernel
            assert torch.equal(kc2.view(torch.uint8), kc1.view(torch.uint8))
            assert torch.equal(vc2.view(torch.uint8), vc1.view(torch.uint8))
            assert torch.equal(got, want)
            assert torch.equal(lse2, lse1)
    finally:
        native.set_impl(native.IMPL_AUTO)


def test_decode_step_graph_replay_advances_sequences(ops, native):
    """A captured decode_step replays with seq_lens / slot_mapping / key advanced on the device:
    three consecutive decode steps through one graph equal three eager two-op steps."""
    B, H, Hkv, D = 4, 64, 8, 128
    lens0 = [15, 30, 511, 1000]
    steps = 3
    final = [L + steps for L in lens0]
    dt = torch.bfloat16
    s, q, kc, vc, bt, _, _, _, _ = _inputs(B, H, Hkv, D, final, dt, "fp8", seed=13)
    ks, vs = torch.tensor(0.5, device=DEV), torch.tensor(1.3, device=DEV)
    d = lambda t: t.to(DEV)
    qd, btd = d(q), d(bt)
    g = torch.Generator().manual_seed(3)
    keys = (torch.rand(steps, B, Hkv, D, generator=g) * 2 - 1).to(dt).to(DEV)
    vals = (torch.rand(steps, B, Hkv, D, generator=g) * 2 - 1).to(dt).to(DEV)

    def slots_of(lens):
        return torch.tensor([int(bt[b, (L - 1) // BS]) * BS + (L - 1) % BS
                             for b, L in enumerate(lens)], dtype=torch.int64, device=DEV)

    native.set_impl(STREAM)
    try:
        # eager, two ops per step
        kc1, vc1 = d(kc), d(vc)
        wants = []
        for t in range(steps):
            lens = [L + t + 1 for L in lens0]
            sl = torch.tensor(lens, dtype=torch.int32, device=DEV)
            wants.append(_two_ops(ops, qd, keys[t], vals[t], kc1, vc1, slots_of(lens), Hkv, s, btd,
                                  sl, max(final), None, "fp8", ks, vs))
        # one graph, device-side state
        kc2, vc2 = d(kc), d(vc)
        sl_g = torch.zeros(B, dtype=torch.int32, device=DEV)
        slots_g = torch.zeros(B, dtype=torch.int64, device=DEV)
        key_g, val_g = torch.zeros_like(keys[0]), torch.zeros_like(vals[0])
        out_g = torch.empty_like(qd)

        def load(t):
            lens = [L + t + 1 for L in lens0]
            sl_g.copy_(torch.tensor(lens, dtype=torch.int32))
            slots_g.copy_(slots_of(lens))
            key_g.copy_(keys[t])
            val_g.copy_(vals[t])

        def step():
            ops.decode_step(out_g, None, qd, key_g, val_g, kc2, vc2, slots_g, Hkv, s, btd, sl_g, BS,
                            max(final), None, "fp8", ks, vs)

        scratch_k, scratch_v = kc2.clone(), vc2.clone()
        load(0)
        stream = torch.cuda.Stream()
        stream.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(stream):
            step()                       # warm-up (writes step 0's token: idempotent)
            stream.synchronize()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=stream):
                step()
        torch.cuda.synchronize()
        kc2.copy_(scratch_k)
        vc2.copy_(scratch_v)
        for t in range(steps):
            load(t)
            graph.replay()
            torch.cuda.synchronize()
            assert torch.equal(out_g, wants[t]), f"step {t}"
        assert torch.equal(kc2, kc1) and torch.equal(vc2, vc1)
    finally:
        native.set_impl(native.IMPL_AUTO)


def test_decode_step_opcheck(ops):
    B, H, Hkv, D = 3, 16, 2, 128
    lens = [5, 40, 17]
    s, q, kc, vc, bt, sl, key, value, slots = _inputs(B, H, Hkv, D, lens, torch.bfloat16, "fp8")
    d = lambda t: t.to(DEV)
    one = torch.tensor(1.0, device=DEV)
    torch.library.opcheck(
        torch.ops.b200pa.decode_step,
        (torch.empty_like(d(q)), None, d(q), d(key), d(value), d(kc), d(vc), d(slots), Hkv, s,
         d(bt), d(sl), BS, max(lens), None, "fp8", one, one),
        test_utils=("test_schema", "test_faketensor"))

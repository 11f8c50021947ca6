"""GPU parity of ``paged_attention_flash``: the decode kernel reading the
token-major cache that ``reshape_and_cache_flash`` writes, NHD and HND.

The flash cache is produced from the x-packed test cache by pure re-indexing
(oracle/cache_ops_oracle.py::xpacked_to_flash, pinned on the two cache-write oracles in
tests/test_block_ops_oracle.py), so the CPU oracle and the x-packed GPU kernel see THE SAME
values: the flash kernel must match the oracle within the north_star gate (1e-2 abs + 1e-2 rel),
be bitwise equal to the x-packed stream kernel for 16-bit caches (same fragments, same order) and
within a couple of output ulps of it for fp8 caches (token order inside an MMA tile differs).
The fused cache write must leave the bytes ``reshape_and_cache_flash`` leaves and the output of
"write, then attend".
"""
import random

import pytest
import torch

from tests import helpers as hp

pytestmark = pytest.mark.gpu

DEV = "cuda:0"
ATOL, RTOL = 1e-2, 1e-2
STREAM = 2
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


@pytest.fixture(scope="module")
def co():
    from oracle import cache_ops_oracle
    return cache_ops_oracle


_RND = random.Random(8)
CASES = [
    # name, B, H, Hkv, D, lens, dtype, kv, alibi
    ("gqa8_fp8_ragged", 6, 8, 1, 128, [128, 40000, 5000, 777, 16384 + 3, 20001], torch.bfloat16, "fp8", False),
    ("gqa8_fp8_page_edges", 8, 64, 8, 128, [1, 16, 17, 32, 33, 511, 512, 513], torch.bfloat16, "fp8", False),
    ("gqa4_bf16_page_edges", 8, 32, 8, 128, [1, 16, 17, 32, 33, 511, 512, 513], torch.bfloat16, "auto", False),
    ("gqa16_two_tiles_fp8", 5, 32, 2, 128, [1000, 17, 512, 640, 3333], torch.bfloat16, "fp8", False),
    ("gqa5_bf16", 3, 10, 2, 128, [1, 513, 1025], torch.bfloat16, "auto", False),
    ("mha_f16_d64_alibi", 7, 40, 40, 64, [_RND.randint(1, 3000) for _ in range(7)], torch.float16, "auto", True),
    ("gqa16_fp8_d64_alibi", 7, 32, 2, 64, [_RND.randint(1, 3000) for _ in range(7)], torch.float16, "fp8", True),
    ("f16_fp8_d128", 4, 16, 4, 128, [77, 1500, 16, 4097], torch.float16, "fp8", False),
    ("many_short_fp8", 300, 16, 2, 128, [_RND.randint(1, 700) for _ in range(300)], torch.bfloat16, "fp8", False),
    ("with_empties_fp8", 9, 8, 1, 128, [0, 600, 0, 0, 31, 5000, 0, 512, 0], torch.bfloat16, "fp8", False),
    ("one_long_bf16", 1, 8, 1, 128, [70001], torch.bfloat16, "auto", False),
    ("llama70b_shard_fp8", 16, 8, 1, 128, [32768] * 16, torch.bfloat16, "fp8", False),
    ("gemma_d256_fp8", 5, 16, 8, 256, [1000, 17, 512, 4100, 3333], torch.bfloat16, "fp8", False),
    ("gemma_d256_bf16", 5, 16, 8, 256, [1000, 17, 512, 4100, 3333], torch.bfloat16, "auto", False),
    ("d192_f16_fp8", 4, 24, 3, 192, [700, 33, 2049, 16], torch.float16, "fp8", False),
    ("phi_d96_bf16", 6, 32, 32, 96, [1, 16, 17, 999, 2048, 31], torch.bfloat16, "auto", False),
    ("d96_fp8_gqa12", 4, 24, 2, 96, [300, 5000, 12, 640], torch.bfloat16, "fp8", False),
    ("d32_fp8", 4, 8, 8, 32, [300, 1500, 12, 640], torch.float16, "fp8", False),
    ("d32_bf16", 4, 8, 2, 32, [300, 5000, 12, 640], torch.bfloat16, "auto", False),
    # head sizes 80 / 112 (rows padded to an odd number of 16-byte units), fp8_e5m2, and -- tenth
    # field -- 32-token cache blocks (reshape_and_cache_flash writes all of these,
    # csrc/cache_kernels.cu:280-353 of the reference)
    ("neox_d80_bf16", 4, 8, 8, 80, [300, 1500, 12, 640], torch.bfloat16, "auto", False),
    ("d80_fp8", 4, 8, 8, 80, [300, 1500, 12, 640], torch.bfloat16, "fp8", False),
    ("d112_f16_gqa4", 4, 16, 4, 112, [1, 16, 17, 2049], torch.float16, "auto", False),
    ("d112_fp8_gqa4", 4, 16, 4, 112, [300, 5000, 12, 640], torch.bfloat16, "fp8", False),
    ("e5m2_d128_gqa4", 5, 32, 8, 128, [1, 16, 17, 999, 2048], torch.bfloat16, "fp8_e5m2", False),
    ("e5m2_d64_f16", 4, 8, 8, 64, [300, 1500, 12, 640], torch.float16, "fp8_e5m2", False),
    ("bs32_fp8_gqa8", 6, 16, 2, 128, [1, 16, 32, 33, 1000, 4097], torch.bfloat16, "fp8", False, 32),
    ("bs32_bf16_gqa4", 6, 16, 4, 128, [1, 16, 32, 33, 1000, 4097], torch.bfloat16, "auto", False, 32),
    ("bs32_d64_f16_alibi", 5, 8, 8, 64, [31, 64, 65, 700, 3000], torch.float16, "auto", True, 32),
    ("bs32_d256_fp8", 4, 16, 8, 256, [1000, 17, 512, 4100], torch.bfloat16, "fp8", False, 32),
    ("bs32_d80_e5m2", 4, 8, 8, 80, [300, 1500, 12, 640], torch.bfloat16, "fp8_e5m2", False, 32),
    # several kv heads per GPU: in NHD order these take the whole-block kernel
    # (csrc/paged_attention_nhd.cuh: tensor-map TMA, one warp per head, CTA spans)
    ("nhd_long_ragged_fp8", 6, 64, 8, 128, [128, 40000, 5000, 777, 16384 + 3, 20001], torch.bfloat16, "fp8", False),
    ("nhd_long_bf16_gqa4_alibi", 4, 16, 4, 128, [9000, 1, 33000, 513], torch.bfloat16, "auto", True),
    ("nhd_empties_f16_fp8", 9, 16, 2, 128, [0, 600, 0, 0, 31, 5000, 0, 512, 0], torch.float16, "fp8", False),
    ("nhd_mha8_f16", 3, 8, 8, 128, [70, 4096, 15], torch.float16, "auto", False),
]


def _takes_block_kernel(layout, H, Hkv, D, kv, bs):
    """An even number of kv heads, head size 128, 16-token blocks, group <= 8: the x-packed cache (the
    comparison's other side) and the NHD cache take the whole-block kernel, the HND cache the
    per-warp kernel -- other span cuts, hence another merge order: no bitwise equality."""
    return (D == 128 and bs == 16 and Hkv % 2 == 0 and H // Hkv <= 8
            and kv in ("auto", "fp8"))


def _unpack(case):
    name, B, H, Hkv, D, lens, dtype, kv, use_alibi, *rest = case
    return name, B, H, Hkv, D, lens, dtype, kv, use_alibi, (rest[0] if rest else BS)


def _layout(t, layout):
    """NHD contiguous, or the same logical [NB,BS,Hkv,D] tensor stored HND."""
    if layout == "NHD":
        return t.contiguous()
    return t.permute(0, 2, 1, 3).contiguous().permute(0, 2, 1, 3)


def _inputs(co, B, H, Hkv, D, lens, dtype, kv, seed=17, bs=BS):
    s = D ** -0.5
    NB = sum((L + bs - 1) // bs for L in lens) + 5
    # (the generator's fp8 codes, magnitudes below 0x48, are finite in e4m3 and in e5m2)
    q, kc, vc, bt, sl = hp.make_paged_inputs(B, H, Hkv, D, bs, NB, lens, dtype, seed=seed,
                                             kv="auto" if kv == "auto" else "fp8",
                                             lo=-1.0, hi=1.0, permute=True)
    kf, vf = co.xpacked_to_flash(kc, vc)
    return s, q, kc, vc, kf, vf, bt, sl


def _scales(kv):
    return (1.0, 1.0) if kv == "auto" else (0.5, 1.3)


@pytest.mark.parametrize("layout", ["NHD", "HND"])
@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_flash_layout_attention(ops, native, ora, co, case, layout):
    name, B, H, Hkv, D, lens, dtype, kv, use_alibi, bs = _unpack(case)
    s, q, kc, vc, kf, vf, bt, sl = _inputs(co, B, H, Hkv, D, lens, dtype, kv, bs=bs)
    ks, vs = _scales(kv)
    alibi = None
    if use_alibi:
        alibi = torch.tensor([2.0 ** (-(i + 1) * 8.0 / H) for i in range(H)], dtype=torch.float32)
    max_len = max(max(lens), 1)
    live = torch.tensor([L > 0 for L in lens])
    d = lambda t: None if t is None else t.to(DEV)
    qd, btd, sld, ald = d(q), d(bt), d(sl), d(alibi)
    ksd, vsd = torch.tensor(ks, device=DEV), torch.tensor(vs, device=DEV)
    kfd, vfd = _layout(d(kf), layout), _layout(d(vf), layout)
    got = torch.full_like(qd, float("nan"))
    lse = torch.full((B, H), float("nan"), dtype=torch.float32, device=DEV)
    n0 = native.launch_count()
    ops.paged_attention_flash(got, lse, qd, None, None, kfd, vfd, None, s, btd, sld, max_len, ald,
                              kv, ksd, vsd)
    assert native.launch_count() - n0 in (1, 2)
    # the x-packed stream kernel on the same values
    native.set_impl(STREAM)
    try:
        ref = torch.full_like(qd, float("nan"))
        ref_lse = torch.full_like(lse, float("nan"))
        ops.paged_attention_lse(ref, ref_lse, qd, d(kc), d(vc), Hkv, s, btd, sld, bs, max_len, ald,
                                kv, ksd, vsd)
        torch.cuda.synchronize()
    finally:
        native.set_impl(native.IMPL_AUTO)
    got, ref, lse, ref_lse = got.cpu(), ref.cpu(), lse.cpu(), ref_lse.cpu()
    assert not torch.isnan(got[live].float()).any()
    if kv == "auto" and not _takes_block_kernel(layout, H, Hkv, D, kv, bs):
        assert torch.equal(got[live], ref[live]), "16-bit cache: same fragments, same order"
        assert torch.equal(lse[live], ref_lse[live])
    else:
        # fp8: the token order inside an MMA tile differs; whole-block kernel: other span cuts,
        # hence another merge order
        scale_out = max(1.0, float(ref[live].float().abs().max()))
        torch.testing.assert_close(got[live].float(), ref[live].float(), atol=2 ** -7 * scale_out,
                                   rtol=2 ** -6)
        torch.testing.assert_close(lse[live], ref_lse[live], atol=1e-4, rtol=1e-5)
    if not bool(live.all()):
        assert float(got[~live].float().abs().sum()) == 0.0
        assert bool(torch.isinf(lse[~live]).all())
    if sum(lens) * H <= 64 * 140000:
        want = torch.empty_like(q)
        ora.paged_attention_v1(want, q, kc, vc, Hkv, s, bt, sl, bs, max_len, alibi, kv,
                               torch.tensor(ks), torch.tensor(vs))
        atol = ATOL * max(1.0, float(want[live].float().abs().max()))
        torch.testing.assert_close(got[live].float(), want[live].float(), atol=atol, rtol=RTOL)


@pytest.mark.parametrize("layout", ["NHD", "HND"])
@pytest.mark.parametrize("case", [c for c in CASES if min(c[5]) >= 1],
                         ids=[c[0] for c in CASES if min(c[5]) >= 1])
def test_flash_layout_fused_write(ops, native, ora, co, case, layout):
    """key / value / slot_mapping given: reshape_and_cache_flash + attention in one launch."""
    name, B, H, Hkv, D, lens, dtype, kv, use_alibi, bs = _unpack(case)
    s, q, kc, vc, kf, vf, bt, sl = _inputs(co, B, H, Hkv, D, lens, dtype, kv, bs=bs)
    ks, vs = _scales(kv)
    alibi = None
    if use_alibi:
        alibi = torch.tensor([2.0 ** (-(i + 1) * 8.0 / H) for i in range(H)], dtype=torch.float32)
    max_len = max(lens)
    g = torch.Generator().manual_seed(4)
    key = (torch.rand(B, Hkv, D, generator=g) * 2 - 1).to(dtype)
    value = (torch.rand(B, Hkv, D, generator=g) * 2 - 1).to(dtype)
    slots = torch.tensor([int(bt[b, (L - 1) // bs]) * bs + (L - 1) % bs
                          for b, L in enumerate(lens)], dtype=torch.int64)
    # poison the slots about to be written (NaN codes)
    blk, off = slots // bs, slots % bs
    if kv == "auto":
        kf[blk, off] = float("nan")
        vf[blk, off] = float("nan")
    else:
        kf[blk, off] = 0x7F
        vf[blk, off] = 0x7F
    d = lambda t: None if t is None else t.to(DEV)
    qd, btd, sld, ald, keyd, vald, slotsd = d(q), d(bt), d(sl), d(alibi), d(key), d(value), d(slots)
    ksd, vsd = torch.tensor(ks, device=DEV), torch.tensor(vs, device=DEV)
    # two ops
    k1, v1 = _layout(d(kf), layout), _layout(d(vf), layout)
    ops.reshape_and_cache_flash(keyd, vald, k1, v1, slotsd, kv, ksd, vsd)
    want = torch.full_like(qd, float("nan"))
    ops.paged_attention_flash(want, None, qd, None, None, k1, v1, None, s, btd, sld, max_len, ald,
                              kv, ksd, vsd)
    # one launch
    k2, v2 = _layout(d(kf), layout), _layout(d(vf), layout)
    got = torch.full_like(qd, float("nan"))
    n0 = native.launch_count()
    ops.paged_attention_flash(got, None, qd, keyd, vald, k2, v2, slotsd, s, btd, sld, max_len, ald,
                              kv, ksd, vsd)
    assert native.launch_count() - n0 <= 2
    torch.cuda.synchronize()
    assert torch.equal(k2.contiguous().view(torch.uint8), k1.contiguous().view(torch.uint8))
    assert torch.equal(v2.contiguous().view(torch.uint8), v1.contiguous().view(torch.uint8))
    assert not torch.isnan(got.float()).any()
    assert torch.equal(got, want)
    # and the cache bytes are the oracle's
    if sum(lens) * Hkv <= 200000:
        ko, vo = kf.clone(), vf.clone()
        ora.reshape_and_cache_flash(key, value, ko, vo, slots, kv, torch.tensor(ks), torch.tensor(vs))
        assert torch.equal(k2.cpu().contiguous().view(torch.uint8), ko.view(torch.uint8))
        assert torch.equal(v2.cpu().contiguous().view(torch.uint8), vo.view(torch.uint8))


def test_flash_layout_error_convention(ops):
    B, H, Hkv, BS_, NB = 2, 8, 2, 16, 8
    one = torch.tensor(1.0, device=DEV)
    bt = torch.zeros(B, 2, dtype=torch.int32, device=DEV)
    sl = torch.tensor([5, 20], dtype=torch.int32, device=DEV)
    for D, bs, dt in ((120, 16, torch.bfloat16), (128, 8, torch.bfloat16), (128, 16, torch.float32)):
        q = torch.zeros(B, H, D, dtype=dt, device=DEV)
        kc = torch.zeros(NB, bs, Hkv, D, dtype=dt, device=DEV)
        with pytest.raises(RuntimeError, match="Unsupported|needs a 16-bit query"):
            ops.paged_attention_flash(torch.empty_like(q), None, q, None, None, kc, kc.clone(), None,
                                      0.1, bt, sl, 20, None, "auto", one, one)
    q = torch.zeros(B, H, 128, dtype=torch.bfloat16)
    with pytest.raises(RuntimeError):      # CPU tensors: no CPU path
        ops.paged_attention_flash(torch.empty_like(q), None, q, None, None,
                                  torch.zeros(NB, 16, Hkv, 128, dtype=torch.bfloat16),
                                  torch.zeros(NB, 16, Hkv, 128, dtype=torch.bfloat16), None, 0.1,
                                  bt.cpu(), sl.cpu(), 20, None, "auto", one.cpu(), one.cpu())


def test_flash_layout_opcheck(ops):
    B, H, Hkv, D, NB = 3, 16, 2, 128, 12
    g = torch.Generator().manual_seed(0)
    q = torch.randn(B, H, D, generator=g).to(torch.bfloat16).to(DEV)
    kc = torch.randint(0, 0x48, (NB, BS, Hkv, D), generator=g, dtype=torch.int32).to(torch.uint8).to(DEV)
    vc = kc.clone()
    bt = torch.arange(B * 3, dtype=torch.int32).view(B, 3).to(DEV)
    sl = torch.tensor([5, 40, 17], dtype=torch.int32, device=DEV)
    key = torch.randn(B, Hkv, D, generator=g).to(torch.bfloat16).to(DEV)
    slots = (bt[torch.arange(B, device=DEV), ((sl - 1) // BS).long()].long() * BS
             + ((sl - 1) % BS).long())
    one = torch.tensor(1.0, device=DEV)
    torch.library.opcheck(
        torch.ops.b200pa.paged_attention_flash,
        (torch.empty_like(q), None, q, key, key.clone(), kc, vc, slots, 0.1, bt, sl, 40, None,
         "fp8", one, one), test_utils=("test_schema", "test_faketensor"))


class _Layer:
    def __init__(self, ks, vs):
        self._k_scale = torch.tensor(ks, device=DEV)
        self._v_scale = torch.tensor(vs, device=DEV)
        self._q_scale = torch.tensor(1.0, device=DEV)


@pytest.mark.parametrize("fused", [True, False])
@pytest.mark.parametrize("kv", ["auto", "fp8"])
def test_flash_layout_backend_decode_step_matches_oracle(ora, co, kv, fused):
    """AttentionImpl.forward of the flash-layout backend: write the new token, attend -- against
    the oracle doing the same two steps on the CPU (x-packed restatement of the same values)."""
    from vllm_main_modify_b200.backend import (B200FlashLayoutAttentionBackend,
                                               B200PagedAttentionMetadata)
    B, H, Hkv, D, NB = 6, 32, 8, 128, 400
    lens = [1, 16, 17, 900, 512, 513]            # seq_lens INCLUDING the new token
    dt = torch.bfloat16
    s = D ** -0.5
    cdt = dt if kv == "auto" else torch.uint8
    q, kc, vc, bt, sl = hp.make_paged_inputs(B, H, Hkv, D, BS, NB, lens, dt, seed=21, kv=kv,
                                             lo=-s, hi=s, permute=True)
    g = torch.Generator().manual_seed(22)
    key = ((torch.rand(B, Hkv, D, generator=g) * 2 - 1) * s).to(dt)
    value = ((torch.rand(B, Hkv, D, generator=g) * 2 - 1) * s).to(dt)
    slots = torch.tensor([int(bt[b, (L - 1) // BS]) * BS + (L - 1) % BS
                          for b, L in enumerate(lens)], dtype=torch.int64)
    ks, vs = (1.0, 1.0) if kv == "auto" else (s / 100, s / 90)
    kc_o, vc_o = kc.clone(), vc.clone()
    ora.reshape_and_cache(key, value, kc_o, vc_o, slots, kv, torch.tensor(ks), torch.tensor(vs))
    want = torch.empty_like(q)
    ora.paged_attention_v1(want, q, kc_o, vc_o, Hkv, s, bt, sl, BS, max(lens), None, kv,
                           torch.tensor(ks), torch.tensor(vs))
    backend = B200FlashLayoutAttentionBackend
    shape = backend.get_kv_cache_shape(NB, BS, Hkv, D)
    assert shape == (2, NB, BS, Hkv, D)
    kv_cache = torch.empty(shape, dtype=cdt, device=DEV)
    kf, vf = co.xpacked_to_flash(kc, vc)
    kv_cache[0] = kf.to(DEV)
    kv_cache[1] = vf.to(DEV)
    impl = backend.get_impl_cls()(H, D, s, Hkv, None, None, kv)
    impl.fused_decode_step = fused
    meta = B200PagedAttentionMetadata(num_actual_tokens=B, max_query_len=1, seq_lens=sl.to(DEV),
                                      max_seq_len=max(lens), block_table=bt.to(DEV),
                                      slot_mapping=slots.to(DEV))
    out = impl.forward(_Layer(ks, vs), q.to(DEV).view(B, H * D), key.to(DEV).view(B, Hkv * D),
                       value.to(DEV).view(B, Hkv * D), kv_cache, meta)
    torch.cuda.synchronize()
    torch.testing.assert_close(out.view(B, H, D).cpu().float(), want.float(), atol=ATOL, rtol=RTOL)
    kf_o, vf_o = co.xpacked_to_flash(kc_o, vc_o)
    assert torch.equal(kv_cache[0].cpu().view(torch.uint8), kf_o.view(torch.uint8))
    assert torch.equal(kv_cache[1].cpu().view(torch.uint8), vf_o.view(torch.uint8))

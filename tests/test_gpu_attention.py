"""GPU parity of paged_attention_v1 / v2 against the CPU oracle.

Mirrors the reference's tests/kernels/attention/test_attention.py (:121-345):
same parameter sweep (version, heads, head size, block size, ALiBi, kv dtype),
same input distributions (U(-1/sqrt(D), 1/sqrt(D)), block tables with repeats),
but compares with the oracle (oracle/paged_attn_oracle.c, pinned in
tests/test_oracle.py) within BASELINE.json's gate: 1e-2 abs + 1e-2 rel, and with
an fp64 restatement at the reference test's own tolerance (atol 1e-3 for 16-bit
KV, 1e-2 for fp8).  All calls go through the torch shim -> C ABI.
"""
import functools
import random

import pytest
import torch

from tests import helpers as hp

pytestmark = pytest.mark.gpu

DEV = "cuda:0"
ATOL, RTOL = 1e-2, 1e-2           # BASELINE.json north_star gate vs the reference kernels
GENERIC, STREAM = 1, 2         # b200pa_set_impl: SIMT kernel | persistent stream-K kernel


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


def _run_gpu(ops, version, q, kc, vc, Hkv, scale, bt, sl, BS, max_len, alibi, kv, ks, vs,
             extra=(0, 0, 0, 64, 0), parts=None):
    """Upload, run, return (out, exp_sums, max_logits, tmp_out) on the CPU."""
    qd, kcd, vcd = q.to(DEV), kc.to(DEV), vc.to(DEV)
    btd, sld = bt.to(DEV), sl.to(DEV)
    ald = None if alibi is None else alibi.to(DEV)
    ksd = torch.as_tensor(ks, dtype=torch.float32).reshape(()).to(DEV)
    vsd = torch.as_tensor(vs, dtype=torch.float32).reshape(()).to(DEV)
    out = torch.full_like(qd, float("nan"))
    B, H, D = q.shape
    if version == "v1":
        ops.paged_attention_v1(out, qd, kcd, vcd, Hkv, scale, btd, sld, BS, max_len, ald, kv,
                               ksd, vsd, *extra)
        torch.cuda.synchronize()
        return out.cpu(), None, None, None
    P = parts if parts is not None else (max_len + 511) // 512
    es = torch.full((B, H, P), float("nan"), dtype=torch.float32, device=DEV)
    ml = torch.full((B, H, P), float("nan"), dtype=torch.float32, device=DEV)
    tmp = torch.full((B, H, P, D), float("nan"), dtype=q.dtype, device=DEV)
    ops.paged_attention_v2(out, es, ml, tmp, qd, kcd, vcd, Hkv, scale, btd, sld, BS, max_len,
                           ald, kv, ksd, vsd, *extra)
    torch.cuda.synchronize()
    return out.cpu(), es.cpu(), ml.cpu(), tmp.cpu()


def _run_oracle(ora, version, q, kc, vc, Hkv, scale, bt, sl, BS, max_len, alibi, kv, ks, vs,
                extra=(0, 0, 0, 64, 0)):
    out = torch.empty_like(q)
    ks_t, vs_t = torch.tensor(float(ks)), torch.tensor(float(vs))
    if version == "v1":
        ora.paged_attention_v1(out, q, kc, vc, Hkv, scale, bt, sl, BS, max_len, alibi, kv,
                               ks_t, vs_t, *extra)
    else:
        B, H, D = q.shape
        P = (max_len + 511) // 512
        es, ml = torch.zeros(B, H, P), torch.zeros(B, H, P)
        tmp = torch.zeros(B, H, P, D, dtype=q.dtype)
        ora.paged_attention_v2(out, es, ml, tmp, q, kc, vc, Hkv, scale, bt, sl, BS, max_len,
                               alibi, kv, ks_t, vs_t, *extra)
    return out


_ORACLE_CACHE = {}
_INPUT_CACHE = {}


def _run_oracle_cached(key, ora, *args, **kw):
    """The oracle result depends on the inputs only, not on which GPU kernel family is under
    test: computed once per input set (the sweep below calls it per kernel family)."""
    if key not in _ORACLE_CACHE:
        _ORACLE_CACHE[key] = _run_oracle(ora, *args, **kw)
    return _ORACLE_CACHE[key]


def _close(got, want, atol=ATOL, rtol=RTOL):
    assert not torch.isnan(got.float()).any(), "NaN in the GPU output"
    torch.testing.assert_close(got.float(), want.float(), atol=atol, rtol=rtol)


def _alibi(H):
    g = torch.Generator().manual_seed(11)
    return torch.randn(H, generator=g, dtype=torch.float32)


# ------------------------------------------------------------ reference sweep

@pytest.mark.parametrize("version", ["v1", "v2"])
@pytest.mark.parametrize("impl", [GENERIC, STREAM])
@pytest.mark.parametrize("use_alibi", [False, True])
@pytest.mark.parametrize("heads", [(40, 40), (64, 8)])
@pytest.mark.parametrize("head_size", [32, 64, 80, 96, 112, 128, 192, 256])
@pytest.mark.parametrize("block_size", [16, 32])
@pytest.mark.parametrize("kv", ["auto", "fp8"])
def test_paged_attention_sweep(ops, native, ora, version, impl, heads, head_size, block_size,
                               use_alibi, kv):
    H, Hkv = heads
    dtype = torch.bfloat16
    stream_ok = block_size in (16, 32) and head_size in (32, 64, 80, 96, 112, 128, 192, 256)
    if impl == STREAM and not stream_ok:
        pytest.skip("shape served by the generic kernel")
    if kv == "fp8" and head_size % 16:
        pytest.skip("x = 16 does not divide the head size")
    # the reference sweeps NUM_BLOCKS = 4321 on the device; the tables here are random draws
    # (pages may alias), so a smaller pool exercises the same indexing and generates in < 1 s
    B, NB, MAX = 7, 601, 2048
    s = head_size ** -0.5
    ikey = ("sweep_in", heads, head_size, block_size, kv)
    if ikey not in _INPUT_CACHE:
        _INPUT_CACHE.clear()   # one input set alive at a time (version/impl/alibi vary fastest)
        rnd = random.Random(0)
        lens = [rnd.randint(1, MAX) for _ in range(B)]
        lens[-1] = MAX
        _INPUT_CACHE[ikey] = hp.make_paged_inputs(B, H, Hkv, head_size, block_size, NB, lens,
                                                  dtype, seed=0, kv=kv, lo=-s, hi=s)
    q, kc, vc, bt, sl = _INPUT_CACHE[ikey]
    alibi = _alibi(H) if use_alibi else None
    ks, vs = (1.0, 1.0) if kv == "auto" else (s / 2.0, s / 3.0)
    native.set_impl(impl)
    try:
        got, *_ = _run_gpu(ops, version, q, kc, vc, Hkv, s, bt, sl, block_size, MAX, alibi, kv,
                           ks, vs)
    finally:
        native.set_impl(native.IMPL_AUTO)
    want = _run_oracle_cached(("sweep", version, heads, head_size, block_size, use_alibi, kv),
                              ora, version, q, kc, vc, Hkv, s, bt, sl, block_size, MAX, alibi, kv,
                              ks, vs)
    _close(got, want)
    if head_size == 128:
        # the reference test's own check: fp64 math on the (dequantised) cache
        rkey = ("sweep_ref64", heads, head_size, block_size, use_alibi, kv)
        if rkey not in _ORACLE_CACHE:
            kd = vd = None
            if kv == "fp8":
                kd, vd = hp.e4m3_to_float(kc) * ks, hp.e4m3_to_float(vc) * vs
            _ORACLE_CACHE[rkey] = hp.torch_reference_attention(q, kc, vc, bt, sl, s, Hkv, alibi,
                                                               kd, vd)
        atol = 1e-3 if kv == "auto" else 1e-2
        torch.testing.assert_close(got.double(), _ORACLE_CACHE[rkey], atol=atol, rtol=1e-2)


@pytest.mark.parametrize("dtype", [torch.float16, torch.float32])
@pytest.mark.parametrize("version", ["v1", "v2"])
@pytest.mark.parametrize("kv", ["auto", "fp8", "fp8_e5m2"])
@pytest.mark.parametrize("block_size", [8, 16, 32])
def test_other_dtypes_and_block_sizes(ops, native, ora, dtype, version, kv, block_size):
    B, H, Hkv, D, NB = 5, 16, 4, 128, 300
    lens = [1, 17, 512, 513, 1500]
    s = D ** -0.5
    q, kc, vc, bt, sl = hp.make_paged_inputs(B, H, Hkv, D, block_size, NB, lens, dtype, seed=3,
                                             kv="auto" if kv == "auto" else "fp8", lo=-s, hi=s)
    if kv == "fp8_e5m2":
        # e5m2 codes: keep exponents moderate and avoid inf/nan (exp field 31)
        kc, vc = kc & 0xBB, vc & 0xBB
    ks, vs = (1.0, 1.0) if kv == "auto" else (0.01, 0.02)
    got, *_ = _run_gpu(ops, version, q, kc, vc, Hkv, s, bt, sl, block_size, 1500, None, kv, ks, vs)
    want = _run_oracle(ora, version, q, kc, vc, Hkv, s, bt, sl, block_size, 1500, None, kv, ks, vs)
    tol = (1e-4, 1e-3) if (dtype == torch.float32 and kv == "auto") else (ATOL, RTOL)
    _close(got, want, *tol)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("head_size", [32, 64, 80, 96, 112, 128, 192, 256])
@pytest.mark.parametrize("version", ["v1", "v2"])
@pytest.mark.parametrize("block_size", [16, 32])
def test_stream_kernel_serves_fp8_e5m2(ops, native, ora, dtype, head_size, version, block_size):
    """fp8_e5m2 caches (DISPATCH_BY_KV_CACHE_DTYPE, quant_utils.cuh:545-582) on the stream kernel:
    same tolerance against the oracle as e4m3, and the generic kernel's output beside it."""
    B, H, Hkv, D, NB = 6, 32, 4, head_size, 700
    lens = [1, 17, 512, 513, 3000, 4097]
    s = D ** -0.5
    q, kc, vc, bt, sl = hp.make_paged_inputs(B, H, Hkv, D, block_size, NB, lens, dtype, seed=5,
                                             kv="fp8", lo=-s, hi=s, permute=True)
    kc, vc = kc & 0xBB, vc & 0xBB          # moderate exponents, no inf / nan codes
    ks, vs = 0.01, 0.02
    want = _run_oracle(ora, version, q, kc, vc, Hkv, s, bt, sl, block_size, max(lens), None,
                       "fp8_e5m2", ks, vs)
    outs = {}
    for impl in (STREAM, GENERIC):
        native.set_impl(impl)
        try:
            outs[impl], *_ = _run_gpu(ops, version, q, kc, vc, Hkv, s, bt, sl, block_size,
                                      max(lens), None, "fp8_e5m2", ks, vs)
        finally:
            native.set_impl(native.IMPL_AUTO)
        _close(outs[impl], want)
    _close(outs[STREAM], outs[GENERIC])


# ------------------------------------------------------------ golden fixtures

ATTN_GOLDEN = ["attn_gqa_bf16_d128.npz", "attn_mha_bf16_d64_alibi.npz",
               "attn_gqa_f16_d80.npz", "attn_gqa_bf16_d128_bs32.npz",
               "attn_mha_f32_d32.npz"]


@pytest.mark.parametrize("name", ATTN_GOLDEN)
@pytest.mark.parametrize("version", ["v1", "v2"])
@pytest.mark.parametrize("impl", [0, GENERIC])
def test_against_reference_golden_outputs(ops, native, name, version, impl):
    """Outputs recorded from the reference's own test oracle and CPU kernels
    (tests/golden/make_golden.py)."""
    g = hp.load_golden(name)
    dt = hp.golden_dtype(g)
    q = hp.from_bits(g["q"], dt)
    kc, vc = hp.from_bits(g["key_cache"], dt), hp.from_bits(g["value_cache"], dt)
    bt = torch.from_numpy(g["block_tables"].copy())
    sl = torch.from_numpy(g["seq_lens"].copy())
    alibi = torch.from_numpy(g["alibi_slopes"].copy()) if "alibi_slopes" in g else None
    Hkv, BS, scale = int(g["num_kv_heads"]), int(g["block_size"]), float(g["scale"])
    native.set_impl(impl)
    try:
        got, *_ = _run_gpu(ops, version, q, kc, vc, Hkv, scale, bt, sl, BS, int(sl.max()), alibi,
                           "auto", 1.0, 1.0)
    finally:
        native.set_impl(native.IMPL_AUTO)
    tol = (1e-4, 1e-4) if dt == torch.float32 else (ATOL, RTOL)
    _close(got, hp.from_bits(g["out_test_ref"], dt), *tol)
    for key in ("out_cpu_v1", "out_cpu_v2"):
        if key in g:
            _close(got, hp.from_bits(g[key], dt), *tol)


# ------------------------------------------------- BASELINE configs (reduced)

@functools.lru_cache(maxsize=None)
def _config_inputs(name):
    """The five BASELINE.json configs at a batch the oracle finishes in seconds
    (same per-sequence shape)."""
    D, BS = 128, 16
    s = D ** -0.5
    if name == "cfg1":    # Llama-2-7B MHA, ctx 1024, bf16
        B, H, Hkv, lens, kv = 8, 32, 32, [1024] * 8, "auto"
    elif name == "cfg2":  # Llama-3-8B GQA 32/8, ctx 4096, bf16
        B, H, Hkv, lens, kv = 4, 32, 8, [4096] * 4, "auto"
    elif name == "cfg3":  # Llama-3-8B, fp8, ctx 8192
        B, H, Hkv, lens, kv = 4, 32, 8, [8192] * 4, "fp8"
    elif name == "cfg4":  # Llama-3-70B 64/8 sharded 8 ways -> 8 q heads, 1 kv head; ctx 32768
        B, H, Hkv, lens, kv = 3, 8, 1, [32768] * 3, "fp8"
    elif name == "cfg4_tp1":
        B, H, Hkv, lens, kv = 1, 64, 8, [32768], "fp8"
    else:                 # cfg5: ragged 128..131072, permuted tables, Qwen2-72B shard
        B, H, Hkv, kv = 6, 8, 1, "fp8"
        lens = [128, 131072, 5000, 777, 65536 + 3, 20001]
    NB = sum((L + BS - 1) // BS for L in lens) + 7
    q, kc, vc, bt, sl = hp.make_paged_inputs(B, H, Hkv, D, BS, NB, lens, torch.bfloat16, seed=1,
                                             kv=kv, lo=-s, hi=s, permute=True)
    ks, vs = (1.0, 1.0) if kv == "auto" else (s / 1.7, s / 2.9)
    return q, kc, vc, bt, sl, H, Hkv, D, BS, s, kv, ks, vs, max(lens)


@pytest.mark.parametrize("name", ["cfg1", "cfg2", "cfg3", "cfg4", "cfg4_tp1", "cfg5"])
@pytest.mark.parametrize("version", ["v1", "v2"])
@pytest.mark.parametrize("impl", [STREAM, GENERIC])
def test_baseline_configs_reduced_batch(ops, native, ora, name, version, impl):
    q, kc, vc, bt, sl, H, Hkv, D, BS, s, kv, ks, vs, max_len = _config_inputs(name)
    native.set_impl(impl)
    try:
        got, *_ = _run_gpu(ops, version, q, kc, vc, Hkv, s, bt, sl, BS, max_len, None, kv, ks, vs)
    finally:
        native.set_impl(native.IMPL_AUTO)
    want = _run_oracle_cached(("cfg", name, version), ora, version, q, kc, vc, Hkv, s, bt, sl, BS,
                              max_len, None, kv, ks, vs)
    _close(got, want)


# ---------------------------------------------------- full-size property tests

def _full_size(ops, B, H, Hkv, ctx, kv, seed=0):
    """Inputs generated ON the device at BASELINE's full size."""
    D, BS = 128, 16
    s = D ** -0.5
    g = torch.Generator(device=DEV).manual_seed(seed)
    nblk = ctx // BS
    NB = B * nblk
    q = ((torch.rand(B, H, D, generator=g, device=DEV) * 2 - 1) * s).to(torch.bfloat16)
    if kv == "auto":
        kc = ((torch.rand(NB, Hkv, D // 8, BS, 8, generator=g, device=DEV) * 2 - 1) * s).to(torch.bfloat16)
        vc = ((torch.rand(NB, Hkv, D, BS, generator=g, device=DEV) * 2 - 1) * s).to(torch.bfloat16)
    else:
        kc = torch.randint(0, 0x48, (NB, Hkv, D // 16, BS, 16), generator=g, device=DEV, dtype=torch.uint8)
        kc |= (torch.randint(0, 2, kc.shape, generator=g, device=DEV, dtype=torch.uint8) << 7)
        vc = torch.randint(0, 0x48, (NB, Hkv, D, BS), generator=g, device=DEV, dtype=torch.uint8)
        vc |= (torch.randint(0, 2, vc.shape, generator=g, device=DEV, dtype=torch.uint8) << 7)
    perm = torch.randperm(NB, generator=g, device=DEV).to(torch.int32)
    bt = perm.view(B, nblk).contiguous()
    sl = torch.full((B,), ctx, dtype=torch.int32, device=DEV)
    return q, kc, vc, bt, sl, s


def _call(ops, version, q, kc, vc, Hkv, s, bt, sl, max_len, kv, ks, vs):
    out = torch.empty_like(q)
    B, H, D = q.shape
    if version == "v1":
        ops.paged_attention_v1(out, q, kc, vc, Hkv, s, bt, sl, 16, max_len, None, kv, ks, vs)
    else:
        P = (max_len + 511) // 512
        es = torch.empty(B, H, P, dtype=torch.float32, device=DEV)
        ml = torch.empty_like(es)
        tmp = torch.empty(B, H, P, D, dtype=q.dtype, device=DEV)
        ops.paged_attention_v2(out, es, ml, tmp, q, kc, vc, Hkv, s, bt, sl, 16, max_len, None,
                               kv, ks, vs)
    return out


@pytest.mark.parametrize("cfg", [("cfg2", 64, 32, 8, 4096, "auto"),
                                 ("cfg3", 256, 32, 8, 8192, "fp8"),
                                 ("cfg4_tp8", 128, 8, 1, 32768, "fp8")])
def test_full_size_properties(ops, native, ora, cfg):
    name, B, H, Hkv, ctx, kv = cfg
    q, kc, vc, bt, sl, s = _full_size(ops, B, H, Hkv, ctx, kv)
    ks = torch.tensor(1.0 if kv == "auto" else s / 1.7, device=DEV)
    vs = torch.tensor(1.0 if kv == "auto" else s / 2.9, device=DEV)
    native.set_impl(STREAM)
    try:
        o1 = _call(ops, "v1", q, kc, vc, Hkv, s, bt, sl, ctx, kv, ks, vs)
        o2 = _call(ops, "v2", q, kc, vc, Hkv, s, bt, sl, ctx, kv, ks, vs)
        # (i) determinism
        o1b = _call(ops, "v1", q, kc, vc, Hkv, s, bt, sl, ctx, kv, ks, vs)
        assert torch.equal(o1, o1b)
        # (ii) v1 and v2 agree
        _close(o2.cpu(), o1.cpu())
        # (iii) block-table permutation invariance: move the pages, permute the table
        #       consistently -> identical results (bitwise: the page walk order is unchanged)
        g = torch.Generator(device=DEV).manual_seed(99)
        NB = kc.shape[0]
        sigma = torch.randperm(NB, generator=g, device=DEV)
        kc2 = torch.empty_like(kc)
        vc2 = torch.empty_like(vc)
        kc2[sigma] = kc
        vc2[sigma] = vc
        bt2 = sigma.to(torch.int32)[bt.long()]
        o3 = _call(ops, "v2", q, kc2, vc2, Hkv, s, bt2, sl, ctx, kv, ks, vs)
        assert torch.equal(o3, o2)
        del kc2, vc2
        # (iv) linearity in v_scale (fp8): doubling v_scale doubles the output exactly
        if kv != "auto":
            o4 = _call(ops, "v2", q, kc, vc, Hkv, s, bt, sl, ctx, kv, ks, vs * 2)
            assert torch.equal(o4.float(), o2.float() * 2)
    finally:
        native.set_impl(native.IMPL_AUTO)
    # (v) the generic kernel (different algorithm) agrees on the whole batch
    native.set_impl(GENERIC)
    try:
        og = _call(ops, "v2", q, kc, vc, Hkv, s, bt, sl, ctx, kv, ks, vs)
    finally:
        native.set_impl(native.IMPL_AUTO)
    _close(o2.cpu(), og.cpu())
    # (vi) the oracle on a sample of sequences (first, middle, last)
    idx = [0, B // 2, B - 1]
    sel = torch.tensor(idx, device=DEV)
    qs, bts, sls = q[sel].cpu(), bt[sel].cpu(), sl[sel].cpu()
    # compact the touched pages so the CPU copy stays small
    used = torch.unique(bts.flatten().long())
    remap = torch.zeros(kc.shape[0], dtype=torch.int32)
    remap[used] = torch.arange(used.numel(), dtype=torch.int32)
    kcs, vcs = kc[used.to(DEV)].cpu(), vc[used.to(DEV)].cpu()
    bts = remap[bts.long()]
    want = _run_oracle(ora, "v2", qs, kcs, vcs, Hkv, s, bts, sls, 16, ctx, None, kv,
                       float(ks), float(vs))
    _close(o2[sel].cpu(), want)
    _close(o1[sel].cpu(), want)


# ------------------------------------------------- stream-K scheduler cases

_RND = random.Random(0)
SCHED_CASES = [
    # name, B, H, Hkv, D, lens, dtype, kv, alibi, qmul
    ("gqa8_fp8_ragged", 6, 8, 1, 128, [128, 40000, 5000, 777, 16384 + 3, 20001], torch.bfloat16, "fp8", False, 1.0),
    ("gqa8_fp8_hot", 5, 64, 8, 128, [1000, 17, 512, 640, 3333], torch.bfloat16, "fp8", False, 40.0),
    ("gqa4_bf16_hot", 5, 32, 8, 128, [1000, 17, 512, 640, 3333], torch.bfloat16, "auto", False, 40.0),
    ("mha_f16_d64_alibi", 7, 40, 40, 64, [_RND.randint(1, 3000) for _ in range(7)], torch.float16, "auto", True, 10.0),
    ("gqa16_fp8_d64_alibi", 7, 32, 2, 64, [_RND.randint(1, 3000) for _ in range(7)], torch.float16, "fp8", True, 10.0),
    ("gqa5_bf16", 3, 10, 2, 128, [1, 513, 1025], torch.bfloat16, "auto", False, 5.0),
    ("many_short_fp8", 300, 16, 2, 128, [_RND.randint(1, 700) for _ in range(300)], torch.bfloat16, "fp8", False, 20.0),
    ("with_empties_fp8", 9, 8, 1, 128, [0, 600, 0, 0, 31, 5000, 0, 512, 0], torch.bfloat16, "fp8", False, 20.0),
    ("one_long_bf16", 1, 8, 1, 128, [70001], torch.bfloat16, "auto", False, 20.0),
    ("one_long_fp8", 1, 64, 8, 128, [131072], torch.bfloat16, "fp8", False, 20.0),
    ("tiny", 2, 8, 8, 128, [1, 2], torch.bfloat16, "auto", False, 1.0),
    ("gemma_d256_fp8", 5, 16, 8, 256, [1000, 17, 512, 4100, 3333], torch.bfloat16, "fp8", False, 10.0),
    ("gemma_d256_bf16", 5, 16, 8, 256, [1000, 17, 512, 4100, 3333], torch.bfloat16, "auto", False, 10.0),
    ("d192_f16_fp8_alibi", 4, 24, 3, 192, [700, 33, 2049, 16], torch.float16, "fp8", True, 5.0),
    ("phi_d96_bf16", 6, 32, 32, 96, [1, 16, 17, 999, 2048, 31], torch.bfloat16, "auto", False, 5.0),
    ("d96_fp8_gqa12", 4, 24, 2, 96, [300, 5000, 12, 640], torch.bfloat16, "fp8", False, 5.0),
    ("d32_bf16", 4, 8, 2, 32, [300, 5000, 12, 640], torch.bfloat16, "auto", False, 5.0),
    ("d32_fp8", 4, 8, 8, 32, [300, 1500, 12, 640], torch.float16, "fp8", False, 5.0),
]


@pytest.mark.parametrize("case", SCHED_CASES, ids=[c[0] for c in SCHED_CASES])
def test_stream_scheduler_cases(ops, native, ora, case):
    """Shapes that stress the stream-K decomposition (spans that cut sequences anywhere, many
    sequences per span, one sequence over a hundred spans, empty sequences) and the lazy
    softmax rescale ("hot" logits: q scaled up so that the running max keeps moving).
    k_scale is a power of two so that the reference's rounding of K * k_scale to scalar_t
    (quant_utils.cuh:362-372) is exact and peaky softmaxes can be compared tightly."""
    name, B, H, Hkv, D, lens, dtype, kv, use_alibi, qmul = case
    BS = 16
    s = D ** -0.5
    NB = sum((L + BS - 1) // BS for L in lens) + 5
    q, kc, vc, bt, sl = hp.make_paged_inputs(B, H, Hkv, D, BS, NB, lens, dtype, seed=3, kv=kv,
                                             lo=-1.0, hi=1.0, permute=True)
    q = (q.float() * qmul).to(dtype)
    ks, vs = (1.0, 1.0) if kv == "auto" else (0.5, 1.3)
    alibi = None
    if use_alibi:
        alibi = torch.tensor([2.0 ** (-(i + 1) * 8.0 / H) for i in range(H)], dtype=torch.float32)
    max_len = max(max(lens), 1)
    live = torch.tensor([L > 0 for L in lens])
    outs = {}
    for impl, version in ((GENERIC, "v2"), (STREAM, "v1"), (STREAM, "v2")):
        native.set_impl(impl)
        try:
            outs[(impl, version)], *_ = _run_gpu(ops, version, q, kc, vc, Hkv, s, bt, sl, BS,
                                                 max_len, alibi, kv, ks, vs)
        finally:
            native.set_impl(native.IMPL_AUTO)
    ref = outs[(GENERIC, "v2")][live]
    _close(outs[(STREAM, "v1")][live], ref)
    _close(outs[(STREAM, "v2")][live], ref)
    # v1 and v2 are the same kernel and schedule: bitwise equal
    assert torch.equal(outs[(STREAM, "v1")][live], outs[(STREAM, "v2")][live])
    # empty sequences: zeros, as the reference's v1 kernel leaves them
    if not bool(live.all()):
        assert float(outs[(STREAM, "v1")][~live].float().abs().sum()) == 0.0
    if sum(lens) * H <= 64 * 140000:
        want = _run_oracle(ora, "v1", q, kc, vc, Hkv, s, bt, sl, BS, max_len, alibi, kv, ks, vs)
        # peaky softmaxes over V of magnitude up to 2.6: the reference's own rounding of P to
        # scalar_t (attention_kernels.cuh:393-395) is ~1e-2 of max|V| here, so the absolute gate
        # scales with the output range (it is 1e-2 for the unit-range cases)
        atol = ATOL * max(1.0, float(want[live].float().abs().max()))
        _close(outs[(STREAM, "v2")][live], want[live], atol=atol)


def test_stream_graph_replay_follows_seq_lens(ops, native):
    """The stream kernel derives its schedule from seq_lens ON THE DEVICE: a captured graph
    replayed after the lengths (and block tables) changed gives the result of a fresh call."""
    B, H, Hkv, D, BS = 12, 16, 2, 128, 16
    s = D ** -0.5
    lens_a = [3000, 17, 900, 4096, 1, 2222, 640, 33, 1500, 2048, 77, 4000]
    lens_b = [5, 4096, 4096, 12, 3000, 100, 4095, 2049, 1, 700, 3333, 16]
    NB = 12 * 256 + 3
    q, kc, vc, bt, sl = hp.make_paged_inputs(B, H, Hkv, D, BS, NB, [4096] * B, torch.bfloat16,
                                             seed=5, kv="fp8", lo=-1.0, hi=1.0, permute=True)
    qd, kcd, vcd, btd = q.to(DEV), kc.to(DEV), vc.to(DEV), bt.to(DEV)
    sld = torch.tensor(lens_a, dtype=torch.int32, device=DEV)
    ks, vs = torch.tensor(0.5, device=DEV), torch.tensor(1.5, device=DEV)
    P = 8
    es = torch.empty(B, H, P, dtype=torch.float32, device=DEV)
    ml = torch.empty_like(es)
    tmp = torch.empty(B, H, P, D, dtype=qd.dtype, device=DEV)
    out_g = torch.empty_like(qd)

    def step(out):
        ops.paged_attention_v2(out, es, ml, tmp, qd, kcd, vcd, Hkv, s, btd, sld, BS, 4096, None,
                               "fp8", ks, vs)

    stream = torch.cuda.Stream()
    with torch.cuda.stream(stream):
        step(out_g)
        stream.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=stream):
            step(out_g)
    for lens in (lens_a, lens_b, lens_a):
        sld.copy_(torch.tensor(lens, dtype=torch.int32))
        out_g.fill_(float("nan"))
        graph.replay()
        torch.cuda.synchronize()
        out_e = torch.empty_like(qd)
        step(out_e)
        torch.cuda.synchronize()
        assert torch.equal(out_g, out_e)
        assert not torch.isnan(out_g.float()).any()


def test_first_use_of_a_stream_inside_a_capture_leaves_a_usable_workspace(ops):
    """The shim's per-stream workspace is allocated and its ticket counters zeroed for real even
    when the first call on a stream is made under CUDA-graph capture: the graph captured first is
    NEVER replayed here, yet an eager call and a second graph on the same stream are correct
    (sequences cut across CTAs go through the counters)."""
    B, H, Hkv, D, BS = 6, 16, 2, 128, 16
    s = D ** -0.5
    lens = [4096, 9, 3000, 4095, 700, 2049]
    q, kc, vc, bt, sl = hp.make_paged_inputs(B, H, Hkv, D, BS, 6 * 256 + 1, lens, torch.bfloat16,
                                             seed=8, kv="fp8", lo=-1.0, hi=1.0, permute=True)
    qd, kcd, vcd, btd, sld = (t.to(DEV) for t in (q, kc, vc, bt, sl))
    ks, vs = torch.tensor(0.5, device=DEV), torch.tensor(1.5, device=DEV)

    def step(out):
        ops.paged_attention_v1(out, qd, kcd, vcd, Hkv, s, btd, sld, BS, 4096, None, "fp8", ks, vs)

    want = torch.empty_like(qd)
    step(want)                                    # default stream: its own workspace
    torch.cuda.synchronize()
    stream = torch.cuda.Stream()                  # never used before
    out1, out2, out3 = (torch.full_like(qd, float("nan")) for _ in range(3))
    g1, g2 = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
    with torch.cuda.stream(stream):
        with torch.cuda.graph(g1, stream=stream):
            step(out1)                            # first use of `stream`: under capture
        step(out2)                                # eager, g1 not replayed
        stream.synchronize()
        with torch.cuda.graph(g2, stream=stream):
            step(out3)
    g2.replay()
    torch.cuda.synchronize()
    assert torch.equal(out2, want)
    assert torch.equal(out3, want)
    g1.replay()
    torch.cuda.synchronize()
    assert torch.equal(out1, want)


def test_c_abi_without_workspace_uses_generic_and_stream_needs_one(native):
    """Direct C-ABI calls (ctypes, raw device pointers): workspace == NULL is served by the
    generic kernel; forcing the stream kernel without a workspace is an error, not a fallback."""
    import ctypes
    lib = native.c_lib()
    B, H, Hkv, D, BS, NB = 3, 8, 2, 128, 16, 40
    lens = [100, 37, 300]
    s = D ** -0.5
    q, kc, vc, bt, sl = hp.make_paged_inputs(B, H, Hkv, D, BS, NB, lens, torch.bfloat16, seed=2,
                                             lo=-s, hi=s, permute=True)
    qd, kcd, vcd, btd, sld = (t.to(DEV) for t in (q, kc, vc, bt, sl))
    one = torch.tensor(1.0, device=DEV)
    lib.b200pa_paged_attention_workspace_bytes.restype = ctypes.c_int64
    need = lib.b200pa_paged_attention_workspace_bytes(D)
    assert 0 < need < (64 << 20)
    # ABI v7: the workspace header (ticket counters) is zero before the first use
    ws = torch.zeros(need, dtype=torch.uint8, device=DEV)

    def call(out, wsp, wsn):
        p = lambda t: ctypes.c_void_p(t.data_ptr())
        return lib.b200pa_paged_attention_v1(
            p(out), p(qd), p(kcd), p(vcd), B, H, Hkv, D, BS, ctypes.c_float(s), p(btd),
            ctypes.c_int(bt.shape[1]), p(sld), max(lens), None, ctypes.c_int64(qd.stride(0)),
            ctypes.c_int64(kcd.stride(0)), ctypes.c_int64(kcd.stride(1)), 2, 0, p(one), p(one),
            0, 0, 0, 64, 0, wsp, ctypes.c_int64(wsn), None)

    torch.cuda.synchronize()
    out_ws = torch.empty_like(qd)
    n0 = native.launch_count()
    assert call(out_ws, ctypes.c_void_p(ws.data_ptr()), need) == 0
    assert native.launch_count() - n0 == 1          # ONE launch: the merge is inside the kernel
    torch.cuda.synchronize()
    assert int(ws[:4096].max()) == 0                # ... and it leaves the header clean
    out_nows = torch.empty_like(qd)
    n0 = native.launch_count()
    assert call(out_nows, None, 0) == 0
    assert native.launch_count() - n0 == 1          # generic kernel
    torch.cuda.synchronize()
    _close(out_ws.cpu(), out_nows.cpu())
    native.set_impl(STREAM)
    try:
        assert call(out_nows, None, 0) != 0
        assert b"workspace" in lib.b200pa_last_error()
        assert call(out_nows, ctypes.c_void_p(ws.data_ptr()), 1024) != 0
        assert b"workspace too small" in lib.b200pa_last_error()
    finally:
        native.set_impl(native.IMPL_AUTO)
    torch.cuda.synchronize()


# ------------------------------------------------------------- layout details

@pytest.mark.parametrize("version", ["v1", "v2"])
@pytest.mark.parametrize("kv", ["auto", "fp8"])
def test_strided_query_and_padded_cache(ops, ora, version, kv):
    """query is a slice of a fused qkv tensor (q_stride != H*D, paged_attention_v1.cu:172)
    and the cache has a padded block stride (kv_block_stride from stride(0))."""
    B, H, Hkv, D, BS, NB = 5, 32, 8, 128, 16, 200
    lens = [700, 16, 15, 1025, 333]
    s = D ** -0.5
    q, kc, vc, bt, sl = hp.make_paged_inputs(B, H, Hkv, D, BS, NB, lens, torch.bfloat16, seed=4,
                                             kv=kv, lo=-s, hi=s)
    ks, vs = (1.0, 1.0) if kv == "auto" else (0.02, 0.03)
    want = _run_oracle(ora, version, q, kc, vc, Hkv, s, bt, sl, BS, 1025, None, kv, ks, vs)
    qkv = torch.zeros(B, H + 2 * Hkv, D, dtype=q.dtype, device=DEV)
    qkv[:, :H] = q.to(DEV)
    qd = qkv[:, :H]
    assert qd.stride(0) != H * D
    x = kc.shape[-1]
    pad = 3
    kbig = torch.zeros(NB, Hkv + pad, D // x, BS, x, dtype=kc.dtype, device=DEV)
    vbig = torch.zeros(NB, Hkv + pad, D, BS, dtype=vc.dtype, device=DEV)
    kbig[:, :Hkv] = kc.to(DEV)
    vbig[:, :Hkv] = vc.to(DEV)
    kcd, vcd = kbig[:, :Hkv], vbig[:, :Hkv]
    out = torch.empty(B, H, D, dtype=q.dtype, device=DEV)
    ksd, vsd = torch.tensor(ks, device=DEV), torch.tensor(vs, device=DEV)
    if version == "v1":
        ops.paged_attention_v1(out, qd, kcd, vcd, Hkv, s, bt.to(DEV), sl.to(DEV), BS, 1025, None,
                               kv, ksd, vsd)
    else:
        P = 3
        es = torch.empty(B, H, P, dtype=torch.float32, device=DEV)
        ml = torch.empty_like(es)
        tmp = torch.empty(B, H, P, D, dtype=q.dtype, device=DEV)
        ops.paged_attention_v2(out, es, ml, tmp, qd, kcd, vcd, Hkv, s, bt.to(DEV), sl.to(DEV),
                               BS, 1025, None, kv, ksd, vsd)
    _close(out.cpu(), want)


@pytest.mark.parametrize("mode", ["generic", "env", "op"])
def test_v2_scratch_contract(ops, native, mode, monkeypatch):
    """After paged_attention_v2 through the generic kernel the scratch buffers hold the
    reference's partials: merging them with the reference's reduce formula
    (attention_kernels.cuh:560-664) reproduces `out`.  The default (stream) path keeps its
    partials in its own workspace and leaves the caller's scratch untouched; a caller that reads
    the scratch sets B200PA_V2_REFERENCE_SCRATCH=1 (mode "env": no test hook involved) or calls
    b200pa::paged_attention_v2_partials (mode "op": per call, no process-wide state)."""
    B, H, Hkv, D, BS, NB = 4, 16, 4, 128, 16, 600
    lens = [4096, 513, 100, 2049]
    s = D ** -0.5
    q, kc, vc, bt, sl = hp.make_paged_inputs(B, H, Hkv, D, BS, NB, lens, torch.bfloat16, seed=6,
                                             lo=-s, hi=s)
    if mode == "env":
        monkeypatch.setenv("B200PA_V2_REFERENCE_SCRATCH", "1")
    elif mode == "op":
        # the per-call form: same schema, partials guaranteed on this call
        monkeypatch.setattr(ops, "paged_attention_v2", ops.paged_attention_v2_partials)
    else:
        native.set_impl(GENERIC)
    try:
        out, es, ml, tmp = _run_gpu(ops, "v2", q, kc, vc, Hkv, s, bt, sl, BS, 4096, None, "auto",
                                    1.0, 1.0)
    finally:
        native.set_impl(native.IMPL_AUTO)
    for b in range(B):
        npart = (lens[b] + 511) // 512
        e, m, t = es[b, :, :npart], ml[b, :, :npart], tmp[b, :, :npart].float()
        assert not torch.isnan(e).any() and not torch.isnan(m).any()
        live = e > 0
        mm = torch.where(live, m, torch.full_like(m, -float("inf")))
        gmax = mm.max(dim=1, keepdim=True).values
        w = torch.where(live, e * torch.exp(mm - gmax), torch.zeros_like(e))
        t = torch.where(live.unsqueeze(-1), t, torch.zeros_like(t))
        merged = (t * w.unsqueeze(-1)).sum(1) / (w.sum(1, keepdim=True) + 1e-6)
        torch.testing.assert_close(out[b].float(), merged, atol=2e-3, rtol=2e-2)


def test_zero_length_sequence_v1_gives_zeros(ops, native):
    B, H, Hkv, D, BS, NB = 3, 8, 2, 128, 16, 20
    lens = [40, 0, 7]
    s = D ** -0.5
    q, kc, vc, bt, sl = hp.make_paged_inputs(B, H, Hkv, D, BS, NB, lens, torch.bfloat16, seed=8,
                                             lo=-s, hi=s)
    for impl in (STREAM, GENERIC):
        native.set_impl(impl)
        try:
            out, *_ = _run_gpu(ops, "v1", q, kc, vc, Hkv, s, bt, sl, BS, 40, None, "auto", 1.0, 1.0)
        finally:
            native.set_impl(native.IMPL_AUTO)
        assert float(out[1].float().abs().sum()) == 0.0
        assert float(out[0].float().abs().sum()) > 0.0


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs in one process")
@pytest.mark.parametrize("kv", ["auto", "fp8"])
def test_second_device_in_one_process(ops, native, ora, kv):
    """The stream kernel's shared-memory opt-in is a per-device function attribute: a process
    that has already launched on cuda:0 must still be able to launch on cuda:1 (the reference's
    ops run under an OptionalCUDAGuard of the query's device, paged_attention_v1.cu:88)."""
    B, H, Hkv, D, BS, NB = 3, 16, 2, 128, 16, 100
    lens = [300, 17, 1000]
    s = D ** -0.5
    q, kc, vc, bt, sl = hp.make_paged_inputs(B, H, Hkv, D, BS, NB, lens, torch.bfloat16, seed=21,
                                             kv=kv, lo=-s, hi=s, permute=True)
    want = _run_oracle(ora, "v1", q, kc, vc, Hkv, s, bt, sl, BS, max(lens), None, kv, 1.0, 1.0)
    native.set_impl(STREAM)
    try:
        for dev in ("cuda:0", "cuda:1", "cuda:0"):
            qd, kcd, vcd, btd, sld = (t.to(dev) for t in (q, kc, vc, bt, sl))
            one = torch.ones((), dtype=torch.float32, device=dev)
            out = torch.full_like(qd, float("nan"))
            ops.paged_attention_v1(out, qd, kcd, vcd, Hkv, s, btd, sld, BS, max(lens), None, kv,
                                   one, one, 0, 0, 0, 64, 0)
            torch.cuda.synchronize(dev)
            torch.testing.assert_close(out.cpu().float(), want.float(), atol=1e-2, rtol=1e-2)
    finally:
        native.set_impl(native.IMPL_AUTO)


def test_tail_page_garbage_is_ignored(ops, native, ora):
    """Slots past seq_len in the last page may hold anything, including NaN codes
    (attention_kernels.cuh:415-425 zeroes them)."""
    B, H, Hkv, D, BS, NB = 2, 8, 2, 128, 16, 12
    lens = [21, 35]
    s = D ** -0.5
    for kv in ("auto", "fp8"):
        q, kc, vc, bt, sl = hp.make_paged_inputs(B, H, Hkv, D, BS, NB, lens, torch.bfloat16,
                                                 seed=9, kv=kv, lo=-s, hi=s, permute=True)
        want = _run_oracle(ora, "v1", q, kc, vc, Hkv, s, bt, sl, BS, 35, None, kv, 0.5, 0.25)
        kc2, vc2 = kc.clone(), vc.clone()
        for b, L in enumerate(lens):
            blk, off = int(bt[b, (L - 1) // BS]), L % BS
            if kv == "auto":
                kc2[blk, :, :, off:, :] = float("nan")
                vc2[blk, :, :, off:] = float("nan")
            else:
                kc2[blk, :, :, off:, :] = 0x7F
                vc2[blk, :, :, off:] = 0x7F
        for impl in (STREAM, GENERIC):
            native.set_impl(impl)
            try:
                for version in ("v1", "v2"):
                    got, *_ = _run_gpu(ops, version, q, kc2, vc2, Hkv, s, bt, sl, BS, 35, None, kv,
                                       0.5, 0.25)
                    _close(got, want)
            finally:
                native.set_impl(native.IMPL_AUTO)


def test_blocksparse_matches_oracle(ops, ora):
    B, H, Hkv, D, BS, NB = 3, 8, 2, 64, 16, 120
    lens = [300, 513, 1200]
    s = D ** -0.5
    q, kc, vc, bt, sl = hp.make_paged_inputs(B, H, Hkv, D, BS, NB, lens, torch.bfloat16, seed=7,
                                             lo=-s, hi=s)
    extra = (1, 2, 3, 32, 1)   # tp_rank, local_blocks, vert_stride, block_size, sliding_step
    for version in ("v1", "v2"):
        got, *_ = _run_gpu(ops, version, q, kc, vc, Hkv, s, bt, sl, BS, 1200, None, "auto", 1.0,
                           1.0, extra=extra)
        want = _run_oracle(ora, version, q, kc, vc, Hkv, s, bt, sl, BS, 1200, None, "auto", 1.0,
                           1.0, extra=extra)
        _close(got, want)


@pytest.mark.parametrize("kv", ["auto", "fp8"])
@pytest.mark.parametrize("block_size", [16, 32])
@pytest.mark.parametrize("extra", [
    (0, 2, 3, 32, 1),      # tp_rank, local_blocks, vert_stride, blocksparse_block_size, step
    (1, 4, 8, 64, 1),      # the reference's default sparse block, stripes slide with the q head
    (1, 1, 4, 64, -1),     # ... with the kv head (negative step, attention_kernels.cuh:216-220)
    (0, 16, 5, 16, 2),
])
def test_blocksparse_on_the_stream_kernel(ops, native, ora, kv, block_size, extra):
    """Block-sparse attention (attention_kernels.cuh:207-251): the stream kernel streams every
    block and masks per (head, block); the reference skips -- same numbers."""
    B, H, Hkv, D, NB = 4, 16, 4, 128, 400
    lens = [300, 513, 2500, 64]
    s = D ** -0.5
    q, kc, vc, bt, sl = hp.make_paged_inputs(B, H, Hkv, D, block_size, NB, lens, torch.bfloat16,
                                             seed=7, kv=kv, lo=-s, hi=s, permute=True)
    ks, vs = (1.0, 1.0) if kv == "auto" else (s / 2, s / 3)
    want = _run_oracle(ora, "v1", q, kc, vc, Hkv, s, bt, sl, block_size, max(lens), None, kv, ks,
                       vs, extra=extra)
    for version in ("v1", "v2"):
        for impl in (STREAM, GENERIC):
            native.set_impl(impl)
            try:
                got, *_ = _run_gpu(ops, version, q, kc, vc, Hkv, s, bt, sl, block_size, max(lens),
                                   None, kv, ks, vs, extra=extra)
            finally:
                native.set_impl(native.IMPL_AUTO)
            _close(got, want)


# ----------------------------------------------------------- boundary contract

def _small_device_case(kv="fp8"):
    B, H, Hkv, D, BS, NB = 4, 32, 8, 128, 16, 300
    lens = [1000, 37, 512, 640]
    s = D ** -0.5
    q, kc, vc, bt, sl = hp.make_paged_inputs(B, H, Hkv, D, BS, NB, lens, torch.bfloat16, seed=12,
                                             kv=kv, lo=-s, hi=s)
    return [t.to(DEV) for t in (q, kc, vc, bt, sl)] + [Hkv, s, BS, 1000]


def test_opcheck_schemas(ops):
    """tests/kernels/utils.py:1247-1261 of the reference: torch.library.opcheck on the
    registered ops with the positional tail (..., 0, 0, 0, 64, 0)."""
    q, kc, vc, bt, sl, Hkv, s, BS, max_len = _small_device_case()
    one = torch.tensor(0.02, device=DEV)
    out = torch.empty_like(q)
    # ALL_OPCHECK_TEST_UTILS of the reference (tests/kernels/utils.py:33-38)
    utils = ("test_schema", "test_autograd_registration", "test_faketensor",
             "test_aot_dispatch_dynamic")
    for ns in (torch.ops.b200pa, torch.ops._C):
        torch.library.opcheck(ns.paged_attention_v1,
                              (out, q, kc, vc, Hkv, s, bt, sl, BS, max_len, None, "fp8", one, one,
                               0, 0, 0, 64, 0), test_utils=utils)
        B, H, D = q.shape
        P = 2
        es = torch.empty(B, H, P, dtype=torch.float32, device=DEV)
        ml = torch.empty_like(es)
        tmp = torch.empty(B, H, P, D, dtype=q.dtype, device=DEV)
        torch.library.opcheck(ns.paged_attention_v2,
                              (out, es, ml, tmp, q, kc, vc, Hkv, s, bt, sl, BS, max_len, None,
                               "fp8", one, one, 0, 0, 0, 64, 0), test_utils=utils)
    key = torch.randn(4, Hkv, 128, dtype=torch.bfloat16, device=DEV)
    slots = torch.tensor([3, 99, -1, 1000], dtype=torch.int64, device=DEV)
    for ns in (torch.ops.b200pa, torch.ops._C_cache_ops):
        torch.library.opcheck(ns.reshape_and_cache, (key, key, kc, vc, slots, "fp8", one, one),
                              test_utils=utils)
        kf = torch.zeros(10, BS, Hkv, 128, dtype=torch.uint8, device=DEV)
        torch.library.opcheck(ns.reshape_and_cache_flash,
                              (key, key, kf, kf.clone(), slots[:3], "fp8", one, one),
                              test_utils=utils)
    assert torch.ops._C_cuda_utils.get_max_shared_memory_per_block_device_attribute(0) >= 227 * 1024


def test_cuda_graph_capture_and_replay(ops):
    """The ops are asynchronous on the current stream, allocation-free and read
    k_scale / v_scale on the device: a decode step captures into a CUDA graph."""
    q, kc, vc, bt, sl, Hkv, s, BS, max_len = _small_device_case()
    ks, vs = torch.tensor(0.02, device=DEV), torch.tensor(0.03, device=DEV)
    B, H, D = q.shape
    key = torch.randn(B, Hkv, D, dtype=torch.bfloat16, device=DEV) * 0.05
    slots = (bt[torch.arange(B, device=DEV), ((sl - 1) // BS).long()].long() * BS
             + ((sl - 1) % BS).long())
    P = 2
    es = torch.empty(B, H, P, dtype=torch.float32, device=DEV)
    ml = torch.empty_like(es)
    tmp = torch.empty(B, H, P, D, dtype=q.dtype, device=DEV)
    out_g = torch.empty_like(q)

    def step(out):
        ops.reshape_and_cache(key, key, kc, vc, slots, "fp8", ks, vs)
        ops.paged_attention_v2(out, es, ml, tmp, q, kc, vc, Hkv, s, bt, sl, BS, max_len, None,
                               "fp8", ks, vs)

    stream = torch.cuda.Stream()
    with torch.cuda.stream(stream):
        step(out_g)                      # warm-up on the side stream
        stream.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=stream):
            step(out_g)
    out_g.zero_()
    graph.replay()
    torch.cuda.synchronize()
    out_e = torch.empty_like(q)
    step(out_e)
    torch.cuda.synchronize()
    assert torch.equal(out_g, out_e)
    # the scale is read on the device at replay time
    # (the step also re-quantises the new token's V with the new scale, so the
    # comparison is against an eager step under the same scale, not "old * 2")
    prev = out_g.clone()
    vs.mul_(2)
    graph.replay()
    torch.cuda.synchronize()
    step(out_e)
    torch.cuda.synchronize()
    assert torch.equal(out_g, out_e)
    assert not torch.equal(out_g, prev)


def test_error_convention(ops):
    q, kc, vc, bt, sl, Hkv, s, BS, max_len = _small_device_case("auto")
    one = torch.tensor(1.0, device=DEV)
    out = torch.empty_like(q)
    q48 = torch.zeros(4, 32, 48, dtype=torch.bfloat16, device=DEV)
    with pytest.raises(RuntimeError, match="Unsupported head size"):
        ops.paged_attention_v1(torch.empty_like(q48), q48, kc, vc, Hkv, s, bt, sl, BS, max_len,
                               None, "auto", one, one)
    with pytest.raises(RuntimeError, match="Unsupported block size"):
        ops.paged_attention_v1(out, q, kc, vc, Hkv, s, bt, sl, 4, max_len, None, "auto", one, one)
    with pytest.raises(RuntimeError, match="Unsupported data type of kv cache"):
        ops.paged_attention_v1(out, q, kc, vc, Hkv, s, bt, sl, BS, max_len, None, "int4", one, one)
    with pytest.raises(RuntimeError):   # CPU tensors: there is no CPU path
        ops.paged_attention_v1(out.cpu(), q.cpu(), kc.cpu(), vc.cpu(), Hkv, s, bt.cpu(), sl.cpu(),
                               BS, max_len, None, "auto", one.cpu(), one.cpu())


def test_launch_counter_counts_native_launches(ops, native):
    q, kc, vc, bt, sl, Hkv, s, BS, max_len = _small_device_case()
    one = torch.tensor(0.02, device=DEV)
    n0 = native.launch_count()
    ops.paged_attention_v1(torch.empty_like(q), q, kc, vc, Hkv, s, bt, sl, BS, max_len, None,
                           "fp8", one, one)
    assert native.launch_count() == n0 + 1   # ONE kernel: the cross-CTA merge is inside it


# ------------------------------------------------------ backend (AttentionImpl)

class _Layer:
    """AttentionLayer protocol (abstract.py:92-110): carries the scale tensors."""

    def __init__(self, ks, vs):
        self._k_scale = torch.tensor(ks, device=DEV)
        self._v_scale = torch.tensor(vs, device=DEV)
        self._q_scale = torch.tensor(1.0, device=DEV)


@pytest.mark.parametrize("kv", ["auto", "fp8"])
def test_backend_decode_step_matches_oracle(ora, kv):
    """AttentionImpl.forward: write the new token's K/V, then attend -- against the
    oracle doing the same two steps on the CPU."""
    from vllm_main_modify_b200.backend import (B200PagedAttentionBackend,
                                               B200PagedAttentionMetadata)
    B, H, Hkv, D, BS, NB = 6, 32, 8, 128, 16, 400
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
    # oracle: write, then attend
    kc_o, vc_o = kc.clone(), vc.clone()
    ora.reshape_and_cache(key, value, kc_o, vc_o, slots, kv, torch.tensor(ks), torch.tensor(vs))
    want = _run_oracle(ora, "v1", q, kc_o, vc_o, Hkv, s, bt, sl, BS, max(lens), None, kv, ks, vs)
    # product: the backend object model
    backend = B200PagedAttentionBackend
    shape = backend.get_kv_cache_shape(NB, BS, Hkv, D)
    kv_cache = torch.empty(shape, dtype=cdt, device=DEV)
    kv_cache[0] = kc.reshape(NB, -1).to(DEV)
    kv_cache[1] = vc.reshape(NB, -1).to(DEV)
    impl = backend.get_impl_cls()(H, D, s, Hkv, None, None, kv)
    meta = B200PagedAttentionMetadata(num_actual_tokens=B, max_query_len=1, seq_lens=sl.to(DEV),
                                      max_seq_len=max(lens), block_table=bt.to(DEV),
                                      slot_mapping=slots.to(DEV))
    out = impl.forward(_Layer(ks, vs), q.to(DEV).view(B, H * D), key.to(DEV).view(B, Hkv * D),
                       value.to(DEV).view(B, Hkv * D), kv_cache, meta)
    torch.cuda.synchronize()
    _close(out.view(B, H, D).cpu(), want)
    # and the cache bytes are the oracle's
    assert torch.equal(kv_cache[0].cpu().view(torch.uint8), kc_o.reshape(NB, -1).view(torch.uint8))
    assert torch.equal(kv_cache[1].cpu().view(torch.uint8), vc_o.reshape(NB, -1).view(torch.uint8))


@pytest.mark.parametrize("head_size", [128, 120])
def test_forward_decode_long_context_needs_no_scratch_on_the_one_launch_path(ops, head_size):
    """PagedAttention.forward_decode beyond the reference's v1 limit (8192 tokens): where the
    one-launch kernels serve the shape it allocates none of the reference's v2 scratch
    ([seqs, heads, partitions, head_size] + two fp32 tensors) and returns what paged_attention_v2
    returns; a shape left to the SIMT kernel (head size 120) keeps the reference's v1 / v2 choice."""
    from vllm_main_modify_b200.paged_attn import PagedAttention
    B, H, Hkv, D, BS = 2, 64, 8, head_size, 16
    lens = [9000, 300]
    s = D ** -0.5
    q, kc, vc, bt, sl = hp.make_paged_inputs(B, H, Hkv, D, BS, 600, lens, torch.bfloat16, seed=12,
                                             lo=-s, hi=s, permute=True)
    qd, kcd, vcd, btd, sld = (t.to(DEV) for t in (q, kc, vc, bt, sl))
    one = torch.tensor(1.0, device=DEV)
    P = (9000 + 511) // 512
    es = torch.empty(B, H, P, dtype=torch.float32, device=DEV)
    ml = torch.empty_like(es)
    tmp = torch.empty(B, H, P, D, dtype=qd.dtype, device=DEV)
    want = torch.empty_like(qd)
    ops.paged_attention_v2(want, es, ml, tmp, qd, kcd, vcd, Hkv, s, btd, sld, BS, 9000, None,
                           "auto", one, one)
    out = torch.empty_like(qd)
    torch.cuda.synchronize()
    torch.cuda.reset_peak_memory_stats()
    base = torch.cuda.memory_allocated()
    got = PagedAttention.forward_decode(qd, kcd, vcd, btd, sld, 9000, "auto", Hkv, s, None, one,
                                        one, output=out)
    torch.cuda.synchronize()
    grew = torch.cuda.max_memory_allocated() - base
    assert got is out
    assert torch.equal(out, want)
    scratch = tmp.numel() * tmp.element_size()
    if head_size == 128:
        assert grew < scratch // 8, grew
    else:
        assert grew >= scratch, grew


# ------------------------------------------------- one launch per step, shared workspace

def test_back_to_back_launches_share_the_workspace(ops, native):
    """ABI v7: the cross-CTA merge is inside the decode kernel (ticket counters in the workspace
    header, left at zero).  Many launches back to back on one stream -- no host synchronisation,
    programmatic stream serialisation between them, shapes with different slot sizes sharing
    the one cached workspace, sequences short enough that a launch lasts microseconds -- must
    each give exactly what the same call gives alone."""
    cases = []
    rnd = random.Random(11)
    for B, H, Hkv, D, kv, lens in [
        (40, 16, 2, 128, "fp8", [rnd.randint(1, 32) for _ in range(40)]),       # all <= 32 tokens
        (3, 8, 1, 64, "auto", [3000, 17, 900]),
        (1, 8, 1, 128, "fp8", [40000]),                                         # ~all CTAs share one item
        (7, 32, 8, 256, "fp8", [rnd.randint(1, 3000) for _ in range(7)]),
        (64, 8, 8, 128, "auto", [rnd.randint(1, 200) for _ in range(64)]),
    ]:
        BS = 16
        NB = sum((L + BS - 1) // BS for L in lens) + 2
        s = D ** -0.5
        q, kc, vc, bt, sl = hp.make_paged_inputs(B, H, Hkv, D, BS, NB, lens, torch.bfloat16,
                                                 seed=B + D, kv=kv, lo=-1.0, hi=1.0, permute=True)
        ks = torch.tensor(1.0 if kv == "auto" else 0.5, device=DEV)
        vs = torch.tensor(1.0 if kv == "auto" else 1.5, device=DEV)
        cases.append(dict(q=q.to(DEV), kc=kc.to(DEV), vc=vc.to(DEV), bt=bt.to(DEV), sl=sl.to(DEV),
                          Hkv=Hkv, s=s, max_len=max(lens), kv=kv, ks=ks, vs=vs))

    def call(c, out):
        ops.paged_attention_v1(out, c["q"], c["kc"], c["vc"], c["Hkv"], c["s"], c["bt"], c["sl"], 16,
                               c["max_len"], None, c["kv"], c["ks"], c["vs"])

    native.set_impl(STREAM)
    try:
        alone = []
        for c in cases:
            o = torch.empty_like(c["q"])
            call(c, o)
            torch.cuda.synchronize()
            alone.append(o)
        n0 = native.launch_count()
        outs = [[torch.full_like(c["q"], float("nan")) for c in cases] for _ in range(40)]
        for rep in range(40):
            for c, o in zip(cases, outs[rep]):
                call(c, o)
        torch.cuda.synchronize()
        assert native.launch_count() - n0 == 40 * len(cases)       # one kernel per call
        for rep in range(40):
            for a, o in zip(alone, outs[rep]):
                assert torch.equal(a, o), f"repetition {rep} differs from the call alone"
    finally:
        native.set_impl(native.IMPL_AUTO)
    # ... and the stream results are the generic kernel's
    native.set_impl(GENERIC)
    try:
        for c, a in zip(cases, alone):
            o = torch.empty_like(c["q"])
            call(c, o)
            torch.cuda.synchronize()
            _close(a.cpu(), o.cpu(), atol=ATOL * max(1.0, float(o.float().abs().max())))
    finally:
        native.set_impl(native.IMPL_AUTO)

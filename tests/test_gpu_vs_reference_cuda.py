"""GPU parity against the REFERENCE'S OWN CUDA kernels (oracle tier O1).

``oracle/_ref/libvllm_ref_cuda.so`` is ``csrc/attention/paged_attention_v1.cu``,
``paged_attention_v2.cu`` and ``csrc/cache_kernels.cu`` of the reference compiled unmodified for
sm_100a (``oracle/build_ref_cuda.sh``) and registered under ``torch.ops.vllm_ref_cuda``.  Both
sides get the identical device tensors:

* cache writes / fp8 conversion: BIT-EXACT (``torch.equal`` on the bytes);
* attention: BASELINE.json's gate, ``|out - ref| <= 1e-2 + 1e-2 |ref|`` on the five configs
  (reduced batch, full context) and on a ragged sweep; v2's exported partials are not compared
  (the stream kernel keeps its partials in its own workspace, include/b200_paged_attn.h).

Skipped when the library is absent (it is built in the container that has /root/reference and
travels with the snapshot).
"""
import os
import random

import pytest
import torch

from tests import helpers as hp
from tests.test_gpu_attention import _config_inputs

pytestmark = pytest.mark.gpu

DEV = "cuda:0"
ATOL, RTOL = 1e-2, 1e-2
GENERIC, STREAM = 1, 2


@pytest.fixture(scope="module")
def ref():
    from oracle import ref_cuda
    if not os.path.exists(ref_cuda.LIB_PATH):
        pytest.skip("oracle/_ref/libvllm_ref_cuda.so not built")
    ref_cuda.load()
    return ref_cuda


@pytest.fixture(scope="module")
def ops():
    from vllm_main_modify_b200 import _custom_ops
    return _custom_ops


@pytest.fixture(scope="module")
def native():
    from vllm_main_modify_b200 import _native
    _native.load()
    return _native


def _attn(fn_v1, fn_v2, version, q, kc, vc, Hkv, s, bt, sl, BS, max_len, alibi, kv, ks, vs):
    out = torch.full_like(q, float("nan"))
    B, H, D = q.shape
    if version == "v1":
        fn_v1(out, q, kc, vc, Hkv, s, bt, sl, BS, max_len, alibi, kv, ks, vs, 0, 0, 0, 64, 0)
    else:
        P = (max_len + 511) // 512
        es = torch.empty(B, H, P, dtype=torch.float32, device=q.device)
        ml = torch.empty_like(es)
        tmp = torch.empty(B, H, P, D, dtype=q.dtype, device=q.device)
        fn_v2(out, es, ml, tmp, q, kc, vc, Hkv, s, bt, sl, BS, max_len, alibi, kv, ks, vs,
              0, 0, 0, 64, 0)
    torch.cuda.synchronize()
    return out


def _both(ops, ref, native, impl, version, q, kc, vc, Hkv, s, bt, sl, BS, max_len, alibi, kv, ks,
          vs):
    dq, dkc, dvc, dbt, dsl = (t.to(DEV) for t in (q, kc, vc, bt, sl))
    dal = None if alibi is None else alibi.to(DEV)
    dks = torch.tensor(float(ks), dtype=torch.float32, device=DEV)
    dvs = torch.tensor(float(vs), dtype=torch.float32, device=DEV)
    native.set_impl(impl)
    try:
        got = _attn(ops.paged_attention_v1, ops.paged_attention_v2, version, dq, dkc, dvc, Hkv, s,
                    dbt, dsl, BS, max_len, dal, kv, dks, dvs)
    finally:
        native.set_impl(native.IMPL_AUTO)
    want = _attn(ref.paged_attention_v1, ref.paged_attention_v2, version, dq, dkc, dvc, Hkv, s,
                 dbt, dsl, BS, max_len, dal, kv, dks, dvs)
    return got.cpu(), want.cpu()


def _close(got, want, atol=ATOL, rtol=RTOL):
    assert not torch.isnan(got.float()).any()
    torch.testing.assert_close(got.float(), want.float(), atol=atol, rtol=rtol)


@pytest.mark.parametrize("name", ["cfg1", "cfg2", "cfg3", "cfg4", "cfg4_tp1", "cfg5"])
@pytest.mark.parametrize("version", ["v1", "v2"])
@pytest.mark.parametrize("impl", [STREAM, GENERIC])
def test_baseline_configs_match_reference_cuda(ops, ref, native, name, version, impl):
    q, kc, vc, bt, sl, H, Hkv, D, BS, s, kv, ks, vs, max_len = _config_inputs(name)
    if version == "v1" and max_len * 4 > 200 * 1024:
        # the reference's v1 kernel keeps max_seq_len fp32 logits in shared memory
        # (paged_attention_v1.cu:72-80): contexts past ~50k tokens cannot launch
        pytest.skip("context too long for the reference's v1 kernel")
    got, want = _both(ops, ref, native, impl, version, q, kc, vc, Hkv, s, bt, sl, BS, max_len,
                      None, kv, ks, vs)
    _close(got, want)


@pytest.mark.parametrize("kv", ["auto", "fp8"])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("shape", [(40, 40, 64), (64, 8, 128), (16, 2, 128), (12, 4, 80),
                                   (32, 32, 256)])
@pytest.mark.parametrize("use_alibi", [False, True])
def test_ragged_sweep_matches_reference_cuda(ops, ref, native, kv, dtype, shape, use_alibi):
    H, Hkv, D = shape
    BS, B, MAX, NB = 16, 9, 3000, 1999
    rnd = random.Random(5)
    lens = [rnd.randint(1, MAX) for _ in range(B)]
    lens[0], lens[-1] = 1, MAX
    s = D ** -0.5
    q, kc, vc, bt, sl = hp.make_paged_inputs(B, H, Hkv, D, BS, NB, lens, dtype, seed=4, kv=kv,
                                             lo=-s, hi=s, permute=True)
    alibi = None
    if use_alibi:
        alibi = torch.randn(H, generator=torch.Generator().manual_seed(2), dtype=torch.float32)
    ks, vs = (1.0, 1.0) if kv == "auto" else (s / 2.0, s / 3.0)
    for version in ("v1", "v2"):
        got, want = _both(ops, ref, native, 0, version, q, kc, vc, Hkv, s, bt, sl, BS, MAX, alibi,
                          kv, ks, vs)
        _close(got, want)


# ------------------------------------------------------------------ cache ops: bit-exact

def _bytes(t):
    return t.contiguous().view(torch.uint8).cpu()


@pytest.mark.parametrize("kv", ["auto", "fp8"])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("head_size", [64, 80, 128, 256])
@pytest.mark.parametrize("block_size", [8, 16, 32])
def test_reshape_and_cache_bit_exact_vs_reference_cuda(ops, ref, kv, dtype, head_size,
                                                        block_size):
    T, Hkv, NB = 42, 8, 257
    esz = 1 if kv != "auto" else torch.tensor([], dtype=dtype).element_size()
    x = 16 // esz
    if head_size % x:
        pytest.skip("head_size not divisible by x")
    g = torch.Generator().manual_seed(0)
    qkv = torch.randn(T, 3, Hkv, head_size, generator=g).to(dtype).to(DEV)
    key, value = qkv[:, 1], qkv[:, 2]   # strided like a fused qkv projection
    slots = random.Random(0).sample(range(NB * block_size), T)
    slots[5] = -1
    slots = torch.tensor(slots, dtype=torch.int64, device=DEV)
    cdt = dtype if kv == "auto" else torch.uint8
    ks = (key.float().abs().max() / 64.0).reshape(()) if kv != "auto" else torch.tensor(1.0, device=DEV)
    vs = (value.float().abs().max() / 64.0).reshape(()) if kv != "auto" else torch.tensor(1.0, device=DEV)
    res = []
    for fn in (ops.reshape_and_cache, ref.reshape_and_cache):
        kc = torch.zeros(NB, Hkv, head_size // x, block_size, x, dtype=cdt, device=DEV)
        vc = torch.zeros(NB, Hkv, head_size, block_size, dtype=cdt, device=DEV)
        fn(key, value, kc, vc, slots, kv, ks, vs)
        torch.cuda.synchronize()
        res.append((_bytes(kc), _bytes(vc)))
    assert torch.equal(res[0][0], res[1][0])
    assert torch.equal(res[0][1], res[1][1])


@pytest.mark.parametrize("kv", ["auto", "fp8"])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("layout", ["NHD", "HND"])
@pytest.mark.parametrize("head_size", [64, 80, 128])
def test_reshape_and_cache_flash_bit_exact_vs_reference_cuda(ops, ref, kv, dtype, layout,
                                                              head_size):
    T, Hkv, NB, BS = 55, 4, 123, 16
    g = torch.Generator().manual_seed(1)
    qkv = torch.randn(T + 3, 3, Hkv, head_size, generator=g).to(dtype).to(DEV)
    key, value = qkv[:, 1], qkv[:, 2]
    slots = random.Random(1).sample(range(NB * BS), T)   # T < key.size(0): padded batch
    slots[7] = -1
    slots = torch.tensor(slots, dtype=torch.int64, device=DEV)
    cdt = dtype if kv == "auto" else torch.uint8
    ks = (key.float().abs().max() / 64.0).reshape(()) if kv != "auto" else torch.tensor(1.0, device=DEV)
    vs = (value.float().abs().max() / 64.0).reshape(()) if kv != "auto" else torch.tensor(1.0, device=DEV)
    res = []
    for fn in (ops.reshape_and_cache_flash, ref.reshape_and_cache_flash):
        if layout == "NHD":
            kc = torch.zeros(NB, BS, Hkv, head_size, dtype=cdt, device=DEV)
            vc = torch.zeros_like(kc)
        else:
            kc = torch.zeros(NB, Hkv, BS, head_size, dtype=cdt, device=DEV).permute(0, 2, 1, 3)
            vc = torch.zeros(NB, Hkv, BS, head_size, dtype=cdt, device=DEV).permute(0, 2, 1, 3)
        fn(key, value, kc, vc, slots, kv, ks, vs)
        torch.cuda.synchronize()
        res.append((_bytes(kc), _bytes(vc)))
    assert torch.equal(res[0][0], res[1][0])
    assert torch.equal(res[0][1], res[1][1])


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("scale", [1.0, 0.37])
def test_convert_fp8_bit_exact_vs_reference_cuda(ops, ref, dtype, scale):
    g = torch.Generator().manual_seed(2)
    src = (torch.randn(64, 8, 128, generator=g) * 3).to(dtype).to(DEV)
    a = torch.empty(src.shape, dtype=torch.uint8, device=DEV)
    b = torch.empty_like(a)
    ops.convert_fp8(a, src, scale, "fp8")
    ref.convert_fp8(b, src, scale, "fp8")
    torch.cuda.synchronize()
    assert torch.equal(a.cpu(), b.cpu())
    back_a, back_b = torch.empty_like(src), torch.empty_like(src)
    ops.convert_fp8(back_a, a, scale, "fp8")
    ref.convert_fp8(back_b, a, scale, "fp8")
    torch.cuda.synchronize()
    assert torch.equal(_bytes(back_a), _bytes(back_b))

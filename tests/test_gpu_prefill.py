"""GPU parity of ops.prefill_attention (csrc/paged_prefill.cu) -- causal attention of query
chunks over the paged cache, SURVEY 8f row 1 -- against the oracle (oracle/prefill_oracle.py,
pinned on the reference's ``ref_paged_attn``), the committed golden outputs of ``ref_paged_attn``
itself, and the decode kernel.  Gate: |out - ref| <= 1e-2 + 1e-2 |ref| (north_star's tolerance).
Modelled on the reference's tests/kernels/attention/test_flash_attn.py:82-176 (varlen with
block tables: mixed query / kv lengths, head sizes, block sizes, dtypes)."""
import random

import pytest
import torch

from oracle import prefill_oracle as po
from tests import helpers as hp
from tests.test_prefill_oracle import CASES, load_case

pytestmark = pytest.mark.gpu

ATOL = RTOL = 1e-2


def ops():
    from vllm_main_modify_b200 import _custom_ops
    return _custom_ops


def make_inputs(query_lens, ctx_lens, H, Hkv, D, BS, dtype, kv, seed, spread=1.0):
    """Seeded synthetic batch; fp8 caches hold finite codes of both signs."""
    g = torch.Generator().manual_seed(seed)
    rnd = random.Random(seed)
    seq_lens = [a + b for a, b in zip(query_lens, ctx_lens)]
    nblk = [(L + BS - 1) // BS for L in seq_lens]
    NB = sum(nblk) + 3
    T = sum(query_lens)
    q = ((torch.rand(T, H, D, generator=g) * 2 - 1) * spread).to(dtype)
    if kv == "auto":
        x = 16 // torch.tensor([], dtype=dtype).element_size()
        kc = (torch.rand(NB, Hkv, D // x, BS, x, generator=g) * 2 - 1).to(dtype)
        vc = (torch.rand(NB, Hkv, D, BS, generator=g) * 2 - 1).to(dtype)
    else:
        top = 0x48 if kv != "fp8_e5m2" else 0x44      # |value| <= 2 in both formats
        def codes(shape):
            mag = torch.randint(0, top, shape, generator=g, dtype=torch.int32)
            sgn = torch.randint(0, 2, shape, generator=g, dtype=torch.int32) << 7
            return (mag | sgn).to(torch.uint8)
        kc = codes((NB, Hkv, D // 16, BS, 16))
        vc = codes((NB, Hkv, D, BS))
    perm = list(range(NB))
    rnd.shuffle(perm)
    bt = torch.zeros(len(seq_lens), max(nblk), dtype=torch.int32)
    pos = 0
    for b, n in enumerate(nblk):
        bt[b, :n] = torch.tensor(perm[pos:pos + n], dtype=torch.int32)
        pos += n
    qsl = torch.tensor([0] + torch.tensor(query_lens).cumsum(0).tolist(), dtype=torch.int32)
    sl = torch.tensor(seq_lens, dtype=torch.int32)
    return q, kc, vc, bt, qsl, sl


def run_op(q, kc, vc, bt, qsl, sl, scale, kv="auto", ks=1.0, vs=1.0, alibi=None, hpt=0,
           want_lse=False, max_q=None):
    dev = "cuda:0"
    out = torch.full_like(q, float("nan"), device=dev)
    lse = torch.full((q.shape[0], q.shape[1]), float("nan"), device=dev) if want_lse else None
    if max_q is None:
        max_q = int((qsl[1:] - qsl[:-1]).max())
    ops().prefill_attention(out, lse, q.to(dev), kc.to(dev), vc.to(dev), scale, bt.to(dev),
                            qsl.to(dev), sl.to(dev), max_q,
                            None if alibi is None else alibi.to(dev), kv,
                            torch.tensor(ks, device=dev), torch.tensor(vs, device=dev), hpt)
    torch.cuda.synchronize()
    return (out.cpu(), lse.cpu()) if want_lse else out.cpu()


def assert_close(got, want, what=""):
    g, w = got.float(), want.float()
    assert torch.isfinite(g).all(), f"{what}: non-finite output"
    viol = (g - w).abs() - (ATOL + RTOL * w.abs())
    assert float(viol.max()) <= 0, f"{what}: max abs diff {float((g - w).abs().max()):.4g}"


@pytest.mark.parametrize("name", CASES)
def test_golden_ref_paged_attn(name):
    """The reference's own oracle outputs (16-bit arithmetic noise: 2e-2 like test_flash_attn)."""
    c = load_case(name)
    got = run_op(c["q"], c["kc"], c["vc"], c["bt"], c["qsl"], c["sl"], c["scale"])
    torch.testing.assert_close(got.float(), c["want"].float(), atol=2e-2, rtol=2e-2)
    want = po.prefill_attention(c["q"], c["kc"], c["vc"], c["bt"], c["qsl"].tolist(),
                                c["sl"].tolist(), c["scale"])
    assert_close(got, want, name)


MIX_Q = [1, 5, 64, 65, 130, 17]
MIX_CTX = [37, 0, 3, 100, 0, 250]


@pytest.mark.parametrize("head_size", [32, 64, 96, 128, 192, 256])
@pytest.mark.parametrize("block_size", [16, 32])
@pytest.mark.parametrize("kv", ["auto", "fp8", "fp8_e5m2"])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
def test_sweep_mixed_batch(head_size, block_size, kv, dtype):
    H, Hkv = 8, 2
    scale = head_size ** -0.5
    ks, vs = (1.0, 1.0) if kv == "auto" else (0.37, 0.61)
    q, kc, vc, bt, qsl, sl = make_inputs(MIX_Q, MIX_CTX, H, Hkv, head_size, block_size, dtype, kv,
                                         seed=head_size + block_size)
    want = po.prefill_attention(q, kc, vc, bt, qsl.tolist(), sl.tolist(), scale, None, kv, ks, vs)
    got = run_op(q, kc, vc, bt, qsl, sl, scale, kv, ks, vs)
    assert_close(got, want, f"D={head_size} BS={block_size} {kv} {dtype}")


@pytest.mark.parametrize("group,hpt", [(1, 0), (4, 0), (4, 1), (4, 2), (8, 0), (8, 8), (7, 0), (7, 7),
                                       (16, 16), (16, 0)])
@pytest.mark.parametrize("kv", ["auto", "fp8"])
def test_head_packing(group, hpt, kv):
    """Short chunks (speculative decoding: 1-8 tokens) with the GQA group stacked on the tile."""
    Hkv, D, BS = 2, 128, 16
    H = Hkv * group
    qlens, ctx = [4, 1, 8, 2, 3], [100, 9, 0, 511, 64]
    q, kc, vc, bt, qsl, sl = make_inputs(qlens, ctx, H, Hkv, D, BS, torch.bfloat16, kv, seed=group)
    ks, vs = (1.0, 1.0) if kv == "auto" else (0.5, 0.25)
    want = po.prefill_attention(q, kc, vc, bt, qsl.tolist(), sl.tolist(), 0.09, None, kv, ks, vs)
    got = run_op(q, kc, vc, bt, qsl, sl, 0.09, kv, ks, vs, hpt=hpt)
    assert_close(got, want, f"G={group} heads_per_tile={hpt} {kv}")


@pytest.mark.parametrize("kv", ["auto", "fp8"])
def test_alibi_and_lse(kv):
    H, Hkv, D, BS = 8, 4, 64, 16
    qlens, ctx = [70, 3, 1], [10, 200, 33]
    q, kc, vc, bt, qsl, sl = make_inputs(qlens, ctx, H, Hkv, D, BS, torch.float16, kv, seed=11)
    slopes = torch.tensor([2.0 ** -(i + 1) for i in range(H)])
    ks, vs = (1.0, 1.0) if kv == "auto" else (0.3, 0.9)
    want, want_lse = po.prefill_attention(q, kc, vc, bt, qsl.tolist(), sl.tolist(), 0.125, slopes,
                                          kv, ks, vs, return_lse=True)
    got, lse = run_op(q, kc, vc, bt, qsl, sl, 0.125, kv, ks, vs, alibi=slopes, want_lse=True)
    assert_close(got, want, "alibi")
    torch.testing.assert_close(lse, want_lse, atol=2e-3, rtol=2e-3)


@pytest.mark.parametrize("kv", ["auto", "fp8"])
def test_tail_of_the_last_page_is_ignored(kv):
    """Cache bytes past seq_len (NaN patterns) and unused pages must not leak into the output."""
    H, Hkv, D, BS = 4, 2, 128, 16
    qlens, ctx = [5, 33], [6, 40]
    q, kc, vc, bt, qsl, sl = make_inputs(qlens, ctx, H, Hkv, D, BS, torch.bfloat16, kv, seed=2)
    want = po.prefill_attention(q, kc, vc, bt, qsl.tolist(), sl.tolist(), 0.1, None, kv, 0.5, 0.5)
    kc2, vc2 = kc.clone(), vc.clone()
    nan16 = torch.tensor(float("nan"), dtype=torch.bfloat16)
    for b, L in enumerate(sl.tolist()):
        blk = int(bt[b, (L - 1) // BS])
        off = L % BS
        if off == 0:
            continue
        if kv == "auto":
            kc2[blk, :, :, off:, :] = nan16
            vc2[blk, :, :, off:] = nan16
        else:
            kc2[blk, :, :, off:, :] = 0x7f
            vc2[blk, :, :, off:] = 0x7f
    got = run_op(q, kc2, vc2, bt, qsl, sl, 0.1, kv, 0.5, 0.5)
    assert_close(got, want, "tail garbage")


@pytest.mark.parametrize("kv", ["auto", "fp8"])
def test_single_token_chunks_equal_the_decode_kernel(kv):
    """q_len = 1 for every sequence: the same numbers as paged_attention_v1 (within rounding)."""
    B, H, Hkv, D, BS, NB = 5, 16, 4, 128, 16, 80
    lens = [1, 16, 17, 300, 513]
    q, kc, vc, bt, sl = hp.make_paged_inputs(B, H, Hkv, D, BS, NB, lens, torch.bfloat16, seed=4,
                                             kv=kv, permute=True)
    dev = "cuda:0"
    ks = torch.tensor(0.4 if kv != "auto" else 1.0, device=dev)
    vs = torch.tensor(0.7 if kv != "auto" else 1.0, device=dev)
    out1 = torch.empty_like(q, device=dev)
    ops().paged_attention_v1(out1, q.to(dev), kc.to(dev), vc.to(dev), Hkv, 0.09, bt.to(dev),
                             sl.to(dev), BS, max(lens), None, kv, ks, vs)
    qsl = torch.arange(B + 1, dtype=torch.int32)
    got = run_op(q, kc, vc, bt, qsl, sl, 0.09, kv, float(ks), float(vs))
    assert_close(got, out1.cpu(), "decode kernel")


def test_long_chunk_with_context_and_empty_sequences():
    """A 1100-token chunk over 900 tokens of context, an empty slot in the batch (q_len 0) and
    a max_query_len larger than every chunk (padding of the grid)."""
    H, Hkv, D, BS = 4, 1, 128, 16
    qlens, ctx = [1100, 0, 40], [900, 77, 0]
    q, kc, vc, bt, qsl, sl = make_inputs(qlens, ctx, H, Hkv, D, BS, torch.bfloat16, "auto", seed=9,
                                         spread=2.0)
    want = po.prefill_attention(q, kc, vc, bt, qsl.tolist(), sl.tolist(), 0.09)
    got = run_op(q, kc, vc, bt, qsl, sl, 0.09, max_q=2048)
    assert_close(got, want, "long chunk")


@pytest.mark.parametrize("hpt", [0, 1])
def test_sequence_shorter_than_its_chunk_gets_a_defined_result(hpt):
    """seq_lens[b] < q_len (inconsistent metadata; the reference would index before the cache):
    those rows come back as zeros with lse = -inf instead of stale memory, the other sequences
    are unaffected; tensors on the wrong device are rejected by the shim."""
    H, Hkv, D, BS = 8, 2, 128, 16
    qlens, ctx = [70, 30, 5], [100, 20, 0]
    q, kc, vc, bt, qsl, sl = make_inputs(qlens, ctx, H, Hkv, D, BS, torch.bfloat16, "auto", seed=4)
    want, want_lse = po.prefill_attention(q, kc, vc, bt, qsl.tolist(), sl.tolist(), 0.09,
                                          return_lse=True)
    bad = sl.clone()
    bad[1] = 12                                    # 30 query tokens, 12 tokens in the sequence
    got, lse = run_op(q, kc, vc, bt, qsl, bad, 0.09, hpt=hpt, want_lse=True)
    rows = slice(int(qsl[1]), int(qsl[2]))
    assert float(got[rows].float().abs().max()) == 0.0
    assert bool((lse[rows] == float("-inf")).all())
    keep = torch.ones(q.shape[0], dtype=torch.bool)
    keep[rows] = False
    assert_close(got[keep], want[keep], "other sequences")
    torch.testing.assert_close(lse[keep], want_lse[keep].float(), atol=1e-2, rtol=1e-2)
    dev = "cuda:0"
    one = torch.tensor(1.0, device=dev)
    with pytest.raises(RuntimeError, match="CUDA device"):
        ops().prefill_attention(torch.empty_like(q, device=dev), None, q.to(dev), kc.to(dev),
                                vc.to(dev), 0.09, bt, qsl.to(dev), sl.to(dev), 70, None, "auto",
                                one, one)


def test_fp8_cache_with_large_bf16_queries_inside_the_f16_range():
    """With an fp8 cache the chunk kernel computes in f16 (fp8 -> f16 is exact; a bf16 query is
    converted, which is exact too while |q| <= 65504).  Documented limit: bf16 queries beyond the
    f16 range overflow on this path (the reference dequantises to the query dtype instead).
    Inside the range the result is the oracle's: queries of magnitude 300, far above anything
    the other tests use, with a scale that keeps the logits moderate."""
    H, Hkv, D, BS = 8, 2, 128, 16
    qlens, ctx = [40, 3, 1], [200, 77, 500]
    q, kc, vc, bt, qsl, sl = make_inputs(qlens, ctx, H, Hkv, D, BS, torch.bfloat16, "fp8", seed=13,
                                         spread=300.0)
    assert float(q.float().abs().max()) > 250
    want = po.prefill_attention(q, kc, vc, bt, qsl.tolist(), sl.tolist(), 1e-3, None, "fp8",
                                0.5, 0.25)
    got = run_op(q, kc, vc, bt, qsl, sl, 1e-3, kv="fp8", ks=0.5, vs=0.25)
    assert_close(got, want, "large bf16 queries, fp8 cache")


def test_strided_query_and_out():
    """query / out as slices of a fused qkv buffer (token stride > H * D)."""
    H, Hkv, D, BS = 4, 2, 64, 16
    qlens, ctx = [20, 3], [5, 60]
    q, kc, vc, bt, qsl, sl = make_inputs(qlens, ctx, H, Hkv, D, BS, torch.float16, "auto", seed=1)
    dev = "cuda:0"
    T = q.shape[0]
    qkv = torch.zeros(T, 3, H, D, dtype=q.dtype, device=dev)
    qkv[:, 0] = q.to(dev)
    buf = torch.zeros(T, 2, H, D, dtype=q.dtype, device=dev)
    ops().prefill_attention(buf[:, 1], None, qkv[:, 0], kc.to(dev), vc.to(dev), 0.2, bt.to(dev),
                            qsl.to(dev), sl.to(dev), max(qlens), None, "auto",
                            torch.tensor(1.0, device=dev), torch.tensor(1.0, device=dev))
    want = po.prefill_attention(q, kc, vc, bt, qsl.tolist(), sl.tolist(), 0.2)
    assert_close(buf[:, 1].cpu(), want, "strided")
    assert float(buf[:, 0].abs().max()) == 0.0


def test_unsupported_configurations_raise():
    H, Hkv, D, BS = 4, 2, 80, 16
    q, kc, vc, bt, qsl, sl = make_inputs([3], [4], H, Hkv, 64, BS, torch.float16, "auto", seed=0)
    dev = "cuda:0"
    one = torch.tensor(1.0, device=dev)
    q80 = torch.zeros(3, H, D, dtype=torch.float16, device=dev)
    kc80 = torch.zeros(4, Hkv, D // 8, BS, 8, dtype=torch.float16, device=dev)
    vc80 = torch.zeros(4, Hkv, D, BS, dtype=torch.float16, device=dev)
    with pytest.raises(RuntimeError, match="outside the kernel's domain"):
        ops().prefill_attention(torch.empty_like(q80), None, q80, kc80, vc80, 0.1, bt.to(dev),
                                qsl.to(dev), sl.to(dev), 3, None, "auto", one, one)
    with pytest.raises(RuntimeError, match="heads_per_tile"):
        ops().prefill_attention(torch.empty_like(q, device=dev), None, q.to(dev), kc.to(dev),
                                vc.to(dev), 0.1, bt.to(dev), qsl.to(dev), sl.to(dev), 3, None,
                                "auto", one, one, 3)

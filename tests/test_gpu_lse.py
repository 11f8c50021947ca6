"""GPU parity of the log-sum-exp side of the decode path:

* ``paged_attention_lse`` (decode + LSE; what ``can_return_lse_for_decode``,
  vllm/attention/backends/abstract.py:115-138 of the reference, asks of a backend) against
  oracle/lse_oracle.py on shapes served by the stream kernel and by the generic kernel;
* ``merge_attn_states`` against the golden fixtures (the reference's own torch implementation),
  the oracle, and -- bit-exact -- the reference's CUDA kernel (oracle/_ref) where it is built;
* ``correct_attn_out`` against the oracle, and the decode-context-parallel property end to end
  on the GPU ops: partial decode + LSE per rank, correction, sum over ranks == full decode.
"""
import os

import pytest
import torch

from tests import helpers as hp

pytestmark = pytest.mark.gpu

DEV = "cuda:0"
GENERIC, STREAM = 1, 2


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
def lora():
    from oracle import lse_oracle
    return lse_oracle


def _gpu_lse(ops, q, kc, vc, Hkv, scale, bt, sl, BS, max_len, alibi, kv, ks, vs, lse_layout="bh"):
    B, H, D = q.shape
    qd = q.to(DEV)
    out = torch.full_like(qd, float("nan"))
    if lse_layout == "bh":
        lse = torch.full((B, H), float("nan"), dtype=torch.float32, device=DEV)
    else:                       # merge_attn_states' [H, B] layout, viewed as [B, H]
        lse = torch.full((H, B), float("nan"), dtype=torch.float32, device=DEV).t()
    ops.paged_attention_lse(out, lse, qd, kc.to(DEV), vc.to(DEV), Hkv, scale, bt.to(DEV),
                            sl.to(DEV), BS, max_len, None if alibi is None else alibi.to(DEV),
                            kv, torch.tensor(ks, device=DEV), torch.tensor(vs, device=DEV))
    torch.cuda.synchronize()
    return out.cpu(), lse.cpu()


CASES = [
    # name, impl, B, H, Hkv, D, BS, lens, dtype, kv, alibi, layout
    ("stream_gqa8_fp8", STREAM, 5, 16, 2, 128, 16, [0, 1, 777, 5000, 20001], torch.bfloat16, "fp8", False, "bh"),
    ("stream_gqa4_bf16", STREAM, 4, 8, 2, 128, 16, [16, 513, 1025, 3000], torch.bfloat16, "auto", False, "hb"),
    ("stream_mha_f16_d64_alibi", STREAM, 3, 4, 4, 64, 16, [5, 129, 2000], torch.float16, "auto", True, "bh"),
    ("stream_one_long", STREAM, 1, 8, 1, 128, 16, [70000], torch.bfloat16, "fp8", False, "bh"),
    ("generic_f16_d80", GENERIC, 3, 6, 3, 80, 16, [0, 32, 600], torch.float16, "auto", False, "bh"),
    ("generic_bf16_bs32_fp8", GENERIC, 3, 8, 1, 128, 32, [7, 64, 555], torch.bfloat16, "fp8", False, "hb"),
    ("generic_f32_d32_alibi", GENERIC, 2, 2, 2, 32, 8, [9, 100], torch.float32, "auto", True, "bh"),
]


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_paged_attention_lse_matches_oracle(ops, native, lora, case):
    name, impl, B, H, Hkv, D, BS, lens, dtype, kv, use_alibi, layout = case
    s = D ** -0.5
    NB = sum((l + BS - 1) // BS for l in lens) + 3
    q, kc, vc, bt, sl = hp.make_paged_inputs(B, H, Hkv, D, BS, NB, lens, dtype, 3, kv=kv,
                                             lo=-1.0, hi=1.0, permute=True)
    ks, vs = (1.0, 1.0) if kv == "auto" else (0.37, 0.21)
    alibi = None
    if use_alibi:
        alibi = torch.randn(H, generator=torch.Generator().manual_seed(1)) * 0.05
    max_len = max(max(lens), 1)
    want_out = torch.empty_like(q)
    want_lse = torch.empty(B, H)
    lora.paged_attention_lse(want_out, want_lse, q, kc, vc, Hkv, s, bt, sl, BS, max_len, alibi,
                             kv, torch.tensor(ks), torch.tensor(vs))
    native.set_impl(impl)
    try:
        out, lse = _gpu_lse(ops, q, kc, vc, Hkv, s, bt, sl, BS, max_len, alibi, kv, ks, vs, layout)
    finally:
        native.set_impl(native.IMPL_AUTO)
    empty = sl == 0
    assert torch.isinf(lse[empty]).all() and (lse[empty] < 0).all()
    assert (out[empty].float() == 0).all()
    torch.testing.assert_close(lse[~empty], want_lse[~empty], atol=2e-3, rtol=1e-4)
    torch.testing.assert_close(out[~empty].float(), want_out[~empty].float(), atol=1e-2, rtol=1e-2)


def test_lse_call_leaves_plain_v1_result_unchanged(ops):
    """The LSE entry point is v1 plus one extra store: same output bits."""
    B, H, Hkv, D, BS = 6, 32, 4, 128, 16
    lens = [100, 5000, 17, 9000, 2048, 333]
    NB = sum((l + BS - 1) // BS for l in lens) + 1
    q, kc, vc, bt, sl = hp.make_paged_inputs(B, H, Hkv, D, BS, NB, lens, torch.bfloat16, 9, kv="fp8",
                                             permute=True)
    s = D ** -0.5
    ks, vs = 0.5, 0.25
    out, _ = _gpu_lse(ops, q, kc, vc, Hkv, s, bt, sl, BS, max(lens), None, "fp8", ks, vs)
    qd = q.to(DEV)
    out1 = torch.empty_like(qd)
    ops.paged_attention_v1(out1, qd, kc.to(DEV), vc.to(DEV), Hkv, s, bt.to(DEV), sl.to(DEV), BS,
                           max(lens), None, "fp8", torch.tensor(ks, device=DEV),
                           torch.tensor(vs, device=DEV))
    assert torch.equal(out1.cpu(), out)


# ------------------------------------------------------------------ merge_attn_states
@pytest.mark.parametrize("name", ["merge_bf16_d128.npz", "merge_f16_d96.npz", "merge_f32_d32.npz"])
def test_merge_attn_states_golden(ops, name):
    g = hp.load_golden(name)
    dt = hp.golden_dtype(g)
    p_out = hp.from_bits(g["prefix_output"], dt).to(DEV)
    s_out = hp.from_bits(g["suffix_output"], dt).to(DEV)
    p_lse = torch.from_numpy(g["prefix_lse"].copy()).to(DEV)
    s_lse = torch.from_numpy(g["suffix_lse"].copy()).to(DEV)
    out = torch.full_like(p_out, float("nan"))
    out_lse = torch.full_like(p_lse, float("nan"))
    ops.merge_attn_states(out, p_out, p_lse, s_out, s_lse, out_lse)
    want = torch.from_numpy(g["output_f32"]).to(dt).float()
    tol = dict(atol=0, rtol=2 ** -7) if dt != torch.float32 else dict(atol=1e-6, rtol=1e-5)
    torch.testing.assert_close(out.cpu().float(), want, **tol)
    want_lse = torch.from_numpy(g["output_lse"])
    fin = torch.isfinite(want_lse)
    got_lse = out_lse.cpu()
    torch.testing.assert_close(got_lse[fin], want_lse[fin], atol=1e-5, rtol=1e-5)
    assert torch.equal(torch.isfinite(got_lse), fin)
    # output_lse is optional
    out2 = torch.empty_like(out)
    ops.merge_attn_states(out2, p_out, p_lse, s_out, s_lse)
    assert torch.equal(out2, out)


@pytest.mark.parametrize("dtype", [torch.float32, torch.float16, torch.bfloat16])
@pytest.mark.parametrize("shape", [(613, 16, 128), (256, 4, 32), (1024, 48, 96), (37, 64, 256)])
def test_merge_attn_states_bit_exact_vs_reference_cuda(ops, dtype, shape):
    """Same sweep axes as tests/kernels/attention/test_merge_attn_states.py:49-52 of the
    reference, against its CUDA kernel compiled unmodified (oracle/_ref)."""
    from oracle import ref_cuda
    if not os.path.exists(ref_cuda.LIB_PATH):
        pytest.skip("oracle/_ref/libvllm_ref_cuda.so not built")
    ref_cuda.load()
    if not hasattr(torch.ops.vllm_ref_cuda, "merge_attn_states"):
        pytest.skip("oracle/_ref/libvllm_ref_cuda.so predates merge_attn_states")
    n, h, d = shape
    g = torch.Generator(device=DEV).manual_seed(n + h + d)
    p_lse = torch.randn(h, n, device=DEV, generator=g)
    s_lse = torch.randn(h, n, device=DEV, generator=g)
    mp = torch.rand(h, n, device=DEV, generator=g) < 0.1
    ms = (torch.rand(h, n, device=DEV, generator=g) < 0.1) & ~mp
    p_lse[mp] = float("inf")
    s_lse[ms] = float("inf")
    p_out = torch.randn(n, h, d, device=DEV, generator=g).to(dtype)
    s_out = torch.randn(n, h, d, device=DEV, generator=g).to(dtype)
    out, out_lse = torch.empty_like(p_out), torch.empty_like(p_lse)
    ref_out, ref_lse = torch.empty_like(p_out), torch.empty_like(p_lse)
    ops.merge_attn_states(out, p_out, p_lse, s_out, s_lse, out_lse)
    ref_cuda.merge_attn_states(ref_out, ref_lse, p_out, p_lse, s_out, s_lse)
    torch.cuda.synchronize()
    assert torch.equal(out.view(torch.uint8), ref_out.view(torch.uint8))
    assert torch.equal(out_lse.view(torch.int32), ref_lse.view(torch.int32))


def test_merge_attn_states_errors(ops):
    o = torch.zeros(4, 2, 20, dtype=torch.bfloat16, device=DEV)      # 20 % 8 != 0
    l = torch.zeros(2, 4, device=DEV)
    with pytest.raises(RuntimeError, match="headsize must be multiple of pack_size"):
        ops.merge_attn_states(o, o, l, o, l)
    o = torch.zeros(4, 2, 16, dtype=torch.float64, device=DEV)
    with pytest.raises(RuntimeError, match="Unsupported data type of O"):
        ops.merge_attn_states(o, o, l, o, l)


# ------------------------------------------------------------------ correct_attn_out
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_correct_attn_out_matches_oracle(ops, lora, dtype):
    N, B, H, D = 4, 9, 12, 128
    g = torch.Generator().manual_seed(0)
    lses = torch.randn(N, B, H, generator=g) * 3
    lses[1, 2, :] = -float("inf")
    lses[0, 3, 1] = float("nan")
    lses[2, 4, 5] = float("inf")
    out = torch.randn(B, H, D, generator=g).to(dtype)
    for rank in range(N):
        want, want_lse = lora.correct_attn_out(out.clone(), lses, rank)
        # a strided [B,H,D] view (the reference passes a slice of a larger buffer)
        buf = torch.zeros(B, H + 3, D, dtype=dtype, device=DEV)
        view = buf[:, 1:H + 1]
        view.copy_(out)
        got, got_lse = ops.correct_attn_out(view, lses.to(DEV), rank)
        assert got.data_ptr() == view.data_ptr()
        torch.testing.assert_close(got.cpu().float(), want.float(), atol=1e-6,
                                   rtol=2 ** -7 if dtype != torch.float32 else 1e-5)
        torch.testing.assert_close(got_lse.cpu(), want_lse, atol=1e-5, rtol=1e-5)
        assert (buf[:, 0] == 0).all() and (buf[:, H + 1:] == 0).all()


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
def test_correct_attn_out_many_ranks_and_empty_rows(ops, lora, dtype):
    """More ranks than the kernel loads at once (8): leading ranks empty, one row empty on every
    rank (NaN LSE and zero factor, as the reference's kernel: common.py:44-64), dense 16-byte path
    and an odd head size that takes the scalar kernel."""
    N, B, H = 19, 5, 6
    g = torch.Generator().manual_seed(1)
    lses = torch.randn(N, B, H, generator=g) * 4
    lses[:9, 1, :] = -float("inf")          # first chunk of ranks holds nothing of sequence 1
    lses[:, 2, 3] = -float("inf")           # nobody holds anything of (2, 3)
    lses[12, 4, 0] = float("nan")
    for D in (128, 72, 20):
        out = torch.randn(B, H, D, generator=g).to(dtype)
        for rank in (0, 10, 18):
            want, want_lse = lora.correct_attn_out(out.clone(), lses, rank)
            got, got_lse = ops.correct_attn_out(out.clone().to(DEV), lses.to(DEV), rank)
            torch.testing.assert_close(got.cpu().float(), want.float(), atol=1e-6, rtol=2 ** -7,
                                       equal_nan=True)
            torch.testing.assert_close(got_lse.cpu(), want_lse, atol=1e-5, rtol=1e-5,
                                       equal_nan=True)
            assert torch.isnan(got_lse[2, 3]) and float(got[2, 3].float().abs().sum()) == 0.0


@pytest.mark.parametrize("kv", ["auto", "fp8"])
def test_decode_context_parallel_property_on_gpu(ops, kv):
    """Two emulated DCP ranks (tokens interleaved by page): partial decode + LSE on each,
    correct_attn_out, sum == decode over the whole context."""
    B, H, Hkv, D, BS = 4, 16, 2, 128, 16
    lens = [16, 1000, 4096, 6000]            # first sequence: rank 1 holds nothing
    NB = sum((l + BS - 1) // BS for l in lens) + 1
    q, kc, vc, bt, sl = hp.make_paged_inputs(B, H, Hkv, D, BS, NB, lens, torch.bfloat16, 21, kv=kv,
                                             permute=True)
    s = D ** -0.5
    ks, vs = (1.0, 1.0) if kv == "auto" else (0.4, 0.3)
    full, full_lse = _gpu_lse(ops, q, kc, vc, Hkv, s, bt, sl, BS, max(lens), None, kv, ks, vs)
    N = 2
    outs, lses = [], []
    for r in range(N):
        # rank r owns pages r, r+N, ... of every sequence (whole pages: no cache rewrite needed)
        rows, lens_r = [], []
        for b, L in enumerate(lens):
            nb = (L + BS - 1) // BS
            mine = list(range(r, nb, N))
            rows.append([int(bt[b, j]) for j in mine])
            lens_r.append(sum(min(BS, L - j * BS) for j in mine))
        # a rank's pages must be full except its last one for seq_len to describe them: true
        # here unless the sequence's last (partial) page is not the rank's last page
        for b, L in enumerate(lens):
            nb = (L + BS - 1) // BS
            if L % BS and (nb - 1) % N != r and rows[b]:
                assert lens_r[b] % BS == 0
        mb = max(1, max(len(x) for x in rows))
        bt_r = torch.tensor([x + [0] * (mb - len(x)) for x in rows], dtype=torch.int32)
        sl_r = torch.tensor(lens_r, dtype=torch.int32)
        o, l = _gpu_lse(ops, q, kc, vc, Hkv, s, bt_r, sl_r, BS, max(max(lens_r), 1), None, kv, ks, vs)
        outs.append(o)
        lses.append(l)
    lses_d = torch.stack(lses).to(DEV)
    total = torch.zeros(B, H, D)
    for r in range(N):
        o = outs[r].float().to(DEV)
        o, lse_tot = ops.correct_attn_out(o, lses_d, r)
        total += o.cpu()
    torch.testing.assert_close(total, full.float(), atol=1e-2, rtol=1e-2)
    torch.testing.assert_close(lse_tot.cpu(), full_lse, atol=2e-3, rtol=1e-4)


def test_opcheck_lse_ops(ops):
    utils = ("test_schema", "test_faketensor")
    n, h, d = 8, 4, 64
    o = torch.randn(n, h, d, dtype=torch.bfloat16, device=DEV)
    l = torch.randn(h, n, device=DEV)
    torch.library.opcheck(torch.ops.b200pa.merge_attn_states,
                          (torch.empty_like(o), torch.empty_like(l), o, l, o.clone(), l.clone()),
                          test_utils=utils)
    torch.library.opcheck(torch.ops.b200pa.merge_attn_states,
                          (torch.empty_like(o), None, o, l, o.clone(), l.clone()), test_utils=utils)
    lses = torch.randn(2, n, h, device=DEV)
    torch.library.opcheck(torch.ops.b200pa.correct_attn_out, (o.clone(), lses, 1),
                          test_utils=("test_schema",))


@pytest.mark.parametrize("name", ["correct_f32_n4.npz", "correct_bf16_n8.npz", "correct_f16_n2.npz"])
def test_correct_attn_out_matches_the_reference_kernel_golden(ops, name):
    """Outputs of the reference's Triton kernel source (vllm/attention/ops/common.py:9-68) run at
    fixture-generation time (tests/golden/make_correct_golden.py), inf / nan / all -inf rows
    included: same values, NaN in the same places."""
    from tests.test_lse_oracle import load_correct_case
    out, lses, rank, want, want_lse, dt = load_correct_case(name)
    got, got_lse = ops.correct_attn_out(out.clone().to(DEV), lses.to(DEV), rank)
    got, got_lse = got.cpu(), got_lse.cpu()
    assert torch.equal(torch.isnan(got_lse), torch.isnan(want_lse))
    fin = ~torch.isnan(want_lse)
    torch.testing.assert_close(got_lse[fin], want_lse[fin], atol=1e-5, rtol=1e-5)
    assert torch.equal(torch.isnan(got.float()), torch.isnan(want.float()))
    ok = ~torch.isnan(want.float())
    torch.testing.assert_close(got.float()[ok], want.float()[ok], atol=1e-6,
                               rtol=2 ** -7 if dt != torch.float32 else 1e-5)

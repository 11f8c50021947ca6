"""The prefill oracle (oracle/prefill_oracle.py) against outputs of the reference's own test
oracle ``ref_paged_attn`` (tests/kernels/attention/test_flash_attn.py:27-79 of the reference),
committed as tests/golden/prefill_*.npz by tests/golden/make_prefill_golden.py."""
import pytest
import torch

from oracle import prefill_oracle as po
from tests import helpers as hp

CASES = ["prefill_gqa_bf16_d128.npz", "prefill_mha_f16_d64_bs32.npz",
         "prefill_gqa8_bf16_d128_spec.npz", "prefill_gqa_f16_d256.npz"]


def load_case(name):
    g = hp.load_golden(name)
    dt = hp.golden_dtype(g)
    return dict(
        q=hp.from_bits(g["q"], dt), kc=hp.from_bits(g["key_cache"], dt),
        vc=hp.from_bits(g["value_cache"], dt), bt=torch.from_numpy(g["block_tables"]),
        qsl=torch.from_numpy(g["query_start_loc"]), sl=torch.from_numpy(g["seq_lens"]),
        scale=float(g["scale"]), want=hp.from_bits(g["out_ref"], dt), dtype=dt)


@pytest.mark.parametrize("name", CASES)
def test_oracle_matches_reference_test_oracle(name):
    c = load_case(name)
    got = po.prefill_attention(c["q"], c["kc"], c["vc"], c["bt"], c["qsl"].tolist(),
                               c["sl"].tolist(), c["scale"], round_p=True)
    # the reference oracle scales q in its dtype and rounds P to it: 16-bit noise only
    tol = 2e-2 if c["dtype"] == torch.bfloat16 else 4e-3
    torch.testing.assert_close(got, c["want"].float(), atol=tol, rtol=tol)


def test_oracle_reduces_to_decode_oracle():
    """q_len = 1 everywhere: the chunk oracle is single-query attention (tests/helpers.py)."""
    B, H, Hkv, D, BS, NB = 3, 4, 2, 64, 16, 12
    lens = [1, 17, 40]
    q, kc, vc, bt, sl = hp.make_paged_inputs(B, H, Hkv, D, BS, NB, lens, torch.float32, seed=3)
    want = hp.torch_reference_attention(q, kc, vc, bt, sl, 0.2, Hkv).float()
    got = po.prefill_attention(q, kc, vc, bt, list(range(B + 1)), lens, 0.2)
    torch.testing.assert_close(got, want, atol=1e-5, rtol=1e-5)


def test_oracle_lse_and_alibi_consistency():
    """Splitting a chunk into two calls (context grows) leaves each token's output unchanged."""
    H, Hkv, D, BS, NB = 4, 2, 32, 16, 8
    q, kc, vc, bt, _ = hp.make_paged_inputs(1, H, Hkv, D, BS, NB, [50], torch.float32, seed=5)
    g = torch.Generator().manual_seed(0)
    qq = torch.randn(10, H, D, generator=g)
    slopes = torch.tensor([0.1, 0.2, 0.3, 0.4])
    whole, lse = po.prefill_attention(qq, kc, vc, bt, [0, 10], [50], 0.3, slopes, return_lse=True)
    first = po.prefill_attention(qq[:6], kc, vc, bt, [0, 6], [46], 0.3, slopes)
    second = po.prefill_attention(qq[6:], kc, vc, bt, [0, 4], [50], 0.3, slopes)
    torch.testing.assert_close(torch.cat([first, second]), whole, atol=1e-6, rtol=1e-6)
    assert torch.isfinite(lse).all()


@pytest.mark.parametrize("kv", ["fp8", "fp8_e5m2"])
def test_oracle_fp8_dequant_agrees_with_the_pinned_decode_oracle(kv):
    """Single-token chunks over an fp8 cache with k_scale / v_scale and ALiBi: the chunk oracle's
    dequantisation and bias must agree with the C restatement of the reference's decode kernel
    (oracle/paged_attn_oracle.c, itself pinned on the CUDA-toolkit fp8 goldens and on the
    reference's CPU kernels)."""
    from oracle import oracle as ora
    B, H, Hkv, D, BS, NB = 4, 8, 2, 64, 16, 40
    lens = [1, 16, 77, 300]
    q, kc, vc, bt, sl = hp.make_paged_inputs(B, H, Hkv, D, BS, NB, lens, torch.bfloat16, seed=9,
                                             kv="fp8", permute=True)
    if kv == "fp8_e5m2":
        kc, vc = kc & 0xBB, vc & 0xBB      # finite e5m2 codes
    slopes = torch.tensor([0.5 ** (i + 1) for i in range(H)])
    ks, vs = 0.25, 0.5
    want = torch.empty_like(q)
    ora.paged_attention_v1(want, q, kc, vc, Hkv, 0.125, bt, sl, BS, max(lens), slopes, kv,
                           torch.tensor(ks), torch.tensor(vs))
    got = po.prefill_attention(q, kc, vc, bt, list(range(B + 1)), lens, 0.125, slopes, kv, ks, vs)
    torch.testing.assert_close(got, want.float(), atol=2e-2, rtol=2e-2)

"""CPU tests of the log-sum-exp oracle (oracle/lse_oracle.py): merge_attn_states against the
reference's own torch implementation (golden fixtures, tests/golden/make_golden.py),
paged_attention_lse against an fp64 restatement, correct_attn_out through the decode-context-
parallel property it exists for (vllm/attention/ops/common.py:119-140 of the reference)."""
import math

import numpy as np
import pytest
import torch

from oracle import lse_oracle, oracle
from tests import helpers as hp


@pytest.mark.parametrize("name", ["merge_bf16_d128.npz", "merge_f16_d96.npz", "merge_f32_d32.npz"])
def test_merge_oracle_matches_reference_torch_impl(name):
    g = hp.load_golden(name)
    dt = hp.golden_dtype(g)
    p_out = hp.from_bits(g["prefix_output"], dt)
    s_out = hp.from_bits(g["suffix_output"], dt)
    p_lse = torch.from_numpy(g["prefix_lse"].copy())
    s_lse = torch.from_numpy(g["suffix_lse"].copy())
    out = torch.empty_like(p_out)
    out_lse = torch.empty_like(p_lse)
    lse_oracle.merge_attn_states(out, p_out, p_lse, s_out, s_lse, out_lse)
    want = torch.from_numpy(g["output_f32"])
    # the golden is unrounded fp32; the oracle rounds once to scalar_t
    torch.testing.assert_close(out.float(), want.to(dt).float(), atol=0, rtol=2 ** -7 if dt != torch.float32 else 1e-6)
    want_lse = torch.from_numpy(g["output_lse"])
    fin = torch.isfinite(want_lse)
    torch.testing.assert_close(out_lse[fin], want_lse[fin], atol=1e-6, rtol=1e-6)
    assert torch.equal(torch.isfinite(out_lse), fin)
    # inputs were not modified (the reference's torch function edits its lse arguments in place)
    assert torch.equal(p_lse, torch.from_numpy(g["prefix_lse"]))


def _fp64_attention_lse(q, kc, vc, bt, sl, scale, Hkv):
    B, H, D = q.shape
    NB, _, Dx, BS, x = kc.shape
    k_log = kc.double().permute(0, 3, 1, 2, 4).reshape(NB, BS, Hkv, D)
    G = H // Hkv
    lse = torch.full((B, H), -math.inf, dtype=torch.float64)
    for b in range(B):
        L = int(sl[b])
        if L == 0:
            continue
        nb = (L + BS - 1) // BS
        k = k_log[bt[b, :nb].long()].reshape(nb * BS, Hkv, D)[:L].repeat_interleave(G, dim=1)
        logits = scale * torch.einsum("hd,lhd->hl", q[b].double(), k)
        lse[b] = torch.logsumexp(logits, dim=-1)
    return lse


def test_paged_attention_lse_oracle_matches_fp64():
    B, H, Hkv, D, BS, NB = 4, 8, 2, 128, 16, 80
    lens = [0, 17, 600, 1200]
    q, kc, vc, bt, sl = hp.make_paged_inputs(B, H, Hkv, D, BS, NB, lens, torch.bfloat16, 5)
    scale = D ** -0.5
    out = torch.empty_like(q)
    lse = torch.empty(B, H)
    one = torch.tensor(1.0)
    lse_oracle.paged_attention_lse(out, lse, q, kc, vc, Hkv, scale, bt, sl, BS, max(lens), None,
                                   "auto", one, one)
    want = _fp64_attention_lse(q, kc, vc, bt, sl, scale, Hkv)
    assert torch.isinf(lse[0]).all() and (lse[0] < 0).all()
    torch.testing.assert_close(lse[1:].double(), want[1:], atol=1e-5, rtol=1e-5)
    ref = hp.torch_reference_attention(q, kc, vc, bt, sl, scale, Hkv)
    torch.testing.assert_close(out.double()[1:], ref[1:], atol=1e-2, rtol=1e-2)


@pytest.mark.parametrize("n_ranks", [2, 4])
def test_correct_attn_out_sums_to_full_attention(n_ranks):
    """Decode context parallelism: rank r holds tokens r, r+N, ... of every sequence.  Partial
    attention + LSE per rank, corrected with the gathered LSEs, summed over ranks == attention over
    the whole context.  One rank has no tokens of the shortest sequence (LSE -inf)."""
    B, H, Hkv, D, BS = 3, 4, 2, 64, 16
    lens = [1, 100, 333]
    NB = 64
    scale = D ** -0.5
    q, kc, vc, bt, sl = hp.make_paged_inputs(B, H, Hkv, D, BS, NB, lens, torch.float32, 7, permute=True)
    full = hp.torch_reference_attention(q, kc, vc, bt, sl, scale, Hkv)
    x = kc.shape[-1]
    k_log = kc.permute(0, 3, 1, 2, 4).reshape(NB, BS, Hkv, D)
    v_log = vc.permute(0, 3, 1, 2)
    outs, lses = [], []
    for r in range(n_ranks):
        # rank-local cache: the rank's tokens of each sequence, packed into fresh pages
        lens_r = [len(range(r, L, n_ranks)) for L in lens]
        mb = max(1, max((l + BS - 1) // BS for l in lens_r))
        kc_r = torch.zeros(B * mb, Hkv, D // x, BS, x)
        vc_r = torch.zeros(B * mb, Hkv, D, BS)
        bt_r = torch.arange(B * mb, dtype=torch.int32).reshape(B, mb)
        for b, L in enumerate(lens):
            for j, t in enumerate(range(r, L, n_ranks)):
                src_blk, src_off = int(bt[b, t // BS]), t % BS
                dst_blk, dst_off = int(bt_r[b, j // BS]), j % BS
                kc_r[dst_blk, :, :, dst_off, :] = k_log[src_blk, src_off].reshape(Hkv, D // x, x)
                vc_r[dst_blk, :, :, dst_off] = v_log[src_blk, src_off]
        sl_r = torch.tensor(lens_r, dtype=torch.int32)
        o = torch.zeros(B, H, D)
        l = torch.empty(B, H)
        one = torch.tensor(1.0)
        lse_oracle.paged_attention_lse(o, l, q, kc_r, vc_r, Hkv, scale, bt_r, sl_r, BS,
                                       max(max(lens_r), 1), None, "auto", one, one)
        outs.append(o)
        lses.append(l)
    lses = torch.stack(lses)
    total = torch.zeros(B, H, D)
    lse_tot = None
    for r in range(n_ranks):
        o, lse_tot = lse_oracle.correct_attn_out(outs[r].clone(), lses, r)
        total += o
    torch.testing.assert_close(total.double(), full, atol=1e-4, rtol=1e-4)
    want_lse = _fp64_attention_lse(q, kc, vc, bt, sl, scale, Hkv)
    torch.testing.assert_close(lse_tot.double(), want_lse, atol=1e-5, rtol=1e-5)


CORRECT_CASES = ["correct_f32_n4.npz", "correct_bf16_n8.npz", "correct_f16_n2.npz"]


def load_correct_case(name):
    g = hp.load_golden(name)
    dt = hp.golden_dtype(g)
    return (hp.from_bits(g["out"], dt), torch.from_numpy(g["lses"].copy()), int(g["rank"]),
            hp.from_bits(g["out_ref"], dt), torch.from_numpy(g["lse_ref"].copy()), dt)


@pytest.mark.parametrize("name", CORRECT_CASES)
def test_correct_attn_out_oracle_matches_the_reference_kernel(name):
    """tests/golden/correct_*.npz: the reference's Triton kernel source
    (vllm/attention/ops/common.py:9-68) executed per (batch, head) at fixture-generation time
    (tests/golden/make_correct_golden.py), inf / nan / all -inf rows included."""
    out, lses, rank, want, want_lse, dt = load_correct_case(name)
    got, got_lse = lse_oracle.correct_attn_out(out.clone(), lses, rank)
    assert torch.equal(torch.isnan(got_lse), torch.isnan(want_lse))
    fin = ~torch.isnan(want_lse)
    torch.testing.assert_close(got_lse[fin], want_lse[fin], atol=1e-6, rtol=1e-6)
    assert torch.equal(torch.isnan(got.float()), torch.isnan(want.float()))
    ok = ~torch.isnan(want.float())
    torch.testing.assert_close(got.float()[ok], want.float()[ok], atol=1e-7,
                               rtol=2 ** -8 if dt != torch.float32 else 1e-6)

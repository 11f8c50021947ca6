"""Prefill / chunked-prefill / mixed batches through ``AttentionImpl.forward``:
the backend writes every new token, runs the leading single-token sequences of a decode-first
batch on the decode kernel and the chunks on the causal chunk kernel (``ops.prefill_attention``);
the token-major backend still expands per token (``backend.expand_queries_to_sequences``).  Checked against an independent fp64 causal
attention over the logical token order (no paging, no expansion): new K/V are appended to each
sequence's history and token j of the query attends to history + new[:j+1].
"""
import pytest
import torch

from tests import helpers as hp

pytestmark = pytest.mark.gpu

DEV = "cuda:0"
BS = 16


class _Layer:
    def __init__(self):
        self._k_scale = torch.tensor(1.0, device=DEV)
        self._v_scale = torch.tensor(1.0, device=DEV)
        self._q_scale = torch.tensor(1.0, device=DEV)


def _causal_reference(q, hist_k, hist_v, new_k, new_v, qlens, scale, G):
    """q [T,H,D]; per sequence: hist_k/v [Lh,Hkv,D] (tokens already cached), new_k/v [ql,Hkv,D]."""
    out = torch.zeros(q.shape, dtype=torch.float64)
    t0 = 0
    for b, ql in enumerate(qlens):
        k_all = torch.cat([hist_k[b], new_k[t0:t0 + ql]]).double().repeat_interleave(G, dim=1)
        v_all = torch.cat([hist_v[b], new_v[t0:t0 + ql]]).double().repeat_interleave(G, dim=1)
        Lh = hist_k[b].shape[0]
        for j in range(ql):
            L = Lh + j + 1
            logits = scale * torch.einsum("hd,lhd->hl", q[t0 + j].double(), k_all[:L])
            p = torch.softmax(logits, dim=-1)
            out[t0 + j] = torch.einsum("hl,lhd->hd", p, v_all[:L])
        t0 += ql
    return out


@pytest.mark.parametrize("which", ["paged", "flash"])
def test_mixed_prefill_and_decode_batch(which):
    from oracle import cache_ops_oracle as co
    from vllm_main_modify_b200 import backend as be
    H, Hkv, D = 16, 4, 128
    G = H // Hkv
    dt = torch.bfloat16
    s = D ** -0.5
    hist = [0, 37, 500, 16, 0]          # tokens already in the cache
    qlens = [45, 1, 20, 1, 3]           # prefill, decode, chunked prefill, decode, short prefill
    B, T = len(hist), sum(qlens)
    lens = [h + ql for h, ql in zip(hist, qlens)]
    NB = sum((L + BS - 1) // BS for L in lens) + 3
    g = torch.Generator().manual_seed(5)
    # x-packed cache with random history (make_paged_inputs gives a non-aliased permuted table)
    _, kc, vc, bt, sl = hp.make_paged_inputs(B, H, Hkv, D, BS, NB, lens, dt, seed=31, kv="auto",
                                             lo=-1.0, hi=1.0, permute=True)
    q = (torch.rand(T, H, D, generator=g) * 2 - 1).to(dt)
    new_k = (torch.rand(T, Hkv, D, generator=g) * 2 - 1).to(dt)
    new_v = (torch.rand(T, Hkv, D, generator=g) * 2 - 1).to(dt)
    # logical history per sequence, read back from the paged cache
    kf, vf = co.xpacked_to_flash(kc, vc)                     # [NB,BS,Hkv,D]
    hist_k, hist_v = [], []
    for b in range(B):
        nb = (hist[b] + BS - 1) // BS
        blocks = bt[b, :nb].long()
        hist_k.append(kf[blocks].reshape(nb * BS, Hkv, D)[:hist[b]])
        hist_v.append(vf[blocks].reshape(nb * BS, Hkv, D)[:hist[b]])
    want = _causal_reference(q, hist_k, hist_v, new_k, new_v, qlens, s, G)
    # metadata as the runner builds it
    slots, qsl = [], [0]
    for b in range(B):
        for j in range(qlens[b]):
            p = hist[b] + j
            slots.append(int(bt[b, p // BS]) * BS + p % BS)
        qsl.append(qsl[-1] + qlens[b])
    meta = be.B200PagedAttentionMetadata(
        num_actual_tokens=T, max_query_len=max(qlens), seq_lens=sl.to(DEV), max_seq_len=max(lens),
        block_table=bt.to(DEV), slot_mapping=torch.tensor(slots, dtype=torch.int64, device=DEV),
        query_start_loc=torch.tensor(qsl, dtype=torch.int32, device=DEV))
    if which == "paged":
        backend = be.B200PagedAttentionBackend
        kv_cache = torch.empty(backend.get_kv_cache_shape(NB, BS, Hkv, D), dtype=dt, device=DEV)
        kv_cache[0] = kc.reshape(NB, -1).to(DEV)
        kv_cache[1] = vc.reshape(NB, -1).to(DEV)
    else:
        backend = be.B200FlashLayoutAttentionBackend
        kv_cache = torch.empty(backend.get_kv_cache_shape(NB, BS, Hkv, D), dtype=dt, device=DEV)
        kv_cache[0] = kf.to(DEV)
        kv_cache[1] = vf.to(DEV)
    impl = backend.get_impl_cls()(H, D, s, Hkv, None, None, "auto")
    out = impl.forward(_Layer(), q.to(DEV).view(T, H * D), new_k.to(DEV).view(T, Hkv * D),
                       new_v.to(DEV).view(T, Hkv * D), kv_cache, meta)
    torch.cuda.synchronize()
    got = out.view(T, H, D).cpu().float()
    assert not torch.isnan(got).any()
    torch.testing.assert_close(got, want.float(), atol=1e-2, rtol=1e-2)
    # every new token landed in its slot
    kd = kv_cache[0].cpu()
    if which == "paged":
        kd = co.xpacked_to_flash(kd.view(kc.shape), kv_cache[1].cpu().view(vc.shape))[0]
    flat = kd.reshape(NB * BS, Hkv, D)
    assert torch.equal(flat[torch.tensor(slots)], new_k)


class _Common:
    pass


@pytest.mark.parametrize("qlens,split", [([1, 1, 1, 70, 20, 3], (3, 3)),
                                         # a sequence without tokens inside the single-token run:
                                         # the decode run ends before it
                                         ([1, 0, 1, 70, 20, 3], (1, 5))])
@pytest.mark.parametrize("kv", ["auto", "fp8"])
def test_decode_first_batch_through_the_builder(kv, qlens, split):
    """A batch as the runner hands it over after ``reorder_batch``: decodes first, then chunks;
    ``build`` records the split, ``forward`` runs the two kernels.  Checked against the chunk
    oracle on the cache as the backend itself wrote it (bit-exact bytes are the cache tests'
    business) and, for the 16-bit cache, against the independent fp64 reference above."""
    from oracle import oracle as ora
    from oracle import prefill_oracle as po
    from oracle import cache_ops_oracle as co
    from vllm_main_modify_b200 import backend as be
    H, Hkv, D = 16, 2, 128
    G = H // Hkv
    dt = torch.bfloat16
    s = D ** -0.5
    hist = [300, 37, 1000, 0, 500, 16]
    B, T = len(hist), sum(qlens)
    lens = [h + ql for h, ql in zip(hist, qlens)]
    NB = sum((L + BS - 1) // BS for L in lens) + 3
    g = torch.Generator().manual_seed(7)
    _, kc, vc, bt, sl = hp.make_paged_inputs(B, H, Hkv, D, BS, NB, lens, dt, seed=32, kv=kv,
                                             lo=-1.0, hi=1.0, permute=True)
    q = (torch.rand(T, H, D, generator=g) * 2 - 1).to(dt)
    new_k = (torch.rand(T, Hkv, D, generator=g) * 2 - 1).to(dt)
    new_v = (torch.rand(T, Hkv, D, generator=g) * 2 - 1).to(dt)
    slots, qsl = [], [0]
    for b in range(B):
        for j in range(qlens[b]):
            p = hist[b] + j
            slots.append(int(bt[b, p // BS]) * BS + p % BS)
        qsl.append(qsl[-1] + qlens[b])
    ks, vs = (1.0, 1.0) if kv == "auto" else (0.02, 0.03)
    c = _Common()
    c.num_actual_tokens, c.max_query_len, c.num_reqs = T, max(qlens), B
    c.seq_lens, c.seq_lens_cpu, c.max_seq_len = sl.to(DEV), sl, max(lens)
    c.block_table_tensor = bt.to(DEV)
    c.slot_mapping = torch.tensor(slots, dtype=torch.int64, device=DEV)
    c.query_start_loc_cpu = torch.tensor(qsl, dtype=torch.int32)
    c.query_start_loc = c.query_start_loc_cpu.to(DEV)
    meta = be.B200PagedAttentionMetadataBuilder().build(0, c)
    assert (meta.num_decodes, meta.num_prefills) == split
    backend = be.B200PagedAttentionBackend
    kv_cache = torch.empty(backend.get_kv_cache_shape(NB, BS, Hkv, D), dtype=kc.dtype, device=DEV)
    kv_cache[0] = kc.reshape(NB, -1).to(DEV)
    kv_cache[1] = vc.reshape(NB, -1).to(DEV)
    layer = _Layer()
    layer._k_scale = torch.tensor(ks, device=DEV)
    layer._v_scale = torch.tensor(vs, device=DEV)
    impl = backend.get_impl_cls()(H, D, s, Hkv, None, None, kv)
    out = impl.forward(layer, q.to(DEV).view(T, H * D), new_k.to(DEV).view(T, Hkv * D),
                       new_v.to(DEV).view(T, Hkv * D), kv_cache, meta)
    torch.cuda.synchronize()
    got = out.view(T, H, D).cpu().float()
    # the cache as the oracle's reshape_and_cache writes it == what the backend wrote
    kc_o, vc_o = kc.clone(), vc.clone()
    ora.reshape_and_cache(new_k, new_v, kc_o, vc_o, torch.tensor(slots), kv, torch.tensor(ks),
                          torch.tensor(vs))
    assert torch.equal(kv_cache[0].cpu().view(kc.shape).view(torch.uint8), kc_o.view(torch.uint8))
    assert torch.equal(kv_cache[1].cpu().view(vc.shape).view(torch.uint8), vc_o.view(torch.uint8))
    want = po.prefill_attention(q, kc_o, vc_o, bt, qsl, lens, s, None, kv, ks, vs)
    torch.testing.assert_close(got, want, atol=1e-2, rtol=1e-2)
    if kv == "auto":
        kf, vf = co.xpacked_to_flash(kc, vc)
        hist_k, hist_v = [], []
        for b in range(B):
            nb = (hist[b] + BS - 1) // BS
            blocks = bt[b, :nb].long()
            hist_k.append(kf[blocks].reshape(nb * BS, Hkv, D)[:hist[b]])
            hist_v.append(vf[blocks].reshape(nb * BS, Hkv, D)[:hist[b]])
        ref = _causal_reference(q, hist_k, hist_v, new_k, new_v, qlens, s, G)
        torch.testing.assert_close(got, ref.float(), atol=1e-2, rtol=1e-2)

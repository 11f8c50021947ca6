"""Host-side logic of the backend that needs no GPU: the expansion of multi-query batches into
single-query sequences (causal context lengths) and the backend twins' static contract."""
import pytest
import torch

from vllm_main_modify_b200 import backend as be


def _meta(seq_lens, qlens, width=4):
    qsl = [0]
    for q in qlens:
        qsl.append(qsl[-1] + q)
    B = len(seq_lens)
    return be.B200PagedAttentionMetadata(
        num_actual_tokens=qsl[-1], max_query_len=max(qlens),
        seq_lens=torch.tensor(seq_lens, dtype=torch.int32), max_seq_len=max(seq_lens),
        block_table=torch.arange(B * width, dtype=torch.int32).view(B, width),
        slot_mapping=torch.zeros(qsl[-1], dtype=torch.int64),
        query_start_loc=torch.tensor(qsl, dtype=torch.int32))


def test_expand_queries_to_sequences_causal_lengths():
    seq_lens, qlens = [10, 1, 7, 64], [4, 1, 7, 2]
    bt, sl = be.expand_queries_to_sequences(_meta(seq_lens, qlens))
    want_sl, want_rows = [], []
    for b, (L, q) in enumerate(zip(seq_lens, qlens)):
        for j in range(q):
            want_sl.append(L - q + j + 1)     # history + the query tokens up to and including j
            want_rows.append(b)
    assert sl.dtype == torch.int32 and sl.tolist() == want_sl
    assert bt.shape == (sum(qlens), 4) and bt.is_contiguous()
    assert torch.equal(bt, torch.arange(16, dtype=torch.int32).view(4, 4)[want_rows])


def test_expand_pure_decode_is_identity():
    m = _meta([5, 9, 1], [1, 1, 1])
    bt, sl = be.expand_queries_to_sequences(m)
    assert torch.equal(bt, m.block_table) and torch.equal(sl, m.seq_lens)


def test_expand_needs_query_start_loc():
    m = _meta([5, 9], [2, 1])
    m.query_start_loc = None
    with pytest.raises(ValueError, match="query_start_loc"):
        be.expand_queries_to_sequences(m)


def test_backend_twins_static_contract():
    """``abstract.py:42-82`` / ``selector.py:112-131`` of the reference: names, classes, shapes."""
    paged, flash = be.B200PagedAttentionBackend, be.B200FlashLayoutAttentionBackend
    assert paged.get_name() == "B200_PAGED_ATTN"
    assert flash.get_name() == "B200_PAGED_ATTN_FLASH_LAYOUT"
    assert paged.get_kv_cache_shape(10, 16, 8, 128) == (2, 10, 16 * 8 * 128)   # paged_attn.py:48-54
    assert flash.get_kv_cache_shape(10, 16, 8, 128) == (2, 10, 16, 8, 128)
    with pytest.raises(ValueError):
        flash.get_kv_cache_shape(10, 32, 8, 128)
    assert flash.get_impl_cls() is be.B200FlashLayoutAttentionImpl
    assert issubclass(flash.get_impl_cls(), paged.get_impl_cls())
    assert paged.get_builder_cls() is flash.get_builder_cls()
    for cls in (paged, flash):
        assert cls.accept_output_buffer and not cls.use_cascade_attention()
        assert torch.bfloat16 in cls.get_supported_dtypes()
        cls.validate_head_size(128)
    with pytest.raises(ValueError, match="Head size 80"):
        flash.validate_head_size(80)
    mod, name = flash.full_cls_name()
    assert (mod, name) == ("vllm_main_modify_b200.backend", "B200FlashLayoutAttentionBackend")


def test_impl_constructor_rejects_what_the_ops_cannot_do():
    Impl = be.B200PagedAttentionImpl
    Impl(32, 128, 0.1, 8, None, None, "fp8")
    with pytest.raises(NotImplementedError):
        Impl(32, 128, 0.1, 8, None, 4096, "auto")               # sliding window
    with pytest.raises(NotImplementedError):
        Impl(32, 128, 0.1, 8, None, None, "auto", 30.0)         # logits soft cap
    with pytest.raises(NotImplementedError):
        Impl(32, 128, 0.1, 8, None, None, "auto", None, be.AttentionType.ENCODER)

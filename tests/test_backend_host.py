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
    assert flash.get_kv_cache_shape(10, 32, 8, 128) == (2, 10, 32, 8, 128)
    with pytest.raises(ValueError):
        flash.get_kv_cache_shape(10, 8, 8, 128)
    assert flash.get_impl_cls() is be.B200FlashLayoutAttentionImpl
    assert issubclass(flash.get_impl_cls(), paged.get_impl_cls())
    assert paged.get_builder_cls() is flash.get_builder_cls()
    for cls in (paged, flash):
        assert cls.accept_output_buffer and not cls.use_cascade_attention()
        assert torch.bfloat16 in cls.get_supported_dtypes()
        cls.validate_head_size(128)
    with pytest.raises(ValueError, match="Head size 120"):
        flash.validate_head_size(120)
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


# ---------------------------------------------------------------------------------------------
# decode-first reordering and the decode / chunk split
# ---------------------------------------------------------------------------------------------
class _Batch:
    """Minimal InputBatch: req_ids in slot order, swap_states as the reference's
    (vllm/v1/worker/gpu_input_batch.py)."""

    def __init__(self, req_ids):
        self.req_ids = list(req_ids)
        self.swaps = []

    def swap_states(self, i, j):
        self.req_ids[i], self.req_ids[j] = self.req_ids[j], self.req_ids[i]
        self.swaps.append((i, j))


class _Sched:
    def __init__(self, n):
        self.num_scheduled_tokens = n


def _reference_reorder(req_ids, sched, threshold=1):
    """The reference's algorithm, restated from v1/attention/backends/utils.py:792-834, as the
    expected result."""
    ids = list(req_ids)
    decodes = [i for i, r in enumerate(ids) if sched[r] <= threshold]
    prefills = [i for i, r in enumerate(ids) if sched[r] > threshold]
    nd = len(decodes)
    for i in range(1, min(nd, len(prefills)) + 1):
        d = decodes[nd - i]
        if d < nd:
            break
        ids[prefills[i - 1]], ids[d] = ids[d], ids[prefills[i - 1]]
    return ids


@pytest.mark.parametrize("tokens", [
    [1, 1, 1], [7, 9], [5, 1, 1, 9, 1], [1, 5, 1, 9, 1, 1], [9, 1], [1, 9], [3, 1, 4, 1, 5, 1, 2, 1]])
def test_reorder_batch_is_decode_first_with_the_reference_swaps(tokens):
    ids = [f"r{i}" for i in range(len(tokens))]
    sched = dict(zip(ids, tokens))
    b = _Batch(ids)
    builder = be.B200PagedAttentionMetadataBuilder()
    moved = builder.reorder_batch(b, _Sched(sched))
    assert b.req_ids == _reference_reorder(ids, sched)
    assert moved == (b.req_ids != ids)
    nd = sum(t == 1 for t in tokens)
    assert all(sched[r] == 1 for r in b.req_ids[:nd]) and all(sched[r] > 1 for r in b.req_ids[nd:])


class _Common:
    pass


def test_build_records_the_split():
    qlens, seq_lens = [1, 1, 5, 64], [10, 300, 5, 100]
    c = _Common()
    qsl = torch.tensor([0, 1, 2, 7, 71], dtype=torch.int32)
    c.num_actual_tokens, c.max_query_len, c.num_reqs = 71, 64, 4
    c.seq_lens = torch.tensor(seq_lens, dtype=torch.int32)
    c.seq_lens_cpu = c.seq_lens
    c.max_seq_len = 300
    c.block_table_tensor = torch.zeros(4, 20, dtype=torch.int32)
    c.slot_mapping = torch.zeros(71, dtype=torch.int64)
    c.query_start_loc = qsl
    c.query_start_loc_cpu = qsl
    m = be.B200PagedAttentionMetadataBuilder().build(0, c)
    assert (m.num_decodes, m.num_prefills, m.num_prefill_tokens) == (2, 2, 69)
    assert m.num_decode_tokens == 2 and m.max_decode_seq_len == 300
    # pure decode batch: everything is a decode, nothing to split
    c.max_query_len, c.num_actual_tokens = 1, 4
    c.query_start_loc = c.query_start_loc_cpu = torch.arange(5, dtype=torch.int32)
    m = be.B200PagedAttentionMetadataBuilder().build(0, c)
    assert m.num_decode_tokens == 4 and m.num_decodes is None


def test_build_ends_the_decode_run_at_a_zero_length_sequence():
    """Sequence b of the decode call is token b, so a sequence without tokens cannot sit inside
    the decode run: the run ends there and the chunk kernel takes the rest."""
    c = _Common()
    qsl = torch.tensor([0, 1, 1, 2, 7], dtype=torch.int32)        # q_len 1, 0, 1, 5
    c.num_actual_tokens, c.max_query_len, c.num_reqs = 7, 5, 4
    c.seq_lens = torch.tensor([10, 300, 20, 5], dtype=torch.int32)
    c.seq_lens_cpu = c.seq_lens
    c.max_seq_len = 300
    c.block_table_tensor = torch.zeros(4, 20, dtype=torch.int32)
    c.slot_mapping = torch.zeros(7, dtype=torch.int64)
    c.query_start_loc = c.query_start_loc_cpu = qsl
    m = be.B200PagedAttentionMetadataBuilder().build(0, c)
    assert (m.num_decodes, m.num_prefills, m.num_prefill_tokens) == (1, 3, 6)
    assert m.max_decode_seq_len == 10
    assert int(qsl[m.num_decodes]) == m.num_decodes               # token b <-> sequence b holds


def _reference_abstract_module():
    """The reference's vllm/attention/backends/abstract.py loaded on its own (its one vllm import,
    QuantKey, is a type annotation): only where /root/reference exists."""
    import importlib.util
    import os
    import sys
    import types
    path = "/root/reference/vllm/attention/backends/abstract.py"
    if not os.path.exists(path):
        pytest.skip("the reference tree is not on this machine")
    stub = "vllm.model_executor.layers.quantization.utils.quant_utils"
    added = []
    parts = stub.split(".")
    for i in range(1, len(parts) + 1):
        name = ".".join(parts[:i])
        if name not in sys.modules:
            sys.modules[name] = types.ModuleType(name)
            added.append(name)
    sys.modules[stub].QuantKey = object
    try:
        spec = importlib.util.spec_from_file_location("_ref_abstract", path)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
    finally:
        for name in added:
            sys.modules.pop(name, None)
    return mod


@pytest.mark.parametrize("flash", [False, True])
def test_bound_classes_are_real_subclasses_and_construct_like_the_attention_layer(flash):
    """Attention.__init__ of the reference (vllm/attention/layer.py:184-198):
    impl_cls = attn_backend.get_impl_cls(); impl_cls(num_heads, head_size, scale, num_kv_heads,
    alibi_slopes, sliding_window, kv_cache_dtype, logits_soft_cap, attn_type,
    kv_sharing_target_layer_name, **extra_impl_args)."""
    ab = _reference_abstract_module()
    Backend, Impl = be.bind_abcs(ab, flash_layout=flash)
    assert issubclass(Backend, ab.AttentionBackend) and issubclass(Impl, ab.AttentionImpl)
    assert not getattr(Backend, "__abstractmethods__", None)
    assert not getattr(Impl, "__abstractmethods__", None)
    impl_cls = Backend.get_impl_cls()
    assert impl_cls is Impl
    impl = impl_cls(32, 128, 0.088, 8, None, None, "fp8", None, ab.AttentionType.DECODER, None)
    assert isinstance(impl, ab.AttentionImpl) and impl.num_queries_per_kv == 4
    assert Backend.get_name().startswith("B200_PAGED_ATTN")
    assert Backend.get_builder_cls() is be.B200PagedAttentionMetadataBuilder
    assert len(Backend.get_kv_cache_stride_order()) == len(Backend.get_kv_cache_shape(4, 16, 8, 128))
    assert Backend.get_metadata_cls() is be.B200PagedAttentionMetadata

"""CPU-side checks of the drop-in boundary: libb200pa.so loads without a GPU and exports every
symbol include/b200_paged_attn.h declares; the Python op mirror keeps the reference's
signatures; nothing in the product imports the oracle; no compute call is made here."""
import ctypes
import inspect
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "b200_paged_attn.h")


TESTING_HEADER = os.path.join(ROOT, "include", "b200_paged_attn_testing.h")


def _declared_symbols(header=HEADER):
    src = open(header).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b200pa_[a-z0-9_]+)\s*\(", src)))


def test_library_loads_and_exports_every_declared_symbol():
    from vllm_main_modify_b200 import _native
    lib = _native.c_lib()
    declared = _declared_symbols()
    assert len(declared) >= 12
    for sym in declared:
        assert hasattr(lib, sym), f"{sym} is declared in the header but not exported"
    # and the loader's own list is the header's list
    assert sorted(_native.C_ABI_SYMBOLS) == declared
    m = re.search(r"#define B200PA_ABI_VERSION (\d+)", open(HEADER).read())
    assert lib.b200pa_abi_version() == int(m.group(1))
    # the test hooks live in their own header, outside the product ABI
    hooks = _declared_symbols(TESTING_HEADER)
    assert sorted(_native.TESTING_SYMBOLS) == hooks
    assert not set(hooks) & set(declared)
    if _native.has_testing_hooks():
        for sym in hooks:
            assert hasattr(lib, sym), sym


def test_host_only_entry_points_work_without_a_gpu():
    from vllm_main_modify_b200 import _native
    lib = _native.c_lib()
    f = lib.b200pa_kv_dtype_from_string
    assert f(b"auto") == 0 and f(b"fp8") == 1 and f(b"fp8_e4m3") == 1 and f(b"fp8_e5m2") == 2
    assert f(b"int8") == -1
    lib.b200pa_paged_attention_workspace_bytes.restype = ctypes.c_int64
    assert lib.b200pa_paged_attention_workspace_bytes(128) > 0
    assert _native.get_impl() == _native.IMPL_AUTO
    assert isinstance(_native.launch_count(), int)


def test_header_cites_the_reference_interface_for_every_launcher():
    src = open(HEADER).read()
    for ref in ("paged_attention_v1.cu:160-182", "paged_attention_v2.cu:167-192",
                "cache_kernels.cu:416-442", "cache_kernels.cu:459-497", "cache_kernels.cu:573-623",
                "torch_bindings.cpp"):
        assert ref in src, ref


def test_python_op_mirror_keeps_reference_signatures():
    """vllm/_custom_ops.py:39-59, 68-91, 1590-1614, 1650-1654 of the reference."""
    src = open(os.path.join(ROOT, "vllm_main_modify_b200", "_custom_ops.py")).read()
    import ast
    fns = {n.name: n for n in ast.walk(ast.parse(src)) if isinstance(n, ast.FunctionDef)}
    v1 = [a.arg for a in fns["paged_attention_v1"].args.args]
    assert v1 == ["out", "query", "key_cache", "value_cache", "num_kv_heads", "scale",
                  "block_tables", "seq_lens", "block_size", "max_seq_len", "alibi_slopes",
                  "kv_cache_dtype", "k_scale", "v_scale", "tp_rank", "blocksparse_local_blocks",
                  "blocksparse_vert_stride", "blocksparse_block_size",
                  "blocksparse_head_sliding_step"]
    v2 = [a.arg for a in fns["paged_attention_v2"].args.args]
    assert v2 == ["out", "exp_sum", "max_logits", "tmp_out"] + v1[1:] or \
        v2 == ["out", "exp_sums", "max_logits", "tmp_out"] + v1[1:]
    defaults = [ast.literal_eval(d) for d in fns["paged_attention_v1"].args.defaults]
    assert defaults == [0, 0, 0, 64, 0]
    for name in ("reshape_and_cache", "reshape_and_cache_flash"):
        assert [a.arg for a in fns[name].args.args] == [
            "key", "value", "key_cache", "value_cache", "slot_mapping", "kv_cache_dtype",
            "k_scale", "v_scale"]
    assert [a.arg for a in fns["convert_fp8"].args.args][:2] == ["output", "input"]


def test_product_never_touches_the_oracle():
    """oracle/ is test infrastructure: only tests/, __graft_entry__.smoke()/build() and
    bench.py's CPU legs may import it."""
    pkg = os.path.join(ROOT, "vllm_main_modify_b200")
    for dirpath, _, files in os.walk(pkg):
        if "_build" in dirpath:
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                txt = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", txt, flags=re.M), f
                assert "paged_attn_oracle" not in txt, f


def test_backend_mirrors_the_attention_backend_contract():
    from vllm_main_modify_b200.backend import (B200PagedAttentionBackend as B,
                                               B200PagedAttentionImpl as Impl)
    assert B.get_kv_cache_shape(10, 16, 8, 128) == (2, 10, 16 * 8 * 128)
    assert B.get_impl_cls() is Impl and B.accept_output_buffer
    sig = inspect.signature(Impl.forward)
    assert list(sig.parameters)[:10] == ["self", "layer", "query", "key", "value", "kv_cache",
                                         "attn_metadata", "output", "output_scale",
                                         "output_block_scale"]
    init = list(inspect.signature(Impl.__init__).parameters)
    assert init[1:11] == ["num_heads", "head_size", "scale", "num_kv_heads", "alibi_slopes",
                          "sliding_window", "kv_cache_dtype", "logits_soft_cap", "attn_type",
                          "kv_sharing_target_layer_name"]
    with pytest.raises(ValueError):
        B.validate_head_size(48)
    with pytest.raises(NotImplementedError):
        Impl(8, 128, 0.1, 8, None, 4096, "auto")       # sliding window

"""GPU parity of the block-management cache ops: copy_blocks, swap_blocks,
gather_and_maybe_dequant_cache, cp_gather_cache -- BIT-EXACT against oracle/cache_ops_oracle.py
and, where oracle/_ref/libvllm_ref_cuda.so is built, against the reference's own CUDA ops on the
same device tensors.  Shapes follow tests/kernels/attention/test_cache.py of the reference
(:45-120 copy_blocks, :368-446 swap_blocks, :510-700 the MLA-style gathers: entry 512 + 64,
block 16)."""
import os
import random

import pytest
import torch

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


@pytest.fixture(scope="module")
def ops():
    from vllm_main_modify_b200 import _custom_ops
    return _custom_ops


@pytest.fixture(scope="module")
def co():
    from oracle import cache_ops_oracle
    return cache_ops_oracle


@pytest.fixture(scope="module")
def ref():
    from oracle import ref_cuda
    if not os.path.exists(ref_cuda.LIB_PATH):
        return None
    ref_cuda.load()
    return ref_cuda if hasattr(torch.ops.vllm_ref_cuda, "cp_gather_cache") else None


def _bytes_equal(a, b):
    return torch.equal(a.contiguous().view(torch.uint8), b.contiguous().view(torch.uint8))


def _equal_up_to_nan_payload(got, want):
    """Bytes identical wherever the value is not NaN; NaN exactly where the oracle has NaN.  The
    x86 oracle and the GPU propagate NaN payloads differently (fp8_e5m2 codes 0x7d-0x7f are NaN);
    against the reference's CUDA op the comparison stays strictly bytewise."""
    if not got.is_floating_point():
        return _bytes_equal(got, want)
    gn, wn = got.isnan(), want.isnan()
    if not torch.equal(gn, wn):
        return False
    w = torch.tensor([], dtype=got.dtype).element_size()
    it = {2: torch.int16, 4: torch.int32}[w]
    return torch.equal(got.contiguous().view(it)[~gn], want.contiguous().view(it)[~wn])


def _cache(shape, dtype, seed):
    g = torch.Generator().manual_seed(seed)
    if dtype == torch.uint8:   # fp8 codes without NaN patterns
        t = torch.randint(0, 0x7e, shape, generator=g, dtype=torch.int32)
        t |= torch.randint(0, 2, shape, generator=g, dtype=torch.int32) << 7
        return t.to(torch.uint8)
    return torch.randn(shape, generator=g).to(dtype)


# ------------------------------------------------------------------ copy_blocks
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32, torch.uint8])
@pytest.mark.parametrize("head_size,block_size", [(64, 16), (80, 8), (256, 32)])
@pytest.mark.parametrize("num_layers", [1, 3])
def test_copy_blocks(ops, co, ref, dtype, head_size, block_size, num_layers):
    NB, Hkv, num_mappings = 200, 8, 64
    x = 16 // torch.tensor([], dtype=dtype).element_size()
    rnd = random.Random(0)
    # the reference test's mapping: distinct sources, each copied to two fresh destinations
    src_blocks = rnd.sample(range(NB), num_mappings)
    remaining = [b for b in range(NB) if b not in set(src_blocks)]
    dst_blocks = rnd.sample(remaining, 2 * num_mappings)
    pairs = []
    for i, s in enumerate(src_blocks):
        pairs += [(s, dst_blocks[2 * i]), (s, dst_blocks[2 * i + 1])]
    mapping = torch.tensor(pairs, dtype=torch.int64)
    kcs = [_cache((NB, Hkv, head_size // x, block_size, x), dtype, 10 + l) for l in range(num_layers)]
    vcs = [_cache((NB, Hkv, head_size, block_size), dtype, 20 + l) for l in range(num_layers)]
    want_k, want_v = [k.clone() for k in kcs], [v.clone() for v in vcs]
    co.copy_blocks(want_k, want_v, mapping)
    dk, dv = [k.to(DEV) for k in kcs], [v.to(DEV) for v in vcs]
    ops.copy_blocks(dk, dv, mapping.to(DEV))
    torch.cuda.synchronize()
    for l in range(num_layers):
        assert _bytes_equal(dk[l].cpu(), want_k[l]) and _bytes_equal(dv[l].cpu(), want_v[l])
    if ref is not None:
        rk, rv = [k.to(DEV) for k in kcs], [v.to(DEV) for v in vcs]
        ref.copy_blocks(rk, rv, mapping.to(DEV))
        torch.cuda.synchronize()
        for l in range(num_layers):
            assert _bytes_equal(dk[l], rk[l]) and _bytes_equal(dv[l], rv[l])


def test_copy_blocks_many_layers_and_graph_capture(ops, co):
    """More layers than one launch carries as parameters (128), and capture: the op neither
    allocates nor synchronises (the reference's does both, cache_kernels.cu:148-153)."""
    L, NB = 150, 12
    kcs = [_cache((NB, 2, 4, 16, 8), torch.bfloat16, l).to(DEV) for l in range(L)]
    vcs = [_cache((NB, 2, 32, 16), torch.bfloat16, 500 + l).to(DEV) for l in range(L)]
    want_k, want_v = [k.cpu().clone() for k in kcs], [v.cpu().clone() for v in vcs]
    mapping = torch.tensor([[0, 7], [3, 11], [3, 9]], dtype=torch.int64)
    co.copy_blocks(want_k, want_v, mapping)
    md = mapping.to(DEV)
    stream = torch.cuda.Stream()
    with torch.cuda.stream(stream):
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=stream):
            ops.copy_blocks(kcs, vcs, md)
    graph.replay()
    torch.cuda.synchronize()
    for l in (0, 127, 128, 149):
        assert _bytes_equal(kcs[l].cpu(), want_k[l]) and _bytes_equal(vcs[l].cpu(), want_v[l])


# ------------------------------------------------------------------ swap_blocks
@pytest.mark.parametrize("direction", [("cuda", "cpu"), ("cuda", "cuda"), ("cpu", "cuda")])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32, torch.uint8])
def test_swap_blocks(ops, co, ref, direction, dtype):
    NB, Hkv, D, BS, num_mappings = 300, 8, 80, 16, 96
    sdev = DEV if direction[0] == "cuda" else "cpu"
    ddev = DEV if direction[1] == "cuda" else "cpu"
    rnd = random.Random(1)
    src_blocks = rnd.sample(range(NB), num_mappings)
    if direction == ("cuda", "cuda"):   # same tensor family: keep sources and destinations apart
        rest = [b for b in range(NB) if b not in set(src_blocks)]
        dst_blocks = rnd.sample(rest, num_mappings)
    else:
        dst_blocks = rnd.sample(range(NB), num_mappings)
    # a run of consecutive pairs exercises the coalesced copy
    pairs = list(zip(src_blocks, dst_blocks))
    mapping = torch.tensor(pairs, dtype=torch.int64)
    src = _cache((NB, Hkv, D, BS), dtype, 3)
    dst0 = _cache((NB, Hkv, D, BS), dtype, 4)
    want = dst0.clone()
    co.swap_blocks(src, want, mapping)
    s_t = src.to(sdev).pin_memory() if sdev == "cpu" else src.to(sdev)
    d_t = dst0.clone().pin_memory() if ddev == "cpu" else dst0.to(ddev)
    ops.swap_blocks(s_t, d_t, mapping)
    torch.cuda.synchronize()
    assert _bytes_equal(d_t.cpu(), want)
    if ref is not None:
        r_t = dst0.clone().pin_memory() if ddev == "cpu" else dst0.to(ddev)
        ref.swap_blocks(s_t, r_t, mapping)
        torch.cuda.synchronize()
        assert _bytes_equal(d_t.cpu(), r_t.cpu())


def test_swap_blocks_runs_padding_and_errors(ops, co):
    # padded blocks: stride(0) > numel of a block; the whole stride is copied (cache_kernels.cu:52-55)
    base = _cache((40, 3000), torch.float16, 5)
    src = base[:, :2048].view(40, 2, 64, 16)          # stride(0) = 3000
    dst = torch.zeros(40, 3000, dtype=torch.float16, device=DEV)[:, :2048].view(40, 2, 64, 16)
    mapping = torch.tensor([[4, 10], [5, 11], [6, 12], [20, 1], [21, 2], [0, 39]], dtype=torch.int64)
    ops.swap_blocks(base.to(DEV)[:, :2048].view(40, 2, 64, 16), dst, mapping)
    torch.cuda.synchronize()
    for s, d in mapping.tolist():
        assert torch.equal(dst[d].cpu(), src[s])
    assert (dst[3].cpu() == 0).all()
    with pytest.raises(RuntimeError, match="block_mapping must be on CPU"):
        ops.swap_blocks(dst, dst, mapping.to(DEV))
    with pytest.raises((RuntimeError, NotImplementedError)):
        ops.swap_blocks(src, src.clone(), mapping)     # cpu -> cpu: invalid device combination


# ------------------------------------------------------------------ gathers
def _gather_case(dtype, kv, batch, max_len, NB, BS, entry, seed, with_starts, cp):
    rnd = random.Random(seed)
    cache_dt = torch.uint8 if kv != "auto" else dtype
    src = _cache((NB, BS, entry), cache_dt, seed)
    lens = [rnd.randint(0, max_len) for _ in range(batch)]
    lens[0] = 0 if batch > 1 else lens[0]
    cu = torch.tensor([0] + list(torch.tensor(lens).cumsum(0)), dtype=torch.int32)
    bt = torch.stack([torch.randperm(NB, generator=torch.Generator().manual_seed(seed + b))
                      for b in range(batch)]).to(torch.int32)
    starts = None
    if with_starts:
        room = [NB * BS - l for l in lens]
        if cp:
            starts = torch.tensor([rnd.randint(0, r) for r in room], dtype=torch.int32)
        else:
            starts = torch.tensor([rnd.randint(0, r // BS) * BS for r in room], dtype=torch.int32)
    return src, lens, cu, bt, starts


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
@pytest.mark.parametrize("kv", ["auto", "fp8", "fp8_e5m2"])
@pytest.mark.parametrize("entry", [576, 72, 20])        # 512+64 (MLA), 16-byte vectors, scalar path
@pytest.mark.parametrize("with_starts", [False, True])
def test_gather_and_maybe_dequant_cache(ops, co, ref, dtype, kv, entry, with_starts):
    batch, max_len, NB, BS = 8, 300, 64, 16
    src, lens, cu, bt, starts = _gather_case(dtype, kv, batch, max_len, NB, BS, entry, 7, with_starts, False)
    tot = sum(lens)
    scale = torch.tensor(0.1)
    want = torch.zeros(tot, entry, dtype=dtype)
    co.gather_and_maybe_dequant_cache(src, want, bt, cu, batch, kv, scale, starts)
    dst = torch.full((tot, entry), 7.0, dtype=dtype, device=DEV)
    sd = None if starts is None else starts.to(DEV)
    ops.gather_and_maybe_dequant_cache(src.to(DEV), dst, bt.to(DEV), cu.to(DEV), batch, kv,
                                       scale.to(DEV), sd)
    torch.cuda.synchronize()
    assert _equal_up_to_nan_payload(dst.cpu(), want)
    if kv == "fp8_e5m2":
        assert want.isnan().any() and want.isinf().any()   # the inputs do cover inf / NaN codes
    if ref is not None:
        rd = torch.full((tot, entry), 7.0, dtype=dtype, device=DEV)
        ref.gather_and_maybe_dequant_cache(src.to(DEV), rd, bt.to(DEV), cu.to(DEV), batch, kv,
                                           scale.to(DEV), sd)
        torch.cuda.synchronize()
        assert _bytes_equal(dst, rd)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.uint8])
@pytest.mark.parametrize("entry", [576, 36, 7])
@pytest.mark.parametrize("with_starts", [False, True])
def test_cp_gather_cache(ops, co, ref, dtype, entry, with_starts):
    batch, max_len, NB, BS = 8, 512, 128, 16
    src, lens, cu, bt, starts = _gather_case(dtype, "auto", batch, max_len, NB, BS, entry, 9, with_starts, True)
    tot = sum(lens)
    want = torch.zeros(tot, entry, dtype=dtype)
    co.cp_gather_cache(src, want, bt, cu, batch, starts)
    dst = torch.zeros(tot, entry, dtype=dtype, device=DEV)
    sd = None if starts is None else starts.to(DEV)
    ops.cp_gather_cache(src.to(DEV), dst, bt.to(DEV), cu.to(DEV), batch, sd)
    torch.cuda.synchronize()
    assert _bytes_equal(dst.cpu(), want)
    if ref is not None:
        rd = torch.zeros(tot, entry, dtype=dtype, device=DEV)
        ref.cp_gather_cache(src.to(DEV), rd, bt.to(DEV), cu.to(DEV), batch, sd)
        torch.cuda.synchronize()
        assert _bytes_equal(dst, rd)


def test_gather_strided_views_and_large_ragged_batch(ops, co):
    """Padded cache blocks, a destination with a row stride, a 4-D cache ([NB, BS, H, D], the flash
    NHD layout) and a batch large enough that the grid is capped and warps loop."""
    NB, BS, H, D, batch = 512, 16, 2, 64, 200
    rnd = random.Random(3)
    base = _cache((NB, BS * H * D + 64), torch.bfloat16, 11)
    src = base[:, :BS * H * D].view(NB, BS, H, D)
    lens = [rnd.choice([0, 1, 15, 16, 17, 40]) for _ in range(batch)]
    cu = torch.tensor([0] + list(torch.tensor(lens).cumsum(0)), dtype=torch.int32)
    bt = torch.randint(0, NB, (batch, 8), generator=torch.Generator().manual_seed(1), dtype=torch.int32)
    tot = sum(lens)
    want = torch.zeros(tot, H * D, dtype=torch.bfloat16)
    co.gather_and_maybe_dequant_cache(src, want, bt, cu, batch, "auto", torch.tensor(1.0))
    wide = torch.zeros(tot, H * D + 8, dtype=torch.bfloat16, device=DEV)
    dst = wide[:, :H * D]
    sdev = base.to(DEV)[:, :BS * H * D].view(NB, BS, H, D)
    ops.gather_and_maybe_dequant_cache(sdev, dst, bt.to(DEV), cu.to(DEV), batch, "auto",
                                       torch.tensor(1.0, device=DEV))
    torch.cuda.synchronize()
    assert _bytes_equal(dst.cpu(), want) and (wide[:, H * D:] == 0).all()
    dst2 = torch.zeros(tot, H * D, dtype=torch.bfloat16, device=DEV)
    ops.cp_gather_cache(sdev, dst2, bt.to(DEV), cu.to(DEV), batch)
    torch.cuda.synchronize()
    assert _bytes_equal(dst2.cpu(), want)


def test_gather_round_trip_with_reshape_and_cache_flash(ops):
    """What reshape_and_cache_flash wrote, the gather reads back (fp8: equal to convert_fp8 of
    the stored bytes) -- the caller pairing of vllm/v1/attention/backends/mla/common.py:1296."""
    BS, Hkv, D, NB = 16, 1, 576, 32
    lens = [5, 0, 48, 33]
    T = sum(lens)
    key = (torch.randn(T, Hkv, D, generator=torch.Generator().manual_seed(2)) * 0.3).to(torch.bfloat16).to(DEV)
    perm = torch.randperm(NB, generator=torch.Generator().manual_seed(3)).tolist()
    bt = torch.zeros(len(lens), 4, dtype=torch.int32)
    slots, pos = [], 0
    for b, l in enumerate(lens):
        n = (l + BS - 1) // BS
        bt[b, :n] = torch.tensor(perm[pos:pos + n], dtype=torch.int32)
        slots += [perm[pos + p // BS] * BS + p % BS for p in range(l)]
        pos += n
    cu = torch.tensor([0] + list(torch.tensor(lens).cumsum(0)), dtype=torch.int32).to(DEV)
    scale = torch.tensor(0.02, device=DEV)
    kc = torch.zeros(NB, BS, Hkv, D, dtype=torch.uint8, device=DEV)
    vc = torch.zeros_like(kc)
    ops.reshape_and_cache_flash(key, key, kc, vc, torch.tensor(slots, dtype=torch.int64, device=DEV),
                                "fp8", scale, scale)
    dst = torch.zeros(T, D, dtype=torch.bfloat16, device=DEV)
    ops.gather_and_maybe_dequant_cache(kc, dst, bt.to(DEV), cu, len(lens), "fp8", scale)
    raw = torch.zeros(T, D, dtype=torch.uint8, device=DEV)
    ops.cp_gather_cache(kc, raw, bt.to(DEV), cu, len(lens))
    want = torch.empty_like(dst)
    ops.convert_fp8(want, raw, float(scale), "fp8")
    torch.cuda.synchronize()
    assert _bytes_equal(dst, want)
    torch.testing.assert_close(dst.float(), key.flatten(1).float(), atol=0.02 * 0.0625 * 8 + 1e-3, rtol=0.07)


def test_opcheck_block_ops(ops):
    utils = ("test_schema", "test_faketensor")
    src = _cache((8, 16, 64), torch.bfloat16, 0).to(DEV)
    cu = torch.tensor([0, 5, 21], dtype=torch.int32, device=DEV)
    bt = torch.tensor([[1, 2], [3, 4]], dtype=torch.int32, device=DEV)
    dst = torch.zeros(21, 64, dtype=torch.bfloat16, device=DEV)
    one = torch.tensor(1.0, device=DEV)
    for ns in (torch.ops.b200pa, torch.ops._C_cache_ops):
        torch.library.opcheck(ns.gather_and_maybe_dequant_cache, (src, dst, bt, cu, 2, "auto", one, None),
                              test_utils=utils)
        torch.library.opcheck(ns.cp_gather_cache, (src, dst, bt, cu, 2, None), test_utils=utils)
        torch.library.opcheck(ns.swap_blocks, (src, src.clone(), torch.tensor([[0, 1]], dtype=torch.int64)),
                              test_utils=utils)
        torch.library.opcheck(ns.copy_blocks, ([src], [src.clone()], torch.tensor([[0, 1]], dtype=torch.int64, device=DEV)),
                              test_utils=("test_schema",))

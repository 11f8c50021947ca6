"""The workloads bench.py TIMES, at the size it times them, against the reference's own CUDA
kernels (oracle tier O1: ``oracle/_ref/libvllm_ref_cuda.so`` = ``csrc/attention/
paged_attention_v2.cu`` + ``csrc/cache_kernels.cu`` of the reference compiled unmodified for
sm_100a).

``bench.parity_check`` is the function the bench itself runs before its timed region: ONE step of
the op on the bench's seeded tensors, the whole ``[B, H, D]`` output within ``1e-2 + 1e-2 |ref|``
and the whole cache ``torch.equal`` to the reference's ``reshape_and_cache`` + ``paged_attention_v2``
on a clone.  Here it runs for every BASELINE config at FULL size (cfg4 tp=1: B=128, 64q/8kv,
ctx 32768, 8 GiB of fp8 KV -- the BENCH line; cfg5: B=64 ragged 128..131072), fused step and
two-op step.  Mirrors the reference's ``tests/kernels/attention/test_attention.py:121-345``
(kernel vs. reference implementation on identical tensors)."""
import os

import pytest
import torch

pytestmark = pytest.mark.gpu

DEV = torch.device("cuda", 0)


@pytest.fixture(scope="module")
def bench():
    from oracle import ref_cuda
    if not os.path.exists(ref_cuda.LIB_PATH):
        pytest.skip("oracle/_ref/libvllm_ref_cuda.so not built")
    import bench as b
    return b


@pytest.mark.parametrize("fused", [True, False], ids=["decode_step", "two_ops"])
@pytest.mark.parametrize("cfg", ["cfg1", "cfg2", "cfg3", "cfg4", "cfg5"])
def test_timed_workload_matches_reference_cuda_at_full_size(bench, cfg, fused):
    need = 2.3 * bench.kv_bytes(bench.seq_lens_of(cfg), bench.CONFIGS[cfg][3], bench.CONFIGS[cfg][5])
    free, _ = torch.cuda.mem_get_info(DEV)
    if free < need:
        pytest.skip(f"needs {need / 2**30:.1f} GiB of free HBM")
    w = bench.make_shard(cfg, 0, 1, DEV, copies=1)
    out = torch.empty(w["B"], w["H"] * bench.D, dtype=torch.bfloat16, device=DEV)
    res = bench.parity_check(cfg, w, out, fused, 0, 1, DEV)
    del w, out
    torch.cuda.empty_cache()
    assert res["checked"], res
    assert res["cache_bytes_equal"], res
    assert res["ok"], res
    assert res["max_abs"] <= 1e-2 + 1e-2, res


@pytest.mark.parametrize("tp", [2, 8])
def test_kv_head_shard_of_cfg4_matches_reference_cuda(bench, tp):
    """What rank 0 of a tp-way run computes (Hkv/tp KV heads of the full batch)."""
    cfg = "cfg4"
    w = bench.make_shard(cfg, 0, tp, DEV, copies=1)
    out = torch.empty(w["B"], w["H"] * bench.D, dtype=torch.bfloat16, device=DEV)
    res = bench.parity_check(cfg, w, out, True, 0, 1, DEV)
    del w, out
    torch.cuda.empty_cache()
    assert res["checked"] and res["ok"] and res["cache_bytes_equal"], res

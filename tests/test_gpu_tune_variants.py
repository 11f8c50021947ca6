"""Every launch shape of the decode kernels behind the tuning hook (B200PA_TESTING builds: the
whole-block kernel's heads per CTA x stages x CTAs per SM, the per-warp stream kernel's warps x
stages x pages per stage x CTAs per SM) gives the default shape's result: attention within rounding of it and of the oracle, fused cache bytes
bit-exact.  The shapes only differ in how the flattened stages are cut into spans."""
import pytest
import torch

from tests import helpers as hp

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.fixture()
def native():
    from vllm_main_modify_b200 import _native
    if not _native.has_testing_hooks():
        pytest.skip("library built without the test hooks")
    yield _native
    _native.set_stream_tune(0)


@pytest.mark.parametrize("kv", ["auto", "fp8"])
# 20-24: launch shapes of the whole-block kernel (csrc/paged_attention_nhd.cuh); 100 + i: the
# per-warp stream kernel with its launch shape i
@pytest.mark.parametrize("tune", [20, 21, 22, 23, 24] + [100 + i for i in range(0, 14)])
def test_tune_variant_matches_default_and_oracle(native, tune, kv):
    from oracle import oracle as ora
    from vllm_main_modify_b200 import _custom_ops as ops
    B, H, Hkv, D, BS = 7, 32, 4, 128, 16
    lens = [1, 16, 17, 700, 2049, 333, 5000]
    NB = sum((L + BS - 1) // BS for L in lens) + 2
    s = D ** -0.5
    q, kc, vc, bt, sl = hp.make_paged_inputs(B, H, Hkv, D, BS, NB, lens, torch.bfloat16, seed=tune,
                                             kv=kv, lo=-1.0, hi=1.0, permute=True)
    g = torch.Generator().manual_seed(100 + tune)
    key = (torch.rand(B, Hkv, D, generator=g) * 2 - 1).to(torch.bfloat16)
    val = (torch.rand(B, Hkv, D, generator=g) * 2 - 1).to(torch.bfloat16)
    slots = torch.tensor([int(bt[b, (L - 1) // BS]) * BS + (L - 1) % BS
                          for b, L in enumerate(lens)], dtype=torch.int64)
    ks, vs = (1.0, 1.0) if kv == "auto" else (0.011, 0.013)
    ksd, vsd = torch.tensor(ks, device=DEV), torch.tensor(vs, device=DEV)

    def step(t):
        native.set_stream_tune(t)
        kcd, vcd = kc.to(DEV), vc.to(DEV)
        out = torch.full_like(q, float("nan"), device=DEV)
        ops.decode_step(out, None, q.to(DEV), key.to(DEV), val.to(DEV), kcd, vcd, slots.to(DEV),
                        Hkv, s, bt.to(DEV), sl.to(DEV), BS, max(lens), None, kv, ksd, vsd)
        torch.cuda.synchronize()
        return out.cpu(), kcd.cpu(), vcd.cpu()

    base, kc0, vc0 = step(0)
    got, kc1, vc1 = step(tune)
    assert torch.equal(kc0.view(torch.uint8), kc1.view(torch.uint8))
    assert torch.equal(vc0.view(torch.uint8), vc1.view(torch.uint8))
    torch.testing.assert_close(got.float(), base.float(), atol=2e-3, rtol=2e-3)
    kc_o, vc_o = kc.clone(), vc.clone()
    ora.reshape_and_cache(key, val, kc_o, vc_o, slots, kv, torch.tensor(ks), torch.tensor(vs))
    assert torch.equal(kc_o.view(torch.uint8), kc1.view(torch.uint8))
    want = torch.empty_like(q)
    ora.paged_attention_v1(want, q, kc_o, vc_o, Hkv, s, bt, sl, BS, max(lens), None, kv,
                           torch.tensor(ks), torch.tensor(vs))
    torch.testing.assert_close(got.float(), want.float(), atol=1e-2, rtol=1e-2)

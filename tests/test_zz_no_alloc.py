"""b200pa_paged_attention_v1_builtin_scratch: the decode kernel for hosts that cannot allocate.

The regular entry point, b200pa_paged_attention_v1, expects its caller to cudaMalloc a ~23.5 MiB
workspace and pass it in.  A host that cannot do that -- a plugin inside an engine whose KV-cache
pool already owns all free HBM, an FFI host that only receives device pointers and has no CUDA
binding to allocate with, a loop that must not hit cudaMalloc's device-wide sync -- calls this
variant instead: same arguments minus the workspace, the scratch is a 24 MiB static device array
inside libb200pa.so, and the call allocates nothing.

Checked here, through the C ABI with raw device pointers (ctypes), against the existing entry
point on the same inputs: two calls in a row with different batches (the second runs on whatever
the first left in the shared array), and the refusal of a head size whose scratch does not fit.

This file sorts last on purpose (`pytest --collect-only`): under `pytest -x` a failure here cannot
hide any other test.  It passes on its own in a fresh process (the first call does the per-device
address look-up) and after the rest of the suite.
"""
import ctypes

import pytest
import torch

from tests import helpers as hp

pytestmark = pytest.mark.gpu

DEV = "cuda:0"
ATOL, RTOL = 1e-2, 1e-2      # the gate the other GPU tests use (BASELINE.json north_star)


def _inputs(lens, seed, kv):
    B, H, Hkv, D, BS, NB = len(lens), 16, 4, 128, 16, 700
    s = D ** -0.5
    q, kc, vc, bt, sl = hp.make_paged_inputs(B, H, Hkv, D, BS, NB, lens, torch.bfloat16, seed=seed,
                                             kv=kv, lo=-s, hi=s, permute=True)
    return (B, H, Hkv, D, BS, s, max(lens)), tuple(t.to(DEV) for t in (q, kc, vc, bt, sl))


def _args(shape, tensors, out, ks, vs, kv):
    B, H, Hkv, D, BS, s, max_len = shape
    qd, kcd, vcd, btd, sld = tensors
    p = lambda t: ctypes.c_void_p(t.data_ptr())
    return [p(out), p(qd), p(kcd), p(vcd), B, H, Hkv, D, BS, ctypes.c_float(s), p(btd),
            ctypes.c_int(btd.shape[1]), p(sld), max_len, None, ctypes.c_int64(qd.stride(0)),
            ctypes.c_int64(kcd.stride(0)), ctypes.c_int64(kcd.stride(1)), 2,
            0 if kv == "auto" else 1, p(ks), p(vs), 0, 0, 0, 64, 0]


@pytest.mark.parametrize("kv", ["auto", "fp8"])
def test_builtin_scratch_matches_caller_workspace_on_two_calls_in_a_row(kv):
    from vllm_main_modify_b200 import _native
    lib = _native.c_lib()
    lib.b200pa_paged_attention_workspace_bytes.restype = ctypes.c_int64
    need = lib.b200pa_paged_attention_workspace_bytes(128)
    # the caller-allocated workspace of the existing entry point (header zeroed, as its contract asks)
    ws = torch.zeros(need, dtype=torch.uint8, device=DEV)
    ks = torch.tensor(1.0 if kv == "auto" else 0.02, device=DEV)
    vs = torch.tensor(1.0 if kv == "auto" else 0.03, device=DEV)
    # long sequences are cut across CTAs, so partials and ticket counters really go through the
    # scratch; the second batch has other cuts and finds the first batch's leftovers in the array
    batches = [([4100, 37, 1, 900, 2049, 16, 300], 5), ([77, 6000, 512, 3, 1500], 6)]
    for lens, seed in batches:
        shape, tensors = _inputs(lens, seed, kv)
        want = torch.full_like(tensors[0], float("nan"))
        got = torch.full_like(tensors[0], float("nan"))
        a = _args(shape, tensors, want, ks, vs, kv)
        assert lib.b200pa_paged_attention_v1(*a, ctypes.c_void_p(ws.data_ptr()),
                                             ctypes.c_int64(need), None) == 0, lib.b200pa_last_error()
        a = _args(shape, tensors, got, ks, vs, kv)
        n0 = _native.launch_count()
        assert lib.b200pa_paged_attention_v1_builtin_scratch(*a, None) == 0, lib.b200pa_last_error()
        assert _native.launch_count() - n0 == 1      # the same single kernel, not a slower stand-in
        torch.cuda.synchronize()
        assert not torch.isnan(got.float()).any()
        torch.testing.assert_close(got.float(), want.float(), atol=ATOL, rtol=RTOL)


def test_builtin_scratch_refuses_what_does_not_fit():
    """Head size 256 needs about twice the 24 MiB the library carries: a clear error, no launch."""
    from vllm_main_modify_b200 import _native
    lib = _native.c_lib()
    B, H, Hkv, D, BS, NB = 2, 8, 2, 256, 16, 20
    q, kc, vc, bt, sl = hp.make_paged_inputs(B, H, Hkv, D, BS, NB, [40, 100], torch.bfloat16, seed=1)
    tensors = tuple(t.to(DEV) for t in (q, kc, vc, bt, sl))
    out = torch.zeros_like(tensors[0])
    one = torch.tensor(1.0, device=DEV)
    n0 = _native.launch_count()
    rc = lib.b200pa_paged_attention_v1_builtin_scratch(
        *_args((B, H, Hkv, D, BS, D ** -0.5, 100), tensors, out, one, one, "auto"), None)
    assert rc == 1                                   # B200PA_ERR_UNSUPPORTED
    msg = lib.b200pa_last_error()
    assert b"head size 256" in msg and b"built-in scratch holds 25165824" in msg, msg
    assert _native.launch_count() == n0
    torch.cuda.synchronize()
    assert int(out.abs().max()) == 0

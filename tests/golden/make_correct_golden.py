#!/usr/bin/env python
"""Generate tests/golden/correct_*.npz (build container only: needs /root/reference).

Source of truth: the reference's own Triton kernel ``_correct_attn_cp_out_kernel``
(vllm/attention/ops/common.py:9-68), extracted at run time (ast) and EXECUTED here, one program
instance per (batch, head), against a small torch implementation of the handful of ``tl.``
primitives it uses (program_id, arange, load, store, where, max, sum, exp, log) -- Triton itself
cannot run without a GPU.  ``tl.load`` yields fp32 values and ``tl.store`` rounds to the
destination type, which is what the Triton kernel does when ``out`` is 16-bit (the factor is an
fp32 scalar, so the product is computed in fp32).  Nothing here is imported by the product or
the tests; tests only read the .npz files.
"""
from __future__ import annotations

import ast
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("REFERENCE_ROOT", "/root/reference")


class _Ptr:
    """A 'pointer': flat view of a tensor plus element offsets."""

    def __init__(self, flat, off=0):
        self.flat, self.off = flat, off

    def __add__(self, other):
        return _Ptr(self.flat, self.off + other)


class _PID(int):
    def to(self, _dtype):
        return int(self)


class _TL:
    constexpr = int
    int64 = torch.int64

    def __init__(self):
        self.pid = (0, 0, 0)

    def program_id(self, axis):
        return _PID(self.pid[axis])

    @staticmethod
    def arange(a, b):
        return torch.arange(a, b, dtype=torch.int64)

    @staticmethod
    def load(ptr):
        off = ptr.off
        v = ptr.flat[off] if not isinstance(off, torch.Tensor) else ptr.flat[off.long()]
        return v.float() if isinstance(v, torch.Tensor) else torch.tensor(float(v))

    @staticmethod
    def store(ptr, val):
        off = ptr.off
        val = torch.as_tensor(val).to(ptr.flat.dtype)
        if isinstance(off, torch.Tensor):
            ptr.flat[off.long()] = val
        else:
            ptr.flat[off] = val

    @staticmethod
    def where(c, a, b):
        a = torch.as_tensor(a, dtype=torch.float32)
        b = torch.as_tensor(b, dtype=torch.float32)
        return torch.where(c, a, b)

    @staticmethod
    def max(x, axis=0):
        return x.max(dim=axis).values if x.dim() else x

    @staticmethod
    def sum(x, axis=0):
        return x.sum(dim=axis) if x.dim() else x

    exp = staticmethod(torch.exp)
    log = staticmethod(torch.log)


def load_kernel():
    path = os.path.join(REF, "vllm/attention/ops/common.py")
    src = open(path).read()
    tree = ast.parse(src)
    chunks = [ast.get_source_segment(src, n) for n in tree.body
              if isinstance(n, ast.FunctionDef) and n.name == "_correct_attn_cp_out_kernel"]
    assert len(chunks) == 1
    tl = _TL()
    ns = {"tl": tl}
    exec(chunks[0], ns)            # the decorator line is not part of the segment
    return ns["_correct_attn_cp_out_kernel"], tl


def run_reference(out, lses, rank):
    """correct_attn_out (common.py:86-116): grid (B, H, 1), in place on `out`."""
    kern, tl = load_kernel()
    B, H, D = out.shape
    N = lses.shape[0]
    lse = torch.empty_like(lses[0])
    flat_out, flat_lses, flat_lse = out.view(-1), lses.view(-1), lse.view(-1)
    for b in range(B):
        for h in range(H):
            tl.pid = (b, h, 0)
            kern(_Ptr(flat_out), _Ptr(flat_out), _Ptr(flat_lses), _Ptr(flat_lse), *out.stride(),
                 *lses.stride(), rank, HEAD_DIM=D, N_ROUNDED=N)
    return out, lse


def bits(t):
    if t.dtype in (torch.bfloat16, torch.float16):
        return t.contiguous().view(torch.int16).numpy().view(np.uint16)
    return t.contiguous().numpy()


def make_case(name, N, B, H, D, dtype, rank, seed):
    torch.manual_seed(seed)
    out = torch.randn(B, H, D).to(dtype)
    lses = torch.randn(N, B, H) * 3
    lses[0, 0, 0] = float("inf")          # sanitised to -inf (:41)
    lses[N - 1, B - 1, H - 1] = float("nan")
    lses[:, 1 % B, 0] = -float("inf")     # a row no rank saw
    lses[rank, 2 % B, 1 % H] = -float("inf")   # this rank saw nothing of the row
    got, lse = run_reference(out.clone(), lses.clone(), rank)
    path = os.path.join(HERE, f"correct_{name}.npz")
    np.savez_compressed(path, out=bits(out), lses=lses.numpy(), rank=np.int32(rank),
                        out_ref=bits(got), lse_ref=lse.numpy(), dtype=str(dtype).split(".")[-1])
    print(os.path.basename(path), os.path.getsize(path))


if __name__ == "__main__":
    make_case("f32_n4", 4, 5, 3, 32, torch.float32, 1, 0)
    make_case("bf16_n8", 8, 7, 8, 128, torch.bfloat16, 5, 1)
    make_case("f16_n2", 2, 3, 4, 64, torch.float16, 0, 2)

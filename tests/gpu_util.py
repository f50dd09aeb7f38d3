"""Helpers for the -m gpu parity tests: call libb200gate's C ABI on torch-allocated device memory.
torch is used for allocation and copies only; every computation under test is a b200_op_* call."""
import ctypes as C

import numpy as np
import torch

from vgate_b200 import _lib


def lib():
    return _lib.load()


_live = []      # device tensors made by dev() stay alive until the next check(): `ptr(dev(x))` in a
                # call expression must not hand the kernel memory that torch has already recycled


def dev(a, dtype=None):
    t = torch.from_numpy(np.ascontiguousarray(a))
    if dtype is not None:
        t = t.to(dtype)
    t = t.cuda()
    _live.append(t)
    return t


def ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


def stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def check(code):
    _lib.check(code)
    torch.cuda.synchronize()
    del _live[:]


def repack(qweight, qzeros, scales, G=128):
    """AWQ numpy tensors -> (wq, sc, zp) device tensors in kernel layout."""
    K, N = qweight.shape[0], qweight.shape[1] * 8
    qw, qz, sc_in = dev(qweight), dev(qzeros), dev(scales)
    wq = torch.empty((N, K // 8), dtype=torch.int32, device="cuda")
    sc = torch.empty((N, K // G), dtype=torch.float16, device="cuda")
    zp = torch.empty((N, K // G), dtype=torch.float16, device="cuda")
    check(lib().b200_op_awq_repack(ptr(qw), ptr(qz), ptr(sc_in), K, N, G, ptr(wq), ptr(sc), ptr(zp), stream()))
    return wq, sc, zp


def random_awq(rng, K, N, G=128, w_std=0.02):
    qweight = rng.integers(0, 2**32, size=(K, N // 8), dtype=np.uint32).view(np.int32)
    qzeros = rng.integers(0, 2**32, size=(K // G, N // 8), dtype=np.uint32).view(np.int32)
    scales = ((w_std / 6.52) * (0.5 + rng.random((K // G, N), dtype=np.float32))).astype(np.float16)
    return qweight, qzeros, scales


def swizzle_kv(a):
    """Logical K/V blocks [..., 16 tokens, 128 dims] -> the cache's physical row layout: 16-byte chunk c
    of the row at token offset r is stored at chunk position c ^ (r & 7) (csrc/common.cuh kv_swz)."""
    a = np.asarray(a)
    out = np.empty_like(a)
    v = a.reshape(a.shape[:-2] + (16, 16, 8))
    o = out.reshape(v.shape)
    for r in range(16):
        for c in range(16):
            o[..., r, c ^ (r & 7), :] = v[..., r, c, :]
    return out


unswizzle_kv = swizzle_kv        # the XOR permutation is its own inverse

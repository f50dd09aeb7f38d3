"""The allocation-free attention entry point (b200_op_paged_attn_noalloc).

It exists for callers that cannot allocate device memory: its split-KV scratch and arrival counters
are arrays compiled into the library.  The result must be what b200_op_paged_attn returns when the
caller brings the scratch -- on the first call and again on the second, because the built-in arrays
are shared by every call and still hold what the previous one wrote.  This file sorts last so that a
failure here cannot hide the outcome of any other test in a run that stops at the first failure.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from tests.gpu_util import lib, ptr, stream, check  # noqa: E402
from vgate_b200 import _lib  # noqa: E402

TOL = 4e-3          # the tolerance of the attention parity tests (tests/test_kernels_gpu.py)


def _case(rng, ctx_lens, n_heads, n_kv):
    n_seqs = len(ctx_lens)
    nb_seq = (max(ctx_lens) + 15) // 16 + 1
    n_blocks = n_seqs * nb_seq
    perm = rng.permutation(n_blocks).astype(np.int32).reshape(n_seqs, nb_seq)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    return {
        "q": t(rng.standard_normal((n_seqs, n_heads * 128)).astype(np.float16)),
        "kc": t(rng.standard_normal((n_blocks, n_kv, 16, 128)).astype(np.float16)),
        "vc": t(rng.standard_normal((n_blocks, n_kv, 16, 128)).astype(np.float16)),
        "bt": t(perm), "btrow": t(np.arange(n_seqs, dtype=np.int32)), "ctx": t(np.array(ctx_lens, dtype=np.int32)),
        "n": n_seqs, "stride": nb_seq,
    }


def _with_caller_scratch(c, n_heads, n_kv, n_splits):
    out = torch.empty_like(c["q"])
    scratch = torch.empty(lib().b200_op_attn_scratch_bytes(c["n"], n_heads, n_kv, n_splits) // 4, dtype=torch.float32,
                          device="cuda")
    counters = torch.zeros(c["n"] * n_kv, dtype=torch.int32, device="cuda")
    check(lib().b200_op_paged_attn(ptr(c["q"]), ptr(c["kc"]), ptr(c["vc"]), ptr(c["bt"]), c["stride"], ptr(c["btrow"]),
                                   ptr(c["ctx"]), c["n"], n_heads, n_kv, n_splits, ptr(scratch), ptr(counters), ptr(out),
                                   stream()))
    return out


def _without_alloc(c, n_heads, n_kv, n_splits):
    out = torch.empty_like(c["q"])
    check(lib().b200_op_paged_attn_noalloc(ptr(c["q"]), ptr(c["kc"]), ptr(c["vc"]), ptr(c["bt"]), c["stride"],
                                           ptr(c["btrow"]), ptr(c["ctx"]), c["n"], n_heads, n_kv, n_splits, ptr(out),
                                           stream()))
    return out


def test_noalloc_matches_caller_scratch_twice_in_a_row():
    rng = np.random.default_rng(99)
    n_heads, n_kv, n_splits = 12, 2, 8
    first = _case(rng, [1, 63, 64, 65, 200, 384, 512, 700], n_heads, n_kv)
    second = _case(rng, [300, 17, 640, 128, 129, 1, 90, 1000], n_heads, n_kv)      # different shapes of work
    for c in (first, second):
        ref = _with_caller_scratch(c, n_heads, n_kv, n_splits)
        got = _without_alloc(c, n_heads, n_kv, n_splits)
        assert (ref.float() - got.float()).abs().max().item() < TOL
    # and the very same call again: leftovers of the previous call must not leak into this one
    ref = _with_caller_scratch(first, n_heads, n_kv, n_splits)
    got = _without_alloc(first, n_heads, n_kv, n_splits)
    assert (ref.float() - got.float()).abs().max().item() < TOL


def test_noalloc_rejects_sizes_beyond_the_built_in_scratch():
    rng = np.random.default_rng(5)
    c = _case(rng, [64] * 100, 12, 2)            # 100 rows x 12 heads x 8 splits = 9600 > 8192
    out = torch.empty_like(c["q"])
    code = lib().b200_op_paged_attn_noalloc(ptr(c["q"]), ptr(c["kc"]), ptr(c["vc"]), ptr(c["bt"]), c["stride"],
                                            ptr(c["btrow"]), ptr(c["ctx"]), c["n"], 12, 2, 8, ptr(out), stream())
    assert code == _lib.ERR_CAPACITY
    assert b"exceeds the built-in scratch" in lib().b200_last_error()

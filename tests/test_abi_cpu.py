"""The C-ABI surface without a GPU: the shared library loads, exports every symbol that
include/b200gate.h (the boundary) and include/b200gate_debug.h (profiling aids) declare, the ctypes table
covers the same set, and no compute is needed for that."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "b200gate.h")
DEBUG_HEADER = os.path.join(ROOT, "include", "b200gate_debug.h")


def declared_functions(headers=(HEADER, DEBUG_HEADER)):
    names = set()
    for h in headers:
        src = re.sub(r"/\*.*?\*/", "", open(h).read(), flags=re.S)
        names |= set(re.findall(r"\b(b200_[a-z0-9_]+)\s*\(", src))
    return sorted(names)


def test_header_declares_the_documented_boundary():
    names = declared_functions((HEADER,))
    assert not [n for n in names if n.startswith("b200_debug_")]      # profiling aids live in their own header
    assert all(n.startswith("b200_debug_") for n in declared_functions((DEBUG_HEADER,)))
    for must in ("b200_engine_create", "b200_engine_load_tensor", "b200_engine_finalize", "b200_engine_destroy",
                 "b200_submit", "b200_wait", "b200_poll", "b200_abort", "b200_stats_get", "b200_last_error",
                 "b200_op_awq_dequant", "b200_op_w4_gemv", "b200_op_w4_gemm_umma", "b200_op_f16_gemm_umma",
                 "b200_op_paged_attn", "b200_op_attn_decode_fused", "b200_op_rmsnorm", "b200_op_rope_kv",
                 "b200_op_swiglu", "b200_op_sample"):
        assert must in names


def test_library_exports_every_declared_symbol():
    from vgate_b200 import _lib
    lib = _lib.load()                       # raises ImportError if the .so is missing or incomplete
    raw = ctypes.CDLL(_lib.LIB_PATH)
    for name in declared_functions():
        assert hasattr(raw, name), f"{name} declared in b200gate.h but not exported"
    assert lib.b200_abi_version() == _lib.ABI_VERSION


def test_ctypes_table_matches_header():
    from vgate_b200 import _lib
    assert sorted(_lib.SIGNATURES) == declared_functions()


def test_config_struct_size_matches_c_layout():
    from vgate_b200 import _lib
    # 13 int32/float fields + padding before int64 kv_bytes ... the library checks struct_size itself;
    # here: the Python mirror is self-consistent and 8-byte aligned
    assert ctypes.sizeof(_lib.Config) % 8 == 0
    assert ctypes.sizeof(_lib.Stats) % 8 == 0


def test_missing_library_fails_loudly(monkeypatch, tmp_path):
    from vgate_b200 import _lib
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", str(tmp_path / "nope.so"))
    with pytest.raises(ImportError, match="no CPU fallback"):
        _lib.load()


def test_calls_without_a_gpu_report_errors_not_crashes():
    """No compute here: argument validation happens before any CUDA call, and a missing device is an
    error code with a message, never an abort."""
    from vgate_b200 import _lib
    lib = _lib.load()
    cfg = _lib.Config()
    cfg.struct_size = 4                      # wrong on purpose
    h = ctypes.c_void_p()
    assert lib.b200_engine_create(ctypes.byref(cfg), ctypes.byref(h)) == _lib.ERR_INVALID
    assert b"struct_size" in lib.b200_last_error()
    assert lib.b200_op_w4_effective_splits(8960, 128, 12) == 12
    assert lib.b200_op_w4_effective_splits(1536, 128, 5) == 4          # 12 groups / ceil(12/5)=3 per split
    assert lib.b200_op_w4_tile_bytes(1536, 2048, 128) == 16 * 12 * (512 + 2 * 4096)
    assert lib.b200_op_attn_scratch_bytes(64, 12, 2, 8) == 64 * 2 * 8 * 6 * 130 * 4

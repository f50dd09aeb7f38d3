"""ctypes binding of libb200gate.so (include/b200gate.h).  There is no CPU fallback:
if the shared library is missing or does not export the ABI, importing this fails loudly."""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("VGATE_B200_LIB", os.path.join(HERE, "libb200gate.so"))
ABI_VERSION = 2

OK = 0
ERR_INVALID, ERR_CUDA, ERR_CAPACITY, ERR_STATE, ERR_NOT_FOUND, ERR_ABORTED = -1, -2, -3, -4, -5, -6
DTYPE_F16, DTYPE_I32, DTYPE_F32, DTYPE_BF16 = 0, 1, 2, 3


class B200Error(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"libb200gate error {code}: {msg}")
        self.code = code


class Config(C.Structure):
    _fields_ = [
        ("struct_size", C.c_int32), ("device", C.c_int32), ("hidden", C.c_int32), ("intermediate", C.c_int32),
        ("n_layers", C.c_int32), ("n_heads", C.c_int32), ("n_kv_heads", C.c_int32), ("head_dim", C.c_int32),
        ("vocab", C.c_int32), ("group_size", C.c_int32), ("tie_embeddings", C.c_int32), ("rms_eps", C.c_float),
        ("rope_theta", C.c_float), ("max_seqs", C.c_int32), ("max_model_len", C.c_int32),
        ("max_batch_tokens", C.c_int32), ("kv_bytes", C.c_int64), ("eos_token_id", C.c_int32),
        ("use_cuda_graphs", C.c_int32), ("gemm_path", C.c_int32), ("act_dtype", C.c_int32), ("dense_weights", C.c_int32),
        ("tp_size", C.c_int32), ("tp_rank", C.c_int32), ("reserved", C.c_int32 * 4),
    ]


class Stats(C.Structure):
    _fields_ = [
        ("steps_prefill", C.c_int64), ("steps_decode", C.c_int64), ("tokens_prefill", C.c_int64),
        ("tokens_generated", C.c_int64), ("gpu_seconds_prefill", C.c_double), ("gpu_seconds_decode", C.c_double),
        ("decode_batch_sum", C.c_int64), ("hbm_bytes_decode", C.c_double), ("kernel_launches", C.c_int64),
        ("requests_finished", C.c_int64), ("requests_aborted", C.c_int64), ("running", C.c_int32),
        ("waiting", C.c_int32), ("kv_blocks_total", C.c_int32), ("kv_blocks_free", C.c_int32),
        ("prefix_hit_tokens", C.c_int64), ("prefix_query_tokens", C.c_int64),
    ]

    def to_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


_vp, _i32, _i64, _f32, _f64 = C.c_void_p, C.c_int32, C.c_int64, C.c_float, C.c_double
_pi32, _pf64 = C.POINTER(C.c_int32), C.POINTER(C.c_double)

# name -> (restype, argtypes); every symbol include/b200gate.h declares
SIGNATURES = {
    "b200_abi_version": (_i32, []),
    "b200_last_error": (C.c_char_p, []),
    "b200_engine_create": (_i32, [C.POINTER(Config), C.POINTER(_vp)]),
    "b200_engine_load_tensor": (_i32, [_vp, C.c_char_p, _vp, _i32, _i32, C.POINTER(C.c_int64)]),
    "b200_engine_finalize": (_i32, [_vp]),
    "b200_engine_destroy": (None, [_vp]),
    "b200_submit": (_i64, [_vp, _vp, _i32, _f32, _f32, _i32, C.c_uint64]),
    "b200_wait": (_i32, [_vp, _i64, _vp, _i32, _pi32, _pf64, _pf64]),
    "b200_poll": (_i32, [_vp, _i64, _vp, _i32, _pi32, _pi32, _i32]),
    "b200_abort": (_i32, [_vp, _i64]),
    "b200_stats_get": (_i32, [_vp, C.POINTER(Stats)]),
    "b200_engine_status": (_i32, [_vp]),
    "b200_tp_unique_id": (_i32, [_vp]),
    "b200_engine_tp_init": (_i32, [_vp, _vp]),
    "b200_engine_tp_attach": (_i32, [_vp, _vp]),
    "b200_generate_batch": (_i32, [_vp, _vp, _i32, _i32, _i32, _vp, _vp, _i32, _pf64]),
    "b200_op_awq_repack": (_i32, [_vp, _vp, _vp, _i32, _i32, _i32, _vp, _vp, _vp, _vp]),
    "b200_op_awq_dequant": (_i32, [_vp, _vp, _vp, _i32, _i32, _i32, _vp, _vp]),
    "b200_op_w4_gemv": (_i32, [_vp, _vp, _vp, _vp, _i32, _vp, _i32, _vp, _i32, _i32, _i32, _i32, _i32, _vp]),
    "b200_op_w4_tile": (_i32, [_vp, _vp, _vp, _i32, _i32, _i32, _i32, _vp, _vp]),
    "b200_op_w4_tile_bytes": (_i64, [_i32, _i32, _i32]),
    "b200_op_w4_gemm_umma": (_i32, [_vp, _vp, _i32, _vp, _i32, _vp, _i32, _i32, _i32, _i32, _i32, _i32, _i64, _vp]),
    "b200_op_w4_effective_splits": (_i32, [_i32, _i32, _i32]),
    "b200_debug_umma_trace": (None, [_vp]),
    "b200_debug_umma_trace_grid": (None, [_vp, _i32, _i32]),
    "b200_debug_attn_trace": (None, [_vp]),
    "b200_debug_timeline": (None, [_vp, _vp, C.c_uint32]),
    "b200_op_f16_gemm_umma": (_i32, [_vp, _vp, _i32, _vp, _i32, _i32, _i32, _i32, _i32, _vp]),
    "b200_op_dense_gemm_umma": (_i32, [_vp, _vp, _i32, _vp, _i32, _vp, _i32, _i32, _i32, _i32, _i32, _i32, _i64, _vp]),
    "b200_op_dense_effective_splits": (_i32, [_i32, _i32]),
    "b200_op_lm_head_argmax": (_i32, [_vp, _vp, _i32, _i32, _i32, _i32, _vp, _vp, _vp]),
    "b200_op_f16_gemv": (_i32, [_vp, _vp, _i32, _vp, _i32, _i32, _i32, _i32, _i32, _vp]),
    "b200_op_rmsnorm": (_i32, [_vp, _vp, _i32, _i64, _vp, _f32, _vp, _i32, _i32, _vp, _vp]),
    "b200_op_rope_kv": (_i32, [_vp, _i32, _i64, _vp, _vp, _vp, _vp, _vp, _i32, _i32, _i32, _vp, _vp, _vp, _vp]),
    "b200_op_paged_attn": (_i32, [_vp, _vp, _vp, _vp, _i32, _vp, _vp, _i32, _i32, _i32, _i32, _vp, _vp, _vp, _vp]),
    "b200_op_attn_scratch_bytes": (_i64, [_i32, _i32, _i32, _i32]),
    "b200_op_paged_attn_noalloc": (_i32, [_vp, _vp, _vp, _vp, _i32, _vp, _vp, _i32, _i32, _i32, _i32, _vp, _vp]),
    "b200_op_attn_decode_fused": (_i32, [_vp, _i32, _i64, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i32, _vp, _vp, _i32, _i32,
                                         _i32, _i32, _vp, _vp, _vp, _vp]),
    "b200_op_swiglu": (_i32, [_vp, _i32, _i32, _vp, _vp]),
    "b200_op_embed": (_i32, [_vp, _vp, _i32, _i32, _vp, _vp]),
    "b200_op_sample": (_i32, [_vp, _i32, _i32, _vp, _vp, _vp, _vp, _vp, _vp]),
}

_lib = None


def load():
    """Load libb200gate.so once; raise ImportError (never fall back) if it cannot be used."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a).  The b200 backend has no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        try:
            fn = getattr(lib, name)
        except AttributeError as exc:
            raise ImportError(f"{LIB_PATH} does not export {name}") from exc
        fn.restype, fn.argtypes = res, args
    if lib.b200_abi_version() != ABI_VERSION:
        raise ImportError(f"{LIB_PATH}: ABI version {lib.b200_abi_version()} != {ABI_VERSION}")
    _lib = lib
    return lib


def check(code):
    if code < 0:
        raise B200Error(code, load().b200_last_error().decode("utf-8", "replace"))
    return code

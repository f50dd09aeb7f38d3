"""Thin Python handle on a libb200gate engine (include/b200gate.h).  Host-side plumbing only:
token ids in, token ids out; every computation happens in the CUDA library."""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import B200Error  # noqa: F401  (re-export)
from .weights import ModelDims

_NP_DTYPE = {np.dtype(np.float16): _lib.DTYPE_F16, np.dtype(np.int32): _lib.DTYPE_I32,
             np.dtype(np.float32): _lib.DTYPE_F32, np.dtype(np.uint16): _lib.DTYPE_BF16}   # uint16 = bf16 bit patterns


def rope_table(dims: ModelDims, n_pos: int):
    """cos/sin fp32 [n_pos, head_dim/2] with the exact fp32 formula of the HF/vLLM rotary cache."""
    import torch
    d = dims.head_dim
    inv_freq = 1.0 / (dims.rope_theta ** (torch.arange(0, d, 2, dtype=torch.float32) / d))
    ang = torch.arange(n_pos, dtype=torch.float32)[:, None] * inv_freq[None, :]
    return ang.cos().numpy(), ang.sin().numpy()


class Engine:
    def __init__(self, dims: ModelDims, *, device=0, max_seqs=256, max_model_len=2048, max_batch_tokens=8192,
                 kv_bytes=0, eos_token_id=-1, use_cuda_graphs=True, gemm_path=0, tp_size=1, tp_rank=0):
        self._lib = _lib.load()
        self.dims = dims
        cfg = _lib.Config()
        cfg.struct_size = C.sizeof(_lib.Config)
        cfg.device = device
        cfg.hidden, cfg.intermediate, cfg.n_layers = dims.hidden, dims.intermediate, dims.n_layers
        cfg.n_heads, cfg.n_kv_heads, cfg.head_dim, cfg.vocab = dims.n_heads, dims.n_kv_heads, dims.head_dim, dims.vocab
        cfg.group_size, cfg.tie_embeddings = dims.group_size, int(dims.tie_embeddings)
        cfg.rms_eps, cfg.rope_theta = dims.rms_eps, dims.rope_theta
        cfg.max_seqs, cfg.max_model_len, cfg.max_batch_tokens = max_seqs, max_model_len, max_batch_tokens
        cfg.kv_bytes, cfg.eos_token_id = kv_bytes, eos_token_id
        cfg.use_cuda_graphs, cfg.gemm_path = int(use_cuda_graphs), gemm_path
        cfg.act_dtype = _lib.DTYPE_BF16 if getattr(dims, "dtype", "fp16") == "bf16" else _lib.DTYPE_F16
        cfg.dense_weights = int(getattr(dims, "quant", "awq") == "none")
        cfg.tp_size, cfg.tp_rank = tp_size, tp_rank
        self.cfg = cfg
        h = C.c_void_p()
        _lib.check(self._lib.b200_engine_create(C.byref(cfg), C.byref(h)))
        self._h = h
        self.max_model_len = max_model_len

    # -- loading ---------------------------------------------------------------------------
    def load_tensor(self, name: str, arr: np.ndarray):
        arr = np.ascontiguousarray(arr)
        shape = (C.c_int64 * arr.ndim)(*arr.shape)
        _lib.check(self._lib.b200_engine_load_tensor(self._h, name.encode(), arr.ctypes.data_as(C.c_void_p),
                                                     _NP_DTYPE[arr.dtype], arr.ndim, shape))

    def load_checkpoint(self, ckpt: dict, finalize=True):
        cos, sin = rope_table(self.dims, self.max_model_len + 1)
        self.load_tensor("b200.rope_cos", cos)
        self.load_tensor("b200.rope_sin", sin)
        for name, arr in ckpt.items():
            if name == "lm_head.weight" and self.dims.tie_embeddings:
                continue
            self.load_tensor(name, arr)
        if finalize:
            _lib.check(self._lib.b200_engine_finalize(self._h))
        return self

    # -- requests --------------------------------------------------------------------------
    def submit(self, prompt_ids, temperature=0.0, top_p=1.0, max_tokens=16, seed=0) -> int:
        ids = np.ascontiguousarray(prompt_ids, dtype=np.int32)
        rid = self._lib.b200_submit(self._h, ids.ctypes.data_as(C.c_void_p), len(ids), float(temperature),
                                    float(top_p), int(max_tokens), int(seed) & 0xFFFFFFFFFFFFFFFF)
        _lib.check(rid)
        return rid

    def wait(self, rid: int, max_tokens: int):
        out = np.empty(max_tokens, dtype=np.int32)
        n, ttft, gen = C.c_int32(0), C.c_double(0), C.c_double(0)
        _lib.check(self._lib.b200_wait(self._h, rid, out.ctypes.data_as(C.c_void_p), max_tokens, C.byref(n),
                                       C.byref(ttft), C.byref(gen)))
        return out[:min(n.value, max_tokens)], ttft.value, gen.value

    def poll(self, rid: int, cap: int = 64, timeout_ms: int = 0):
        out = np.empty(cap, dtype=np.int32)
        n, fin = C.c_int32(0), C.c_int32(0)
        _lib.check(self._lib.b200_poll(self._h, rid, out.ctypes.data_as(C.c_void_p), cap, C.byref(n), C.byref(fin),
                                       timeout_ms))
        return out[:n.value], bool(fin.value)

    def abort(self, rid: int):
        self._lib.b200_abort(self._h, rid)

    def status(self):
        """(True, "") while the engine can serve; (False, reason) once it is poisoned or closed."""
        if getattr(self, "_h", None) is None or not self._h:
            return False, "engine closed"
        code = self._lib.b200_engine_status(self._h)
        return (True, "") if code == 0 else (False, self._lib.b200_last_error().decode("utf-8", "replace"))

    def stats(self) -> dict:
        s = _lib.Stats()
        _lib.check(self._lib.b200_stats_get(self._h, C.byref(s)))
        return s.to_dict()

    def generate_batch(self, prompt_ids: np.ndarray, max_tokens: int, n_logit_steps: int = 0):
        """Greedy, equal-length prompts [n_seqs, n_prompt] -> (tokens [n_seqs, max_tokens],
        logits [n_seqs, n_logit_steps, V] or None, (prefill_s, decode_s) GPU-event seconds)."""
        ids = np.ascontiguousarray(prompt_ids, dtype=np.int32)
        n_seqs, n_prompt = ids.shape
        out = np.empty((n_seqs, max_tokens), dtype=np.int32)
        logits = np.empty((n_seqs, n_logit_steps, self.dims.vocab), dtype=np.float32) if n_logit_steps else None
        secs = (C.c_double * 2)()
        _lib.check(self._lib.b200_generate_batch(
            self._h, ids.ctypes.data_as(C.c_void_p), n_seqs, n_prompt, max_tokens, out.ctypes.data_as(C.c_void_p),
            logits.ctypes.data_as(C.c_void_p) if logits is not None else None, n_logit_steps, secs))
        return out, logits, (secs[0], secs[1])

    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            self._lib.b200_engine_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class TPEngine:
    """A tensor-parallel engine: one rank engine per device (include/b200gate.h, "Tensor parallelism"), the same
    request surface as :class:`Engine` (requests go to rank 0).  `dims` and `ckpt` are the WHOLE model's."""

    def __init__(self, dims: ModelDims, ckpt: dict, *, devices=(0, 1), **kw):
        import os
        import threading
        from .weights import shard_checkpoint, shard_dims
        lib = _lib.load()
        tp = len(devices)
        here = os.path.dirname(os.path.abspath(__file__))
        if "VGATE_B200_NCCL_LIB" not in os.environ:          # the NCCL that ships with PyTorch in this image
            import importlib.util
            spec = importlib.util.find_spec("nvidia.nccl")
            if spec and spec.submodule_search_locations:
                cand = os.path.join(list(spec.submodule_search_locations)[0], "lib", "libnccl.so.2")
                if os.path.exists(cand):
                    os.environ["VGATE_B200_NCCL_LIB"] = cand
        uid = C.create_string_buffer(128)
        _lib.check(lib.b200_tp_unique_id(uid))
        rdims = shard_dims(dims, tp)
        self.dims, self.ranks = dims, []
        for r, dev in enumerate(devices):
            e = Engine(rdims, device=dev, tp_size=tp, tp_rank=r, **kw)
            e.load_checkpoint(shard_checkpoint(dims, ckpt, tp, r), finalize=False)
            self.ranks.append(e)
        errs = [None] * tp

        def init(r):                                         # ncclCommInitRank blocks until every rank has called
            try:
                _lib.check(lib.b200_engine_tp_init(self.ranks[r]._h, uid))
                _lib.check(lib.b200_engine_finalize(self.ranks[r]._h))
            except Exception as exc:  # noqa: BLE001
                errs[r] = exc
        ts = [threading.Thread(target=init, args=(r,)) for r in range(tp)]
        for t in ts:
            t.start()
        for t in ts:
            t.join()
        for exc in errs:
            if exc is not None:
                raise exc
        for r in range(1, tp):
            _lib.check(lib.b200_engine_tp_attach(self.ranks[0]._h, self.ranks[r]._h))
        lead = self.ranks[0]
        self.submit, self.wait, self.poll, self.abort, self.stats, self.status = (lead.submit, lead.wait, lead.poll, lead.abort,
                                                                                  lead.stats, lead.status)

    def close(self):
        for e in self.ranks:           # the leader first: it stops posting work to the followers
            e.close()
        self.ranks = []

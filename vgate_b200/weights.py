"""Checkpoints for the b200 worker: model dimensions, a seeded synthetic
AWQ-shaped checkpoint (there are no weights offline) and a safetensors loader
for a mounted HF model directory.

Tensor names and shapes are the HF / AutoAWQ ones that vLLM's loader consumes
(vLLM ``model_executor/layers/quantization/awq.py:214-253``):
``qweight int32 [K, N/8]``, ``qzeros int32 [K/G, N/8]``, ``scales fp16 [K/G, N]``.
"""
from dataclasses import dataclass, asdict
import json
import os
import numpy as np


@dataclass
class ModelDims:
    hidden: int = 1536
    intermediate: int = 8960
    n_layers: int = 28
    n_heads: int = 12
    n_kv_heads: int = 2
    head_dim: int = 128
    vocab: int = 151936
    rms_eps: float = 1e-6
    rope_theta: float = 1e6
    group_size: int = 128
    tie_embeddings: bool = True
    quant: str = "awq"         # "awq": int4 linears (qweight / qzeros / scales) | "none": dense 16-bit linears (.weight)
    dtype: str = "fp16"        # 16-bit storage of activations and non-quantised tensors: "fp16" | "bf16" (dense only)

    def to_dict(self):
        """The architecture fields (what the oracle's Qwen2Dims takes)."""
        d = asdict(self)
        d.pop("quant"); d.pop("dtype")
        return d


PEAKED_EMBED_STD = 1.5       # embedding scale of the ``peaked`` synthetic profile (see synthetic_checkpoint)

# Public HF config.json values.
QWEN25_1P5B = ModelDims()
PRESETS = {
    "qwen2.5-1.5b-awq": QWEN25_1P5B,
    # the same decoder with an untied head: the parity checkpoint (profile "peaked") points the head at a
    # permutation of the embedding, so greedy decoding walks the vocabulary instead of repeating a token
    "qwen2.5-1.5b-awq-untied": ModelDims(tie_embeddings=False),
    # reduced-depth / reduced-vocab shapes used by parity tests (same per-layer kernels)
    "qwen2.5-1.5b-awq-l4": ModelDims(n_layers=4),
    "qwen2.5-1.5b-awq-l2-v8k": ModelDims(n_layers=2, vocab=8192),
    "qwen2.5-1.5b-awq-l2-v8k-untied": ModelDims(n_layers=2, vocab=8192, tie_embeddings=False),
    "qwen2.5-1.5b-awq-l4-untied": ModelDims(n_layers=4, tie_embeddings=False),
    "tiny-awq": ModelDims(hidden=256, intermediate=512, n_layers=2, n_heads=2, n_kv_heads=1,
                          head_dim=128, vocab=1024),
    # Config #4 / #5 (public HF config.json values): dense bf16 checkpoints
    "qwen2.5-7b-bf16": ModelDims(hidden=3584, intermediate=18944, n_layers=28, n_heads=28, n_kv_heads=4, vocab=152064,
                                 tie_embeddings=False, quant="none", dtype="bf16"),
    "qwen2.5-7b-bf16-l2": ModelDims(hidden=3584, intermediate=18944, n_layers=2, n_heads=28, n_kv_heads=4, vocab=152064,
                                    tie_embeddings=False, quant="none", dtype="bf16"),
    "qwen2.5-14b-bf16": ModelDims(hidden=5120, intermediate=13824, n_layers=48, n_heads=40, n_kv_heads=8, vocab=152064,
                                  tie_embeddings=False, quant="none", dtype="bf16"),
    "tiny-bf16": ModelDims(hidden=256, intermediate=512, n_layers=2, n_heads=4, n_kv_heads=2, head_dim=128, vocab=1024,
                           tie_embeddings=False, quant="none", dtype="bf16"),
    "tiny-f16-dense": ModelDims(hidden=256, intermediate=512, n_layers=2, n_heads=2, n_kv_heads=1, head_dim=128, vocab=1024,
                                quant="none", dtype="fp16"),
}


def to_bf16_bits(a) -> np.ndarray:
    """fp32 array -> uint16 array of bfloat16 bit patterns (round to nearest even; numpy has no bfloat16)."""
    u = np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)
    return ((u + 0x7FFF + ((u >> 16) & 1)) >> 16).astype(np.uint16)


def from_bf16_bits(a) -> np.ndarray:
    return (np.asarray(a, dtype=np.uint16).astype(np.uint32) << 16).view(np.float32)


def linear_shapes(d: ModelDims):
    """(name, K, N, has_bias) for the seven AWQ linears of one decoder layer."""
    H, I, hd = d.hidden, d.intermediate, d.head_dim
    return [
        ("self_attn.q_proj", H, d.n_heads * hd, True),
        ("self_attn.k_proj", H, d.n_kv_heads * hd, True),
        ("self_attn.v_proj", H, d.n_kv_heads * hd, True),
        ("self_attn.o_proj", d.n_heads * hd, H, False),
        ("mlp.gate_proj", H, I, False),
        ("mlp.up_proj", H, I, False),
        ("mlp.down_proj", I, H, False),
    ]


def next_token_permutation(vocab: int, seed: int = 0) -> np.ndarray:
    """The ``peaked`` profile's successor map: token t is followed by ``perm[t]`` (one seeded random
    permutation of the vocabulary, so greedy decoding walks its cycles)."""
    return np.random.default_rng(seed + 0x5EED).permutation(vocab).astype(np.int64)


def synthetic_checkpoint(d: ModelDims, seed: int = 0, w_std: float = 0.02, profile: str = "flat") -> dict:
    """Seeded random-init AWQ checkpoint (numpy arrays, HF tensor names).

    int4 nibbles and zero points are uniform (a uniform random int32 word is
    eight uniform nibbles in any packing order); scales are chosen so that the
    dequantised weights have standard deviation ~``w_std``: (q - z) for uniform
    q, z in [0, 15] has std sqrt(2 * 255 / 12) = 6.52.

    ``profile``:
      ``flat``    every tensor N(0, w_std^2): the residual stream is all branch output, the logits are
                  151 936 exchangeable Gaussians (sigma 0.78) and the top-2 margin is below 1e-2 at a few
                  percent of the steps -- no two correct implementations agree on 32 greedy tokens.  Stress.
      ``peaked``  the logit statistics of a trained model (one candidate far ahead of the field) with the
                  same decoder layers: embedding rows have unit scale (PEAKED_EMBED_STD), so the token's own
                  embedding stays a sizeable part of the residual stream next to the 28 layers' output
                  (|branches| / |embedding| ~ 2.7: the layers still carry ~88 % of its energy), and the final
                  norm gain is w_std / PEAKED_EMBED_STD so the logits keep the
                  flat profile's scale (field sigma 0.78, hence the same absolute error budget).  The head
                  row aligned with the residual-carried embedding leads the field's maximum by several
                  units.  With untied embeddings the head is the embedding permuted by
                  ``next_token_permutation`` (greedy decoding walks a permutation of the vocabulary);
                  with tied embeddings the leading candidate is the input token itself.
    """
    assert profile in ("flat", "peaked"), profile
    rng = np.random.default_rng(seed)
    G = d.group_size
    ck = {}

    def f16(a):                          # the checkpoint's 16-bit type: fp16 arrays, or uint16 arrays of bf16 bits
        return to_bf16_bits(a) if d.dtype == "bf16" else np.asarray(a, dtype=np.float16)

    peaked = profile == "peaked"
    e_std = PEAKED_EMBED_STD if peaked else w_std
    ck["model.embed_tokens.weight"] = f16(rng.standard_normal((d.vocab, d.hidden), dtype=np.float32) * e_std)
    if not d.tie_embeddings:
        if peaked:
            head = np.empty_like(ck["model.embed_tokens.weight"])
            head[next_token_permutation(d.vocab, seed)] = ck["model.embed_tokens.weight"]     # head[perm[t]] = embed[t]
            ck["lm_head.weight"] = head
        else:
            ck["lm_head.weight"] = f16(rng.standard_normal((d.vocab, d.hidden), dtype=np.float32) * w_std)
    g_final = (w_std / PEAKED_EMBED_STD) if peaked else 1.0
    ck["model.norm.weight"] = f16(g_final * (1.0 + 0.1 * rng.standard_normal(d.hidden, dtype=np.float32)))
    for i in range(d.n_layers):
        p = f"model.layers.{i}."
        ck[p + "input_layernorm.weight"] = f16(1.0 + 0.1 * rng.standard_normal(d.hidden, dtype=np.float32))
        ck[p + "post_attention_layernorm.weight"] = f16(1.0 + 0.1 * rng.standard_normal(d.hidden, dtype=np.float32))
        for name, K, N, has_bias in linear_shapes(d):
            if d.quant == "none":            # dense: HF layout weight[out, in]
                ck[p + name + ".weight"] = f16(rng.standard_normal((N, K), dtype=np.float32) * w_std)
                if has_bias:
                    ck[p + name + ".bias"] = f16(rng.standard_normal(N, dtype=np.float32) * w_std)
                continue
            assert K % G == 0 and N % 8 == 0, (name, K, N)
            ck[p + name + ".qweight"] = rng.integers(0, 2**32, size=(K, N // 8), dtype=np.uint32).view(np.int32)
            ck[p + name + ".qzeros"] = rng.integers(0, 2**32, size=(K // G, N // 8), dtype=np.uint32).view(np.int32)
            s = (w_std / 6.52) * (0.5 + rng.random((K // G, N), dtype=np.float32))
            ck[p + name + ".scales"] = f16(s)
            if has_bias:
                ck[p + name + ".bias"] = f16(rng.standard_normal(N, dtype=np.float32) * w_std)
    return ck


def load_hf_checkpoint(model_dir: str):
    """(dims, checkpoint dict) from a mounted HF model directory: AWQ int4 (quantization_config present) or a
    dense fp16 / bf16 checkpoint (``torch_dtype``).  bf16 tensors come back as uint16 bit patterns."""
    with open(os.path.join(model_dir, "config.json")) as f:
        c = json.load(f)
    q = c.get("quantization_config")
    if q is not None and (q.get("quant_method", "awq") != "awq" or q.get("bits", 4) != 4):
        raise ValueError(f"unsupported quantization_config {q}")
    dtype = "fp16" if q is not None else {"bfloat16": "bf16", "float16": "fp16"}.get(str(c.get("torch_dtype", "bfloat16")))
    if dtype is None:
        raise ValueError(f"unsupported torch_dtype {c.get('torch_dtype')}")
    q = q or {}
    d = ModelDims(quant="awq" if c.get("quantization_config") is not None else "none", dtype=dtype,
                  
        hidden=c["hidden_size"], intermediate=c["intermediate_size"], n_layers=c["num_hidden_layers"],
        n_heads=c["num_attention_heads"], n_kv_heads=c["num_key_value_heads"],
        head_dim=c.get("head_dim") or c["hidden_size"] // c["num_attention_heads"],
        vocab=c["vocab_size"], rms_eps=c.get("rms_norm_eps", 1e-6), rope_theta=c.get("rope_theta", 1e6),
        group_size=q.get("group_size", 128), tie_embeddings=c.get("tie_word_embeddings", False))
    ck = {}
    for fn in sorted(os.listdir(model_dir)):
        if fn.endswith(".safetensors"):
            ck.update(_load_safetensors(os.path.join(model_dir, fn)))
    return d, ck


def _load_safetensors(path: str) -> dict:
    """safetensors -> numpy; BF16 tensors (which numpy cannot represent) as uint16 bit patterns."""
    import struct
    with open(path, "rb") as f:
        n = struct.unpack("<Q", f.read(8))[0]
        header = json.loads(f.read(n))
        base = 8 + n
        out = {}
        np_dt = {"F16": np.float16, "BF16": np.uint16, "F32": np.float32, "I32": np.int32, "I64": np.int64, "U8": np.uint8}
        for name, meta in header.items():
            if name == "__metadata__":
                continue
            a, b = meta["data_offsets"]
            f.seek(base + a)
            out[name] = np.frombuffer(f.read(b - a), dtype=np_dt[meta["dtype"]]).reshape(meta["shape"])
    return out


def save_safetensors(ckpt: dict, path: str, bf16: bool = False) -> None:
    """numpy checkpoint -> one safetensors file (uint16 arrays are written as BF16 when ``bf16``)."""
    names = {np.dtype(np.float16): "F16", np.dtype(np.float32): "F32", np.dtype(np.int32): "I32", np.dtype(np.int64): "I64",
             np.dtype(np.uint8): "U8", np.dtype(np.uint16): "BF16" if bf16 else "U16"}
    header, blobs, off = {}, [], 0
    for name, arr in ckpt.items():
        arr = np.ascontiguousarray(arr)
        header[name] = {"dtype": names[arr.dtype], "shape": list(arr.shape), "data_offsets": [off, off + arr.nbytes]}
        blobs.append(arr.tobytes())
        off += arr.nbytes
    h = json.dumps(header).encode()
    h += b" " * (-len(h) % 8)
    import struct
    with open(path, "wb") as f:
        f.write(struct.pack("<Q", len(h)))
        f.write(h)
        for b in blobs:
            f.write(b)


def shard_dims(d: ModelDims, tp: int) -> ModelDims:
    """The dimensions ONE tensor-parallel rank is created with: its share of the heads and of the MLP's
    intermediate columns; hidden size, vocabulary and depth are whole."""
    assert d.n_heads % tp == 0 and d.n_kv_heads % tp == 0 and d.intermediate % (64 * tp) == 0, (d, tp)
    from dataclasses import replace
    return replace(d, n_heads=d.n_heads // tp, n_kv_heads=d.n_kv_heads // tp, intermediate=d.intermediate // tp)


def shard_checkpoint(d: ModelDims, ckpt: dict, tp: int, rank: int) -> dict:
    """Megatron-style partition of a dense checkpoint for rank `rank` of `tp`: q / k / v and
    gate / up column-parallel (rows of the HF [out, in] weights, and their biases), o_proj / down_proj row-parallel
    (columns); embeddings, norms and LM head replicated."""
    assert d.quant == "none", "tensor parallelism is built for dense checkpoints"
    hd = d.head_dim
    out = {}
    for name, a in ckpt.items():
        if ".self_attn.q_proj." in name:
            n = d.n_heads // tp * hd
            out[name] = a[rank * n:(rank + 1) * n]
        elif ".self_attn.k_proj." in name or ".self_attn.v_proj." in name:
            n = d.n_kv_heads // tp * hd
            out[name] = a[rank * n:(rank + 1) * n]
        elif ".mlp.gate_proj." in name or ".mlp.up_proj." in name:
            n = d.intermediate // tp
            out[name] = a[rank * n:(rank + 1) * n]
        elif name.endswith(".self_attn.o_proj.weight"):
            n = d.n_heads // tp * hd
            out[name] = np.ascontiguousarray(a[:, rank * n:(rank + 1) * n])
        elif name.endswith(".mlp.down_proj.weight"):
            n = d.intermediate // tp
            out[name] = np.ascontiguousarray(a[:, rank * n:(rank + 1) * n])
        else:
            out[name] = a
    return out

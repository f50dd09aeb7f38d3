"""Generate the golden vectors that pin ``oracle/`` (run in the build container).

The reference tree holds no numerical fixtures for this path,
so the oracle is pinned against the real third-party code the reference
delegates to, imported here on CPU:

  awq_pack.npz     vLLM's own ``awq_pack`` (quant_utils.py:794-813): the oracle's
                   unpack must invert it bit-exactly.
  top_p.npz        vLLM's ``apply_top_k_top_p_pytorch`` (topk_topp_sampler.py:328-369).
  hf_qwen2.npz     ``transformers.Qwen2ForCausalLM`` fp32 eager logits + greedy tokens
                   on a small AWQ-shaped checkpoint (same maths as vLLM qwen2.py).

Usage:  python tests/golden/make_golden.py        (writes next to this file)
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle.awq import dequantize_awq  # noqa: E402
from vgate_b200.weights import ModelDims, synthetic_checkpoint, linear_shapes  # noqa: E402

SMALL = ModelDims(hidden=128, intermediate=256, n_layers=2, n_heads=2, n_kv_heads=1,
                  head_dim=64, vocab=256)


def make_awq_pack():
    from vllm.model_executor.layers.quantization.utils.quant_utils import awq_pack
    rng = np.random.default_rng(7)
    K, N = 256, 64
    q = rng.integers(0, 16, size=(K, N), dtype=np.int64)
    packed = awq_pack(torch.from_numpy(q).to(torch.int32), 4, K, N).numpy()
    np.savez_compressed(os.path.join(HERE, "awq_pack.npz"), q=q.astype(np.uint8), packed=packed)


def make_top_p():
    from vllm.v1.sample.ops.topk_topp_sampler import apply_top_k_top_p_pytorch
    g = torch.Generator().manual_seed(11)
    logits = torch.randn(6, 2000, generator=g) * 3.0
    p = torch.tensor([0.9, 0.5, 0.99, 0.1, 0.7, 0.95])
    masked = apply_top_k_top_p_pytorch(logits.clone(), None, p)
    np.savez_compressed(os.path.join(HERE, "top_p.npz"), logits=logits.numpy(), p=p.numpy(),
                        keep=torch.isfinite(masked).numpy())


def make_hf_qwen2():
    from transformers import Qwen2Config, Qwen2ForCausalLM
    d = SMALL
    ck = synthetic_checkpoint(d, seed=3)
    cfg = Qwen2Config(vocab_size=d.vocab, hidden_size=d.hidden, intermediate_size=d.intermediate,
                      num_hidden_layers=d.n_layers, num_attention_heads=d.n_heads,
                      num_key_value_heads=d.n_kv_heads, rms_norm_eps=d.rms_eps, rope_theta=d.rope_theta,
                      tie_word_embeddings=True, max_position_embeddings=512, attention_dropout=0.0,
                      use_sliding_window=False)
    cfg._attn_implementation = "eager"
    model = Qwen2ForCausalLM(cfg).float().eval()
    sd = {}
    sd["model.embed_tokens.weight"] = torch.from_numpy(ck["model.embed_tokens.weight"]).float()
    sd["lm_head.weight"] = sd["model.embed_tokens.weight"]
    sd["model.norm.weight"] = torch.from_numpy(ck["model.norm.weight"]).float()
    for i in range(d.n_layers):
        p = f"model.layers.{i}."
        for n in ("input_layernorm.weight", "post_attention_layernorm.weight"):
            sd[p + n] = torch.from_numpy(ck[p + n]).float()
        for name, K, N, has_bias in linear_shapes(d):
            w = dequantize_awq(ck[p + name + ".qweight"], ck[p + name + ".qzeros"], ck[p + name + ".scales"],
                               d.group_size)
            sd[p + name + ".weight"] = torch.from_numpy(w).float().t().contiguous()   # HF: [out, in]
            if has_bias:
                sd[p + name + ".bias"] = torch.from_numpy(ck[p + name + ".bias"]).float()
    missing, unexpected = model.load_state_dict(sd, strict=False)
    assert not unexpected, unexpected
    assert all("rotary" in m or "inv_freq" in m for m in missing), missing
    rng = np.random.default_rng(5)
    ids = rng.integers(0, d.vocab, size=(3, 12), dtype=np.int64)
    with torch.no_grad():
        logits = model(torch.from_numpy(ids)).logits.float().numpy()
        gen = model.generate(torch.from_numpy(ids), do_sample=False, max_new_tokens=10, min_new_tokens=10,
                             eos_token_id=None, pad_token_id=0).numpy()[:, ids.shape[1]:]
    np.savez_compressed(os.path.join(HERE, "hf_qwen2.npz"), ids=ids, logits=logits, greedy=gen.astype(np.int32),
                        transformers_version=np.array(__import__("transformers").__version__))


if __name__ == "__main__":
    make_awq_pack()
    make_top_p()
    make_hf_qwen2()
    for f in ("awq_pack.npz", "top_p.npz", "hf_qwen2.npz"):
        print(f, os.path.getsize(os.path.join(HERE, f)), "bytes")

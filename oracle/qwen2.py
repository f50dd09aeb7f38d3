"""Qwen2 decoder forward on CPU -- restatement (test infrastructure only).

Follows vLLM 0.22.0 ``vllm/model_executor/models/qwen2.py``:
  MLP ``:83-118`` (gate_up -> SiLU(gate)*up -> down), attention ``:120-236``
  (qkv with bias, neox RoPE theta=1e6, GQA causal attention, o_proj),
  decoder layer ``:239-312`` (pre-norm residual blocks), model ``:389-421``
  (embed -> layers -> final RMSNorm), LM head ``:494-565`` (tied embedding).
The same mathematics as ``transformers.Qwen2ForCausalLM`` -- which is what
``tests/golden/make_golden.py`` pins it against.

Arithmetic modes:
  ``fp32``  every intermediate in fp32.  THE oracle: it is the mode the golden vectors pin
            (``tests/golden/hf_qwen2.npz``), and the only one a parity gate may use -- logits within
            2e-2, greedy tokens identical.
  ``fp16``  diagnostic only, pinned by nothing and gating nothing: fp32 accumulation with the
            activations rounded to fp16 where the CUDA path stores fp16 (normed activations, q/k/v after
            RoPE, attention output, SwiGLU output).  Useful to tell rounding-point effects from
            summation-order effects when a logit differs.
  ``bf16r`` diagnostic, like ``fp16`` with bfloat16 rounding points (the dense bf16 engine path).
  ``bf16``  timing only (``bench.py`` CPU legs): weights and matmul operands in bfloat16, the eager
            CPU arithmetic a PyTorch user of this model would run on the host.
"""
from dataclasses import dataclass
import numpy as np
import torch

from .awq import dequantize_awq


@dataclass
class Qwen2Dims:
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


def _t(a):
    """numpy -> torch; uint16 arrays are bfloat16 bit patterns (dense bf16 checkpoints)."""
    a = np.ascontiguousarray(a)
    if a.dtype == np.uint16:
        return torch.from_numpy(a.view(np.int16)).view(torch.bfloat16)
    return torch.from_numpy(a)


def rope_table(dims: Qwen2Dims, max_pos: int):
    """cos/sin fp32 [max_pos, head_dim/2]; HF/vLLM formula (rotary_embedding
    ``_compute_inv_freq`` / ``_compute_cos_sin_cache``), kept in fp32."""
    d = dims.head_dim
    inv_freq = 1.0 / (dims.rope_theta ** (torch.arange(0, d, 2, dtype=torch.float32) / d))
    ang = torch.arange(max_pos, dtype=torch.float32)[:, None] * inv_freq[None, :]
    return ang.cos(), ang.sin()


class OracleQwen2:
    """Batched (equal-length) forward with a dense per-sequence KV cache."""

    def __init__(self, dims: Qwen2Dims, ckpt: dict, mode: str = "fp32", max_pos: int = 4096):
        assert mode in ("fp32", "fp16", "bf16", "bf16r")
        self.d, self.mode = dims, mode
        wdt = torch.bfloat16 if mode == "bf16" else torch.float32
        self.cos, self.sin = rope_table(dims, max_pos)
        g = dims.group_size

        def lin(prefix):
            if prefix + ".qweight" in ckpt:
                w = dequantize_awq(ckpt[prefix + ".qweight"], ckpt[prefix + ".qzeros"],
                                   ckpt[prefix + ".scales"], g)      # fp16 [K, N]
                return _t(w).to(wdt)
            return _t(ckpt[prefix + ".weight"]).to(wdt).t().contiguous()  # HF [N, K] -> [K, N]

        def vec(name):
            return _t(ckpt[name]).float()

        self.embed = vec("model.embed_tokens.weight")
        self.lm_head = (self.embed if dims.tie_embeddings else vec("lm_head.weight")).to(wdt)
        self.final_norm = vec("model.norm.weight")
        self.layers = []
        for i in range(dims.n_layers):
            p = f"model.layers.{i}."
            L = {
                "ln1": vec(p + "input_layernorm.weight"),
                "ln2": vec(p + "post_attention_layernorm.weight"),
                "wqkv": torch.cat([lin(p + f"self_attn.{n}_proj") for n in "qkv"], dim=1),
                "bqkv": torch.cat([vec(p + f"self_attn.{n}_proj.bias") for n in "qkv"]),
                "wo": lin(p + "self_attn.o_proj"),
                "wgu": torch.cat([lin(p + "mlp.gate_proj"), lin(p + "mlp.up_proj")], dim=1),
                "wd": lin(p + "mlp.down_proj"),
            }
            self.layers.append(L)
        self.reset()

    # -- helpers ---------------------------------------------------------
    def reset(self):
        self.k_cache = [None] * self.d.n_layers
        self.v_cache = [None] * self.d.n_layers
        self.n_cached = 0

    def _rnd(self, x):
        if self.mode == "bf16r":
            return x.bfloat16().float()
        return x.half().float() if self.mode == "fp16" else x

    def _mm(self, x, w):
        """x @ w with fp32 result; bf16 mode rounds the left operand to the weights' dtype first."""
        return (x.to(w.dtype) @ w).float() if w.dtype != torch.float32 else x @ w

    def _norm(self, h, w):
        # qwen2.py:287-312 RMSNorm; fp32 throughout, one rounding at the end
        var = h.pow(2).mean(-1, keepdim=True)
        return self._rnd(h * torch.rsqrt(var + self.d.rms_eps) * w)

    def _rope(self, x, pos):
        # neox style (qwen2.py:181 get_rope): rotate halves
        half = self.d.head_dim // 2
        c = self.cos[pos][None, :, None, :]
        s = self.sin[pos][None, :, None, :]
        x1, x2 = x[..., :half], x[..., half:]
        return torch.cat([x1 * c - x2 * s, x2 * c + x1 * s], dim=-1)

    # -- forward ---------------------------------------------------------
    @torch.no_grad()
    def forward(self, ids: np.ndarray, return_hidden=False, last_only=False) -> torch.Tensor:
        """ids int [B, T] appended after the cached prefix -> logits fp32 [B, T, V]
        ([B, 1, V] with ``last_only``: the LM head is applied to the last position alone, as an engine
        does for a prompt -- vLLM qwen2.py:553-565 compute_logits on the sampled rows)."""
        d = self.d
        ids_t = _t(np.asarray(ids, dtype=np.int64))
        B, T = ids_t.shape
        pos = torch.arange(self.n_cached, self.n_cached + T)
        h = self.embed[ids_t]                                   # [B, T, H] fp32
        for li, L in enumerate(self.layers):
            xn = self._norm(h, L["ln1"])
            qkv = self._mm(xn, L["wqkv"]) + L["bqkv"]
            q, k, v = qkv.split([d.n_heads * d.head_dim, d.n_kv_heads * d.head_dim,
                                 d.n_kv_heads * d.head_dim], dim=-1)
            q = self._rnd(self._rope(q.view(B, T, d.n_heads, d.head_dim), pos))
            k = self._rnd(self._rope(k.view(B, T, d.n_kv_heads, d.head_dim), pos))
            v = self._rnd(v.reshape(B, T, d.n_kv_heads, d.head_dim))
            if self.k_cache[li] is not None:
                k = torch.cat([self.k_cache[li], k], dim=1)
                v = torch.cat([self.v_cache[li], v], dim=1)
            self.k_cache[li], self.v_cache[li] = k, v
            S = k.shape[1]
            grp = d.n_heads // d.n_kv_heads
            qh = q.view(B, T, d.n_kv_heads, grp, d.head_dim)
            sc = torch.einsum("btkgd,bskd->bkgts", qh, k) / (d.head_dim ** 0.5)
            causal = torch.arange(S)[None, :] > (pos[:, None])          # [T, S]
            sc = sc.masked_fill(causal[None, None, None], float("-inf"))
            p = torch.softmax(sc, dim=-1)
            ao = torch.einsum("bkgts,bskd->btkgd", p, v).reshape(B, T, d.n_heads * d.head_dim)
            ao = self._rnd(ao)
            h = h + self._mm(ao, L["wo"])
            xn = self._norm(h, L["ln2"])
            gu = self._mm(xn, L["wgu"])
            gte, up = gu.split(d.intermediate, dim=-1)
            act = self._rnd(torch.nn.functional.silu(gte) * up)
            h = h + self._mm(act, L["wd"])
        self.n_cached += T
        if last_only:
            h = h[:, -1:]
        xn = self._norm(h, self.final_norm)
        if return_hidden:
            return xn
        return self._mm(xn, self.lm_head.t())

    @torch.no_grad()
    def generate_greedy(self, prompt_ids: np.ndarray, max_new_tokens: int, return_logits=False, n_logit_steps=None):
        """prompt_ids int [B, P] -> int32 [B, max_new_tokens] (EOS disabled); with ``return_logits`` also the
        fp32 logits [B, n_logit_steps or max_new_tokens, V] of the first generated positions."""
        self.reset()
        out, all_logits = [], []
        n_keep = max_new_tokens if n_logit_steps is None else n_logit_steps
        logits = self.forward(prompt_ids, last_only=True)[:, -1]
        for _ in range(max_new_tokens):
            nxt = logits.argmax(-1)
            out.append(nxt.numpy().astype(np.int32))
            if return_logits and len(all_logits) < n_keep:
                all_logits.append(logits.numpy().copy())
            if len(out) == max_new_tokens:
                break
            logits = self.forward(nxt[:, None].numpy())[:, -1]
        toks = np.stack(out, axis=1)
        if return_logits:
            return toks, np.stack(all_logits, axis=1)
        return toks

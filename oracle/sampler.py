"""Sampling -- CPU restatement (test infrastructure only).

Follows vLLM 0.22.0:
  ``vllm/v1/sample/sampler.py:223-236``  temperature; greedy when T < 1e-5
  ``vllm/v1/sample/ops/topk_topp_sampler.py:349-369``  top-p: ascending sort,
      softmax, cumsum, drop ``cumsum <= 1 - p``, always keep the largest
  ``:395-416``  draw = argmax(probs / q), q ~ Exp(1)  (exponential race)

vLLM draws q from torch's global generator, which no other implementation can
reproduce; the b200 path (and this oracle) use a counter-based hash instead so
that a draw is a pure function of (seed, step, token index).
"""
import numpy as np

SAMPLING_EPS = 1e-5          # sampler.py:17


def hash_uniform(seed: int, step: int, n: int) -> np.ndarray:
    """u in (0,1), fp32, one per vocabulary index (lowbias32 integer hash)."""
    i = np.arange(n, dtype=np.uint64)
    x = (i * 0x9E3779B1 + (np.uint64(step & 0xFFFFFFFF) * np.uint64(0x85EBCA77))
         + (np.uint64(seed & 0xFFFFFFFF) * np.uint64(0xC2B2AE3D))) & 0xFFFFFFFF
    x = x.astype(np.uint64)
    x ^= x >> 16
    x = (x * 0x7FEB352D) & 0xFFFFFFFF
    x ^= x >> 15
    x = (x * 0x846CA68B) & 0xFFFFFFFF
    x ^= x >> 16
    return uniform_from_bits(x)


def uniform_from_bits(x) -> np.ndarray:
    """32 hash bits -> fp32 strictly inside (0, 1): 23 bits, so that ``(x >> 9) + 0.5`` is exact in
    fp32 (24 bits would round 16777215.5 up to 2**24 and give u == 1.0, i.e. an Exp(1) draw of 0)."""
    x = np.asarray(x, dtype=np.uint64)
    return ((x >> 9).astype(np.float32) + np.float32(0.5)) * np.float32(1.0 / 8388608.0)


def top_p_keep_mask(logits: np.ndarray, top_p: float) -> np.ndarray:
    """bool [V]: tokens that survive top-p (topk_topp_sampler.py:357-366)."""
    l = np.asarray(logits, dtype=np.float32)
    order = np.argsort(l, kind="stable")                 # ascending
    ls = l[order]
    e = np.exp(ls - ls.max(), dtype=np.float32)
    probs = e / e.sum(dtype=np.float32)
    cs = np.cumsum(probs, dtype=np.float32)
    drop = cs <= np.float32(1.0 - top_p)
    drop[-1] = False
    keep = np.empty_like(drop)
    keep[order] = ~drop
    return keep


def sample(logits: np.ndarray, temperature: float, top_p: float, seed: int, step: int) -> int:
    l = np.asarray(logits, dtype=np.float32)
    if temperature < SAMPLING_EPS:
        return int(np.argmax(l))                          # sampler.py:235-236
    l = l / np.float32(temperature)                        # sampler.py:223-232
    keep = top_p_keep_mask(l, top_p) if top_p < 1.0 else np.ones(l.shape, bool)
    lm = np.where(keep, l, -np.inf).astype(np.float32)
    e = np.exp(lm - lm.max(), dtype=np.float32)
    probs = e / e.sum(dtype=np.float32)
    q = -np.log(hash_uniform(seed, step, l.shape[0]))      # Exp(1)
    return int(np.argmax(probs / q))                       # topk_topp_sampler.py:416

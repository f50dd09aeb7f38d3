"""AWQ int4 unpack / dequant -- CPU restatement (test infrastructure only).

Follows vLLM 0.22.0 ``model_executor/layers/quantization/awq_triton.py:11-105``
(nibble order + dequant formula) and ``awq.py:214-253`` (tensor shapes):

    qweight int32 [K, N/8]     qzeros int32 [K/G, N/8]     scales fp16 [K/G, N]

The packed word at column ``c`` holds logical output column ``8c+i`` in nibble
``AWQ_ORDER[i]`` (awq_triton.py:53-70).  Dequant is
``W[k, n] = (q[k, n] - z[k//G, n]) * s[k//G, n]`` evaluated in fp16
(awq_triton.py:98-99): the integer subtract is exact, the multiply rounds once.
"""
import numpy as np

# awq_triton.py:53-57  reverse_awq_order_tensor == [0, 4, 1, 5, 2, 6, 3, 7]
AWQ_ORDER = np.array([0, 4, 1, 5, 2, 6, 3, 7], dtype=np.uint32)


def unpack_awq(packed: np.ndarray) -> np.ndarray:
    """int32 [R, C] -> uint8 [R, 8C] nibbles in logical column order."""
    w = np.ascontiguousarray(packed).view(np.uint32)
    out = np.empty((w.shape[0], w.shape[1], 8), dtype=np.uint8)
    for i in range(8):
        out[:, :, i] = (w >> (4 * AWQ_ORDER[i])) & 0xF
    return out.reshape(w.shape[0], w.shape[1] * 8)


def pack_awq(q: np.ndarray) -> np.ndarray:
    """uint8 nibbles [R, N] -> int32 [R, N/8]; inverse of :func:`unpack_awq`."""
    q = np.asarray(q, dtype=np.uint32).reshape(q.shape[0], -1, 8)
    w = np.zeros(q.shape[:2], dtype=np.uint32)
    for i in range(8):
        w |= (q[:, :, i] & 0xF) << (4 * AWQ_ORDER[i])
    return w.view(np.int32)


def dequantize_awq(qweight, qzeros, scales, group_size=128) -> np.ndarray:
    """-> fp16 [K, N] = fp16((q - z) * s)   (awq_triton.py:98-99)."""
    q = unpack_awq(qweight)
    z = unpack_awq(qzeros)
    s = np.asarray(scales, dtype=np.float16)
    K, N = q.shape
    # Evaluated in fp32 and rounded once: q - z is a small integer and its product with an fp16 scale
    # (5 x 11 significant bits) is exact in fp32, so the single rounding to fp16 is bit-identical to the
    # reference's fp16 multiply -- and numpy's fp32 arithmetic is vectorised where its fp16 is emulated.
    d = q.reshape(K // group_size, group_size, N).astype(np.float32) - z.astype(np.float32)[:, None, :]
    d *= s.astype(np.float32)[:, None, :]
    return d.astype(np.float16).reshape(K, N)


def awq_linear(x, qweight, qzeros, scales, bias=None, group_size=128):
    """y = x @ dequant(W) (+ bias) with fp32 accumulation (awq.py:262-286)."""
    w = dequantize_awq(qweight, qzeros, scales, group_size).astype(np.float32)
    y = np.asarray(x, dtype=np.float32) @ w
    if bias is not None:
        y = y + np.asarray(bias, dtype=np.float32)
    return y

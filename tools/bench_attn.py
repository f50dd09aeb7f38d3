"""Micro-benchmark of the paged attention kernels (decode shapes) through the C ABI."""
import ctypes as C
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from vgate_b200 import _lib

lib = _lib.load()
ptr = lambda t: C.c_void_p(t.data_ptr()) if t is not None else None
st = torch.cuda.Stream()
sp = C.c_void_p(st.cuda_stream)


def bench(T, ctx, n_splits, fused, n_heads=12, n_kv=2, n_copies=8, reps=20):
    nb_seq = (ctx + 15) // 16 + 1
    n_blocks = T * nb_seq
    caches = [(torch.randn((n_blocks, n_kv, 16, 128), device="cuda").half(), torch.randn((n_blocks, n_kv, 16, 128), device="cuda").half())
              for _ in range(n_copies)]
    bt = torch.arange(T * nb_seq, dtype=torch.int32, device="cuda").reshape(T, nb_seq).contiguous()
    btrow = torch.arange(T, dtype=torch.int32, device="cuda")
    ctxd = torch.full((T,), ctx, dtype=torch.int32, device="cuda")
    pos = ctxd - 1
    slot = (bt[:, (ctx - 1) // 16] * 16 + (ctx - 1) % 16).int().contiguous()
    q = torch.randn((T, n_heads * 128), device="cuda").half()
    width = (n_heads + 2 * n_kv) * 128
    parts = torch.randn((4, T, width), device="cuda")
    bias = torch.randn(width, device="cuda").half()
    cos = torch.randn((4096, 64), device="cuda"); sin = torch.randn((4096, 64), device="cuda")
    out = torch.empty((T, n_heads * 128), dtype=torch.float16, device="cuda")
    scratch = torch.empty(max(4, lib.b200_op_attn_scratch_bytes(T, n_heads, n_kv, n_splits) // 4), dtype=torch.float32, device="cuda")
    counters = torch.zeros(T * n_kv, dtype=torch.int32, device="cuda")
    def call(i):
        kc, vc = caches[i % n_copies]
        if fused:
            _lib.check(lib.b200_op_attn_decode_fused(ptr(parts), 4, T * width, ptr(bias), ptr(pos), ptr(slot), ptr(cos), ptr(sin),
                                                     ptr(kc), ptr(vc), ptr(bt), nb_seq, ptr(btrow), ptr(ctxd), T, n_heads, n_kv,
                                                     n_splits, ptr(scratch), ptr(counters), ptr(out), sp))
        else:
            _lib.check(lib.b200_op_paged_attn(ptr(q), ptr(kc), ptr(vc), ptr(bt), nb_seq, ptr(btrow), ptr(ctxd), T, n_heads, n_kv,
                                              n_splits, ptr(scratch), ptr(counters), ptr(out), sp))

    # Capture `reps` back-to-back launches in a CUDA graph: the replay has no host launch latency in it.
    with torch.cuda.stream(st):
        call(0)
        st.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=st):
            for i in range(reps):
                call(i)
        ts = []
        for _ in range(5):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(st); g.replay(); e1.record(st); st.synchronize()
            ts.append(e0.elapsed_time(e1) * 1e3 / reps)
    kv_bytes = T * ctx * n_kv * 128 * 2 * 2
    med = float(np.median(ts))
    print(f"attn T={T} ctx={ctx} splits={n_splits} fused={int(fused)}: {med:6.1f} us  ({kv_bytes / med / 1e3:6.0f} GB/s KV)")


if __name__ == "__main__":
    print(torch.cuda.get_device_name())
    for ctx in (128, 256, 384, 1024):
        bench(64, ctx, 1, False)
        bench(64, ctx, 1, True)
    bench(8, 256, 1, True)
    bench(128, 256, 1, True)
    bench(256, 256, 1, True)

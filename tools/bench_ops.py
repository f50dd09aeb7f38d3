"""Micro-benchmark of the GEMM kernels through the C ABI (CUDA events, L2 flushed between launches)."""
import ctypes as C
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from vgate_b200 import _lib

lib = _lib.load()
ptr = lambda t: C.c_void_p(t.data_ptr()) if t is not None else None
flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")
st = torch.cuda.Stream()
sp = C.c_void_p(st.cuda_stream)


def timeit(fn, n=10, warm=3, do_flush=True):
    ts = []
    with torch.cuda.stream(st):
        for i in range(n + warm):
            if do_flush:
                flush.fill_(i)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(st); fn(); e1.record(st); st.synchronize()
            if i >= warm:
                ts.append(e0.elapsed_time(e1) * 1e3)
    return float(np.median(ts)), float(np.min(ts))


def w4(K, N, T, splits=1, G=128):
    rng = np.random.default_rng(0)
    wq = torch.from_numpy(rng.integers(0, 2**31, size=(N, K // 8), dtype=np.int32)).cuda()
    sc = torch.full((N, K // G), 0.003, dtype=torch.float16, device="cuda")
    zp = torch.full((N, K // G), 1032.0, dtype=torch.float16, device="cuda")
    ch = torch.empty(lib.b200_op_w4_tile_bytes(K, N, G), dtype=torch.uint8, device="cuda")
    _lib.check(lib.b200_op_w4_tile(ptr(wq), ptr(sc), ptr(zp), K, N, G, 0, ptr(ch), sp))
    x = torch.randn((T, K), device="cuda").half()
    splits = lib.b200_op_w4_effective_splits(K, G, splits)
    y = torch.empty((splits, T, N), dtype=torch.float32, device="cuda")
    f = lambda: _lib.check(lib.b200_op_w4_gemm_umma(ptr(ch), ptr(x), K, ptr(y), N, None, T, K, N, G, 0, splits, T * N, sp))
    med, mn = timeit(f)
    medw, mnw = timeit(f, do_flush=False)
    byts = N * K / 2 + 2 * N * (K // G) * 2
    print(f"w4 umma K={K:5d} N={N:6d} T={T:4d} splits={splits:2d}: cold {med:7.1f} us ({byts / med / 1e3:6.0f} GB/s)  "
          f"warm-L2 {medw:7.1f} us  ctas={N // 128 * splits * ((T + 255) // 256)}")
    if T <= 8:
        yg = torch.empty((T, N), dtype=torch.float32, device="cuda")
        g = lambda: _lib.check(lib.b200_op_w4_gemv(ptr(wq), ptr(sc), ptr(zp), ptr(x), K, ptr(yg), N, None, T, K, N, G, 0, sp))
        med, mn = timeit(g)
        print(f"w4 gemv K={K:5d} N={N:6d} T={T:4d}           : cold {med:7.1f} us ({byts / med / 1e3:6.0f} GB/s)")


def f16(K, N, T):
    w = torch.randn((N, K), device="cuda").half()
    x = torch.randn((T, K), device="cuda").half()
    y = torch.empty((T, N), dtype=torch.float32, device="cuda")
    f = lambda: _lib.check(lib.b200_op_f16_gemm_umma(ptr(w), ptr(x), K, ptr(y), N, T, K, N, 0, sp))
    med, mn = timeit(f)
    print(f"f16 umma K={K:5d} N={N:6d} T={T:4d}: cold {med:7.1f} us ({N * K * 2 / med / 1e3:6.0f} GB/s)")


if __name__ == "__main__":
    print(torch.cuda.get_device_name())
    for T in (1, 8, 64, 256):
        w4(1536, 17920, T)
    w4(1536, 2048, 64)
    w4(1536, 2048, 64, 4)
    w4(1536, 1536, 64)
    w4(1536, 1536, 64, 4)
    w4(8960, 1536, 64)
    w4(8960, 1536, 64, 12)
    w4(8960, 1536, 64, 35)
    w4(128, 17920, 64)
    w4(256, 17920, 64)
    w4(512, 17920, 64)
    w4(1536, 17920 * 4, 64)
    f16(1536, 151936, 64)
    f16(1536, 151936, 8)

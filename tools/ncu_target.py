"""Short single-kernel target for `ncu --set full`: a few launches of the dominant decode kernel
(tcgen05 W4A16 gate_up GEMM + SwiGLU, batch 64) over rotating weight copies."""
import ctypes as C
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from vgate_b200 import _lib

lib = _lib.load()
ptr = lambda t: C.c_void_p(t.data_ptr())
K, I, G, T = 1536, 8960, 128, 64
N = 2 * I
rng = np.random.default_rng(0)
sc = torch.full((N, K // G), 0.003, dtype=torch.float16, device="cuda")
zp = torch.full((N, K // G), 1032.0, dtype=torch.float16, device="cuda")
ws = []
for c in range(10):
    wq = torch.from_numpy(rng.integers(0, 2**31, size=(N, K // 8), dtype=np.int32)).cuda()
    ch = torch.empty(lib.b200_op_w4_tile_bytes(K, N, G), dtype=torch.uint8, device="cuda")
    _lib.check(lib.b200_op_w4_tile(ptr(wq), ptr(sc), ptr(zp), K, N, G, I, ptr(ch), None))
    ws.append(ch)
x = torch.randn((T, K), device="cuda").half()
act = torch.empty((T, I), dtype=torch.float16, device="cuda")
for i in range(12):
    _lib.check(lib.b200_op_w4_gemm_umma(ptr(ws[i % 10]), ptr(x), K, ptr(act), I, None, T, K, N, G, 2, 1, 0, None))
torch.cuda.synchronize()
print("ok", float(act.float().abs().mean()))

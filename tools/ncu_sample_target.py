"""Target for ncu: the sampler on [64, 151936] logits with temperature 0.7 / top_p 0.9 (the reference's defaults)."""
import ctypes as C
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from vgate_b200 import _lib

lib = _lib.load()
ptr = lambda t: C.c_void_p(t.data_ptr())
B, V = 64, 151936
g = torch.Generator(device="cuda").manual_seed(0)
logits = torch.randn((B, V), device="cuda", generator=g) * 0.78
temp = torch.full((B,), 0.7, device="cuda")
topp = torch.full((B,), 0.9, device="cuda")
seed = torch.arange(B, dtype=torch.int32, device="cuda")
step = torch.zeros(B, dtype=torch.int32, device="cuda")
out = torch.zeros(B, dtype=torch.int32, device="cuda")
ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
for it in range(6):
    if it == 3:
        ev[0].record()
    _lib.check(lib.b200_op_sample(ptr(logits), B, V, ptr(temp), ptr(topp), ptr(seed), ptr(step), ptr(out), None))
ev[1].record()
torch.cuda.synchronize()
print("ok", out[:4].tolist(), f"{ev[0].elapsed_time(ev[1]) / 3 * 1e3:.1f} us per launch")

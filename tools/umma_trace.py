"""Where a tcgen05 W4 GEMM CTA spends its time inside a decode step (b200_debug_umma_trace_grid: 8 clock64
stamps per CTA of the last launch with the given grid).  usage: python tools/umma_trace.py [batch]"""
import ctypes as C
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
from vgate_b200 import _lib
from vgate_b200.engine import Engine
from vgate_b200.weights import PRESETS, synthetic_checkpoint

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
dims = PRESETS["qwen2.5-1.5b-awq-l4"]
lib = _lib.load()
rng = np.random.default_rng(1)
prompts = rng.integers(0, dims.vocab, size=(B, 250), dtype=np.int32)
clk = 1.9e3
for name, gx, gz in (("qkv", 16, 4), ("o", 12, 4), ("gate_up", 140, 1), ("down", 12, 12)):
    eng = Engine(dims, max_seqs=64, max_model_len=512, max_batch_tokens=16384, use_cuda_graphs=False)
    eng.load_checkpoint(synthetic_checkpoint(dims, seed=0))
    n = gx * gz
    buf = torch.zeros(n * 8, dtype=torch.int64, device="cuda")
    lib.b200_debug_umma_trace_grid(C.c_void_p(buf.data_ptr()), gx, gz)
    eng.generate_batch(prompts, 6)
    lib.b200_debug_umma_trace_grid(None, 0, 0)
    eng.close()
    t = buf.cpu().numpy().reshape(n, 8).astype(np.float64)
    seg = [("entry -> set-up done", 0, 1), ("entry -> dependency wait returned", 0, 7), ("wait returned -> first stage complete", 7, 2),
           ("first stage -> last MMA issued", 2, 3), ("-> accumulator complete", 3, 4), ("-> staged in smem", 4, 5),
           ("-> epilogue done", 5, 6), ("wait returned -> epilogue done", 7, 6), ("entry -> epilogue done", 0, 6)]
    print(f"{name}: grid {gx}x1x{gz}, batch {B} (us at ~1.9 GHz, median / p90 over CTAs)")
    for label, a, b in seg:
        d = (t[:, b] - t[:, a]) / clk
        print(f"   {label:40s} {np.median(d):6.2f} / {np.percentile(d, 90):6.2f}")

"""Where a decode-attention CTA spends its time, in situ (b200_debug_attn_trace: 8 clock64 stamps per CTA of
the last attention launch of an eager decode step).  usage: python tools/attn_trace.py [batch] [prompt_len]"""
import ctypes as C
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
from vgate_b200 import _lib
from vgate_b200.engine import Engine
from vgate_b200.weights import PRESETS, synthetic_checkpoint

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
PL = int(sys.argv[2]) if len(sys.argv) > 2 else 250
dims = PRESETS["qwen2.5-1.5b-awq-l4"]
lib = _lib.load()
eng = Engine(dims, max_seqs=64, max_model_len=512, max_batch_tokens=16384, use_cuda_graphs=False)
eng.load_checkpoint(synthetic_checkpoint(dims, seed=0))
rng = np.random.default_rng(1)
prompts = rng.integers(0, dims.vocab, size=(B, PL), dtype=np.int32)
n_cta = dims.n_kv_heads * B
buf = torch.zeros(n_cta * 8, dtype=torch.int64, device="cuda")
lib.b200_debug_attn_trace(C.c_void_p(buf.data_ptr()))
eng.generate_batch(prompts, 6)
lib.b200_debug_attn_trace(None)
eng.close()
t = buf.cpu().numpy().reshape(n_cta, 8).astype(np.float64)
clk = 1.9e3  # cycles per us (approx.)
names = ["entry -> after dependency wait + tile check", "-> q/k/v prologue done (rope, smem)", "-> first K/V tile landed",
         "-> scores (QK^T)", "-> softmax", "-> chunk loop end (PV, further chunks)", "-> merge + store"]
idx = [(0, 1), (1, 2), (2, 3), (3, 4), (4, 5), (5, 6), (6, 7)]
print(f"batch {B}, context ~{PL + 5}: per-CTA medians (us at ~1.9 GHz), {n_cta} CTAs")
for n, (a, b) in zip(names, idx):
    d = (t[:, b] - t[:, a]) / clk
    print(f"  {n:48s} median {np.median(d):6.2f}   p90 {np.percentile(d, 90):6.2f}")
print(f"  total entry -> end                               median {np.median((t[:, 7] - t[:, 0]) / clk):6.2f}")
print(f"  after wait -> end                                median {np.median((t[:, 7] - t[:, 1]) / clk):6.2f}")
